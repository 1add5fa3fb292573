"""Times pa_prefill_attention (tcgen05) on prompt batches and, where the image has them, the
library kernels a port would have used (flash_attn 2.x varlen = mma.sync recompiled for sm_100,
torch SDPA) on the same tensors.  Writes gpurun_out/prefill_bench.json.

    python tools/prefill_bench.py [--json gpurun_out/prefill_bench.json]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = [
    # name, heads, head size, prompt lengths, dtype
    ("llama13b_5x512", 40, 128, [512] * 5, torch.bfloat16),          # v0.1.0 max_num_batched_tokens = 2560
    ("llama13b_ragged_2560", 40, 128, [37, 911, 128, 1024, 460], torch.bfloat16),
    ("llama13b_8x2048", 40, 128, [2048] * 8, torch.bfloat16),
    ("llama13b_8x4096", 40, 128, [4096] * 8, torch.bfloat16),
    ("llama7b_fp16_8x2048", 32, 128, [2048] * 8, torch.float16),
    ("neox20b_d96_8x2048", 64, 96, [2048] * 8, torch.float16),
    ("opt_d64_8x2048", 32, 64, [2048] * 8, torch.float16),
]


def flops(heads, d, lens):
    return sum(4 * d * heads * (n * (n + 1) // 2) for n in lens)


def time_fn(fn, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=os.path.join(ROOT, "gpurun_out", "prefill_bench.json"))
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    from vllm_b200 import attention_ops
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
    except Exception:
        peak = None
    rows = []
    for name, h, d, lens, dtype in SHAPES:
        if args.only and args.only not in name:
            continue
        n = sum(lens)
        g = torch.Generator(device="cuda").manual_seed(0)
        qkv = torch.empty((n, 3, h, d), dtype=dtype, device="cuda").normal_(generator=g)
        q, k, v = qkv[:, 0], qkv[:, 1], qkv[:, 2]
        out = torch.empty((n, h, d), dtype=dtype, device="cuda")
        scale = d ** -0.5
        ms = time_fn(lambda: attention_ops.multi_query_kv_attention(out, q, k, v, lens, scale))
        fl = flops(h, d, lens)
        row = {"shape": name, "heads": h, "head_size": d, "prompt_lens": lens, "dtype": str(dtype).split(".")[-1],
               "ours_ms": round(ms, 4), "ours_tflops": round(fl / ms / 1e9, 1),
               "frac_of_measured_bf16_peak": round(fl / ms / 1e9 / peak, 3) if peak else None}
        try:
            from flash_attn import flash_attn_varlen_func
            cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device="cuda")
            fa = lambda: flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), softmax_scale=scale, causal=True)
            ref = fa()
            ms_fa = time_fn(fa)
            row["flash_attn2_ms"] = round(ms_fa, 4)
            row["flash_attn2_tflops"] = round(fl / ms_fa / 1e9, 1)
            row["speedup_vs_flash_attn2"] = round(ms_fa / ms, 2)
            row["max_abs_diff_vs_flash_attn2"] = float((ref.float() - out.float()).abs().max())
        except Exception as e:  # comparator only
            row["flash_attn2_error"] = repr(e)[:200]
        if len(set(lens)) == 1:
            try:
                import torch.nn.functional as F
                b, L = len(lens), lens[0]
                q4 = q.reshape(b, L, h, d).transpose(1, 2)
                k4 = k.reshape(b, L, h, d).transpose(1, 2)
                v4 = v.reshape(b, L, h, d).transpose(1, 2)
                sd = lambda: F.scaled_dot_product_attention(q4, k4, v4, is_causal=True, scale=scale)
                ms_sd = time_fn(sd)
                row["torch_sdpa_ms"] = round(ms_sd, 4)
                row["torch_sdpa_tflops"] = round(fl / ms_sd / 1e9, 1)
            except Exception as e:
                row["torch_sdpa_error"] = repr(e)[:200]
        print(json.dumps(row), flush=True)
        rows.append(row)
        del qkv, out
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(args.json), exist_ok=True)
    json.dump(rows, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
