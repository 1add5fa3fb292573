"""Times pa_multi_query_cached_kv_attention (prompt chunks over the paged cache; 16-bit types: one launch whose
producer warps read the pages) next to a stand-alone gather of the same pages (what a gather-then-attend design
would pay on top) and to recomputing the whole prompt.  Writes gpurun_out/cached_prefill_bench.json.

    python tools/cached_prefill_bench.py

Flops counted: 4 * head_size per (query row, visible key, head).  The comparator is what a v0.1.0 user
would have to do without this operator: recompute the whole prompt with multi_query_kv_attention.
"""
from __future__ import annotations

import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools.prefill_bench import time_fn  # noqa: E402

SHAPES = [
    # name, heads, head size, block size, [(query_len, context_len)], dtype
    ("llama13b_chunk512_of_2048_x8", 40, 128, 16, [(512, 2048)] * 8, torch.bfloat16),
    ("llama13b_chunk2048_of_8192_x4", 40, 128, 16, [(2048, 8192)] * 4, torch.bfloat16),
    ("llama13b_prefix1024_new256_x10", 40, 128, 16, [(256, 1280)] * 10, torch.bfloat16),
    ("llama7b_fp16_ragged", 32, 128, 16, [(37, 1500), (911, 911), (128, 4096), (460, 600)], torch.float16),
    ("neox20b_d96_chunk512_of_2048_x8", 64, 96, 16, [(512, 2048)] * 8, torch.float16),
]


def main():
    from vllm_b200 import attention_ops, cache_ops
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
    except Exception:
        peak = None
    rows = []
    for name, h, d, b, pairs, dtype in SHAPES:
        x = 8
        ctx = [p[1] for p in pairs]
        qlens = [p[0] for p in pairs]
        cu = [0]
        for n in qlens:
            cu.append(cu[-1] + n)
        nblocks = sum((c + b - 1) // b for c in ctx)
        g = torch.Generator(device="cuda").manual_seed(0)
        kc = torch.empty((nblocks, h, d // x, b, x), dtype=dtype, device="cuda").normal_(generator=g)
        vc = torch.empty((nblocks, h, d, b), dtype=dtype, device="cuda").normal_(generator=g)
        q = torch.empty((cu[-1], h, d), dtype=dtype, device="cuda").normal_(generator=g)
        out = torch.empty_like(q)
        perm = torch.randperm(nblocks, generator=torch.Generator().manual_seed(1)).int()
        mb = max((c + b - 1) // b for c in ctx)
        bt = torch.zeros((len(pairs), mb), dtype=torch.int32)
        at = 0
        for i, c in enumerate(ctx):
            n = (c + b - 1) // b
            bt[i, :n] = perm[at:at + n]
            at += n
        btd = bt.cuda()
        scale = d ** -0.5
        ms = time_fn(lambda: attention_ops.multi_query_cached_kv_attention(cu, out, q, kc, vc, scale, btd, ctx, b))
        # for scale: a stand-alone gather of the same pages into packed rows (not part of the operator)
        slots = torch.cat([bt[i, torch.arange(c) // b] * b + (torch.arange(c) % b).int() for i, c in enumerate(ctx)]).int().cuda()
        kr = torch.empty((sum(ctx), h, d), dtype=dtype, device="cuda")
        vr = torch.empty_like(kr)
        ms_g = time_fn(lambda: cache_ops.gather_cached_kv(kr, vr, kc, vc, slots))
        fl = sum(4 * d * h * sum(min(c, r + 1 + c - n) for r in range(n)) for n, c in pairs)
        kv_bytes = 2 * sum(ctx) * h * d * 2
        row = {"shape": name, "heads": h, "head_size": d, "block_size": b, "pairs": pairs if len(set(pairs)) > 1 else [pairs[0], len(pairs)],
               "dtype": str(dtype).split(".")[-1], "ms": round(ms, 4), "tflops": round(fl / ms / 1e9, 1),
               "frac_of_measured_bf16_peak": round(fl / ms / 1e9 / peak, 3) if peak else None,
               "standalone_gather_of_the_same_pages_ms": round(ms_g, 4),
               "gather_gbs_read_plus_write": round(2 * kv_bytes / ms_g / 1e6, 1)}
        # without the operator: recompute the whole prompt (only valid to compare when every context is a full prompt)
        qf = torch.empty((sum(ctx), 3, h, d), dtype=dtype, device="cuda").normal_(generator=g)
        of = torch.empty((sum(ctx), h, d), dtype=dtype, device="cuda")
        ms_full = time_fn(lambda: attention_ops.multi_query_kv_attention(of, qf[:, 0], qf[:, 1], qf[:, 2], ctx, scale))
        row["recompute_whole_prompt_ms"] = round(ms_full, 4)
        row["speedup_vs_recompute"] = round(ms_full / ms, 2)
        print(json.dumps(row), flush=True)
        rows.append(row)
        del kc, vc, q, out, kr, vr, qf, of
        torch.cuda.empty_cache()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "cached_prefill_bench.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
