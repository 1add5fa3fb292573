#!/usr/bin/env python
"""Parallel sampling / beam workload (BASELINE configs[3] shape): groups of 4 siblings sharing the prompt's
pages.  Times the decode kernel with sibling groups served together (variant 2) and one by one
(variant 3), and the reference kernel."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools.sweep import time_gpu_only
from vllm_b200 import _cabi, attention_ops
DEV = torch.device("cuda", 0)

def main():
    try:
        from tools.compare_reference_kernel import load
        ref = load("ref_attention_ops")
    except Exception:
        ref = None
    rows = []
    H, D, B = 40, 128, 16
    for name, groups, width, prompt, gen in (("n4_prompt1024_gen32", 64, 4, 1024, 32), ("n4_prompt512_gen128", 64, 4, 512, 128),
                                             ("n4_prompt2048_gen16", 32, 4, 2048, 16), ("beam8_prompt1024_gen32", 32, 8, 1024, 32),
                                             ("n2_prompt1024_gen32", 128, 2, 1024, 32), ("n1_prompt1024_gen32", 256, 1, 1024, 32)):
        S = groups * width
        L = prompt + gen
        npp, npg = prompt // B, (gen + B - 1) // B
        NB = groups * npp + S * npg
        g = torch.Generator(device=DEV).manual_seed(0)
        kc = torch.empty((NB, H, D // 8, B, 8), dtype=torch.float16, device=DEV).normal_(generator=g)
        vc = torch.empty((NB, H, D, B), dtype=torch.float16, device=DEV).normal_(generator=g)
        q = torch.empty((S, 3, H, D), dtype=torch.float16, device=DEV).normal_(generator=g)[:, 0]
        perm = torch.randperm(NB, generator=g, device=DEV).to(torch.int32)
        prompt_blocks = perm[:groups * npp].view(groups, npp)
        gen_blocks = perm[groups * npp:].view(S, npg)
        tables = torch.cat([prompt_blocks.repeat_interleave(width, dim=0), gen_blocks], dim=1).contiguous()
        lens = torch.full((S,), L, dtype=torch.int32, device=DEV)
        out = torch.empty((S, H, D), dtype=torch.float16, device=DEV)
        unique = (groups * npp + S * npg) * H * 2 * D * B * 2
        per_seq = S * H * 2 * L * D * 2
        row = dict(shape=name, seqs=S, width=width, per_sequence_mbytes=round(per_seq / 1e6), unique_mbytes=round(unique / 1e6))
        res = {}
        for label, tun in (("grouped", _cabi.DecodeTuning(2, 0, 0, 0)), ("ungrouped", _cabi.DecodeTuning(3, 0, 0, 0))):
            ms = time_gpu_only(lambda: attention_ops.single_query_cached_kv_attention(out, q, kc, vc, D ** -0.5, tables, lens, B, L,
                                                                                     tuning=tun), iters=20)
            res[label] = out.clone()
            row[label + "_us"] = round(ms * 1e3, 1)
        row["speedup_grouped"] = round(row["ungrouped_us"] / row["grouped_us"], 2)
        row["max_abs_diff"] = float((res["grouped"].float() - res["ungrouped"].float()).abs().max())
        if ref is not None and kc.numel() < 2 ** 31:
            ms = time_gpu_only(lambda: ref.single_query_cached_kv_attention(out, q, kc, vc, D ** -0.5, tables, lens, B, L), iters=10)
            row["reference_kernel_us"] = round(ms * 1e3, 1)
            row["speedup_vs_reference_kernel"] = round(row["reference_kernel_us"] / row["grouped_us"], 2)
        rows.append(row)
        print(json.dumps(row), flush=True)
        del kc, vc
        torch.cuda.empty_cache()
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "sampling_bench.json"), "w"), indent=1)

if __name__ == "__main__":
    main()
