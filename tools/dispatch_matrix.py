#!/usr/bin/env python
"""Times the decode kernel over the reference's whole dispatch set on one B200:
{fp16, bf16, fp32} x block {8, 16, 32} x head size {64, 80, 96, 128}
(attention_kernels.cu:370-400, 416-448, 460-468), at 256 sequences x 32 heads x 1024 tokens, next to the
reference's own kernel (oracle/_ref, compiled unmodified for sm_100) on the same tensors.

    python tools/dispatch_matrix.py [--seqs 256 --heads 32 --ctx 1024]  -> gpurun_out/dispatch_matrix.json
    python tools/dispatch_matrix.py --cells fp16:8:64,fp32:32:128 --tunings 2.4.2,2.3.2,1.8.1 [--no-ref]
        only these cells (dtype:block:head), each under every forced launch shape (variant.warps.stages) too

Per cell: ours_us / ours_gbs (algorithmic bytes of SURVEY.md section 8(d) over device time), frac_of_8tbs,
reference_kernel_gbs, speedup, the automatic plan, and the largest difference between the two outputs.
Every cell's KV footprint is far above the 126 MB L2, so no flush is needed between calls.
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from tools.sweep import time_gpu_only  # noqa: E402
from vllm_b200 import attention_ops  # noqa: E402

DEV = torch.device("cuda", 0)


def load_ref():
    path = os.path.join(ROOT, "oracle", "_ref", "ref_attention_ops.so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("ref_attention_ops", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seqs", type=int, default=256)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--ctx", type=int, default=1024)
    ap.add_argument("--cells", default="", help="dtype:block:head,... (default: all 36)")
    ap.add_argument("--tunings", default="", help="variant.warps.stages,... timed beside the automatic plan")
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "dispatch_matrix.json"))
    a = ap.parse_args()
    from vllm_b200 import _cabi
    ref = None if a.no_ref else load_ref()
    rows = []
    only = [tuple(c.split(":")) for c in a.cells.split(",") if c]
    tunings = [tuple(int(x) for x in t.split(".")) for t in a.tunings.split(",") if t]
    for dt in ("fp16", "bf16", "fp32"):
        for B in (8, 16, 32):
            for D in (64, 80, 96, 128):
                if only and (dt, str(B), str(D)) not in only:
                    continue
                wl = dict(heads=a.heads, head_size=D, block_size=B, seqs=a.seqs, ctx=a.ctx, dtype=dt)
                p = bench.make_problem(wl, a.heads, DEV)
                es = p["kc"].element_size()
                nbytes = bench.algorithmic_bytes(a.seqs, a.heads, D, B, a.ctx, es)
                out_a = torch.empty(p["q"].shape, dtype=p["dtype"], device=DEV)
                out_b = torch.empty_like(out_a)
                plan = attention_ops.decode_plan(a.seqs, a.heads, D, B, a.ctx, p["dtype"])
                ours = lambda: attention_ops.single_query_cached_kv_attention(
                    out_a, p["q"], p["kc"], p["vc"], p["scale"], p["tables"], p["lens"], B, a.ctx)
                ms = time_gpu_only(ours, iters=20, warm=3)
                row = dict(dtype=dt, block_size=B, head_size=D, page_bytes=D * B * es, ours_us=round(ms * 1e3, 1),
                           ours_gbs=round(nbytes / ms / 1e6, 1), frac_of_8tbs=round(nbytes / ms / 1e6 / 8000, 3),
                           plan=[plan.variant, plan.num_warps, plan.num_stages, plan.num_splits])
                forced = {}
                for (v, w, st) in tunings:
                    tun = _cabi.DecodeTuning(v, w, st, 1)
                    try:
                        fms = time_gpu_only(lambda: attention_ops.single_query_cached_kv_attention(
                            out_a, p["q"], p["kc"], p["vc"], p["scale"], p["tables"], p["lens"], B, a.ctx, tuning=tun),
                            iters=20, warm=3)
                        forced[f"{v}.{w}.{st}"] = round(nbytes / fms / 1e6, 1)
                    except RuntimeError as e:
                        forced[f"{v}.{w}.{st}"] = str(e)[:60]
                if forced:
                    row["forced_gbs"] = forced
                if ref is not None and p["kc"].numel() < 2 ** 31:
                    theirs = lambda: ref.single_query_cached_kv_attention(
                        out_b, p["q"], p["kc"], p["vc"], p["scale"], p["tables"], p["lens"], B, a.ctx)
                    rms = time_gpu_only(theirs, iters=10, warm=2)
                    row.update(reference_kernel_us=round(rms * 1e3, 1), reference_kernel_gbs=round(nbytes / rms / 1e6, 1),
                               speedup=round(rms / ms, 3),
                               max_rel_diff=float((out_a.float() - out_b.float()).abs().max() / out_b.float().abs().max()))
                rows.append(row)
                print(json.dumps(row), flush=True)
                del p, out_a, out_b
                torch.cuda.empty_cache()
    worst = min(rows, key=lambda r: r["frac_of_8tbs"])
    slow = min((r for r in rows if "speedup" in r), key=lambda r: r["speedup"], default=None)
    res = dict(workload=dict(seqs=a.seqs, heads=a.heads, ctx=a.ctx), cells=rows, worst_frac_cell=worst,
               worst_speedup_cell=slow)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(res, open(a.out, "w"), indent=1)
    print(json.dumps(dict(worst_frac_cell=worst, worst_speedup_cell=slow)))


if __name__ == "__main__":
    main()
