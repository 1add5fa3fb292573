#!/usr/bin/env python
"""Decode time against the number of sequences (H heads per GPU fixed), around wave boundaries of the grid kernel:
separates the per-launch fixed cost (intercept) from wave quantisation (steps) and the marginal rate (slope)."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from tools.sweep import time_gpu_only
from vllm_b200 import _cabi, attention_ops
DEV = torch.device("cuda", 0)

def main():
    H = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    variants = [int(v) for v in (sys.argv[2].split(",") if len(sys.argv) > 2 else ["0", "4"])]
    L, D, B = 2048, 128, 16
    wl = dict(heads=H, head_size=D, block_size=B, seqs=1100, ctx=L, dtype="bf16")
    p = bench.make_problem(wl, H, DEV)
    rows = []
    for S in (22, 44, 88, 89, 90, 133, 177, 178, 179, 222, 256, 266, 267, 268, 311, 355, 356, 444, 532, 533, 534, 710, 888, 1065, 1066):
        row = {"seqs": S, "items": S * H, "waves444": round(S * H / 444, 3)}
        out = torch.empty((S, H, D), dtype=p["dtype"], device=DEV)
        nbytes = bench.algorithmic_bytes(S, H, D, B, L, 2)
        for v in variants:
            tun = _cabi.DecodeTuning(v, 0, 0, 0) if v else None
            ms = time_gpu_only(lambda: attention_ops.single_query_cached_kv_attention(
                out, p["q"][:S], p["kc"], p["vc"], p["scale"], p["tables"][:S], p["lens"][:S], B, L, tuning=tun), iters=30)
            row[f"v{v}_us"] = round(ms * 1e3, 1)
            row[f"v{v}_gbs"] = round(nbytes / ms / 1e6)
        rows.append(row)
        print(json.dumps(row), flush=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", f"wave_probe_h{H}.json"), "w"), indent=1)

if __name__ == "__main__":
    main()
