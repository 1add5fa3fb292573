#!/usr/bin/env python
"""One call of the decode kernel per named cell (dtype:block:head[:variant.warps.stages]) at 256 x 32 x 1024 -- the
target of `ncu --set full` when a cell of the dispatch matrix needs explaining.

    ncu --set full --clock-control none --import-source on -o gpurun_out/cells python tools/dispatch_ncu.py fp16:8:64 fp16:8:80
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from vllm_b200 import _cabi, attention_ops  # noqa: E402

DEV = torch.device("cuda", 0)
for cell in sys.argv[1:]:
    parts = cell.split(":")
    dt, B, D = parts[0], int(parts[1]), int(parts[2])
    tun = None
    if len(parts) > 3:
        v, w, st = (int(x) for x in parts[3].split("."))
        tun = _cabi.DecodeTuning(v, w, st, 1)
    p = bench.make_problem(dict(heads=32, head_size=D, block_size=B, seqs=256, ctx=1024, dtype=dt), 32, DEV)
    out = torch.empty(p["q"].shape, dtype=p["dtype"], device=DEV)
    for _ in range(2):
        attention_ops.single_query_cached_kv_attention(out, p["q"], p["kc"], p["vc"], p["scale"], p["tables"], p["lens"],
                                                       B, 1024, tuning=tun)
    torch.cuda.synchronize()
    del p, out
    torch.cuda.empty_cache()
