#!/usr/bin/env python
"""reshape_and_cache at 128 / 256 / 512 tokens x 40 heads x 128 (fp16, block 16): the tile mover ("runs") against the
per-token element sweep ("scattered") on consecutive slots (a short prompt) and on scattered slots (a decode step),
with the reference's kernel beside them when oracle/_ref is built.  Device time, host running ahead.

    python tools/reshape_hint_bench.py  -> gpurun_out/reshape_hint_bench.json
"""
from __future__ import annotations

import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools.sweep import time_gpu_only  # noqa: E402
from vllm_b200 import cache_ops  # noqa: E402

DEV = torch.device("cuda", 0)


def main():
    H, D, B, NB = 40, 128, 16, 1024
    kc = torch.randn((NB, H, D // 8, B, 8), device=DEV).half()
    vc = torch.randn((NB, H, D, B), device=DEV).half()
    ref = None
    path = os.path.join(ROOT, "oracle", "_ref", "ref_cache_ops.so")
    if os.path.exists(path):
        import importlib.util
        spec = importlib.util.spec_from_file_location("ref_cache_ops", path)
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
    rows = []
    for N in (128, 256, 512):
        qkv = torch.randn((N, 3, H, D), device=DEV).half()
        layouts = {
            "consecutive": torch.arange(3 * B, 3 * B + N, device=DEV).to(torch.int32),
            "scattered": (torch.randperm(NB, device=DEV)[:N] * B + torch.randint(0, B, (N,), device=DEV)).to(torch.int32),
        }
        for lname, slots in layouts.items():
            row = dict(tokens=N, slots=lname)
            for hint in ("runs", "scattered", None):
                us = 1e3 * time_gpu_only(lambda: cache_ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, slots, slots=hint),
                                         iters=50)
                row[f"ours_{hint or 'no_hint'}_us"] = round(us, 2)
            if ref is not None:
                row["reference_kernel_us"] = round(1e3 * time_gpu_only(
                    lambda: ref.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, slots), iters=50), 2)
            rows.append(row)
            print(json.dumps(row), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "reshape_hint_bench.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
