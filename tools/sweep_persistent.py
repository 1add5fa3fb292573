#!/usr/bin/env python
"""Persistent work-queue decode kernel (variant 4) against the grid kernel's automatic choice, GPU-only time.

    python tools/sweep_persistent.py [quick|full] [shape,shape,...]   -> gpurun_out/sweep_persistent.json

Development tool: the heuristic in csrc/pa_api.cu (make_persistent_plan, variant selection) is fit to this."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from tools.sweep import time_gpu_only  # noqa: E402
from vllm_b200 import _cabi, attention_ops  # noqa: E402

DEV = torch.device("cuda", 0)

# name -> (seqs, heads on this GPU, head_size, block_size, ctx, dtype, ragged)
SHAPES = {
    "cfg3_g8_h5": (256, 5, 128, 16, 2048, "bf16", False),
    "cfg3_g4_h10": (256, 10, 128, 16, 2048, "bf16", False),
    "cfg3_g2_h20": (256, 20, 128, 16, 2048, "bf16", False),
    "cfg3_g1_h40": (256, 40, 128, 16, 2048, "bf16", False),
    "cfg2_h32": (256, 32, 128, 16, 1024, "fp16", False),
    "cfg2_g8_h4": (256, 4, 128, 16, 1024, "fp16", False),
    "cfg5a_neox_h8_d96": (64, 8, 96, 16, 8192, "fp16", False),
    "cfg5b_opt66_h9": (64, 9, 128, 16, 8192, "fp16", False),
    "S16_L512": (16, 32, 128, 16, 512, "fp16", False),
    "L64": (256, 32, 128, 16, 64, "fp16", False),
    "L256": (256, 32, 128, 16, 256, "fp16", False),
    "ragged_1_4096": (256, 32, 128, 16, 4096, "fp16", True),
    "S64_ragged_1_2048": (64, 32, 128, 16, 2048, "fp16", True),
    "small_s8_h40": (8, 40, 128, 16, 2048, "fp16", False),
    "b32_h32": (256, 32, 128, 32, 1024, "fp16", False),
    "d64_b8_h32": (256, 32, 64, 8, 1024, "fp16", False),
    "f32_h32": (128, 32, 128, 16, 1024, "fp32", False),
}

QUICK = [(0, 0, 0, 0), (4, 0, 0, 0), (4, 0, 0, 1), (4, 0, 0, 2), (4, 0, 0, 4), (4, 0, 0, 8), (4, 0, 0, 16), (4, 8, 3, 0),
         (4, 6, 4, 0), (4, 12, 2, 8), (4, 16, 1, 0)]
FULL = QUICK + [(4, w, s, x) for w in (6, 8, 10, 12) for s in (2, 3, 4) for x in (2, 4, 8, 16, 32)
                if w * s <= 26]


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "quick"
    names = sys.argv[2].split(",") if len(sys.argv) > 2 else list(SHAPES)
    tunings = FULL if mode == "full" else QUICK
    rows = []
    for name in names:
        S, H, D, B, L, dt, ragged = SHAPES[name]
        wl = dict(heads=H, head_size=D, block_size=B, seqs=S, ctx=L, dtype=dt)
        es = 4 if dt == "fp32" else 2
        kv_bytes = 2 * S * H * L * D * es
        ncopies = 1 if kv_bytes > 600e6 else 3        # rotate over copies when one fits the 126 MB L2 twice over
        ps = [bench.make_problem(wl, H, DEV, seed=i) for i in range(ncopies)]
        if ragged:
            g = torch.Generator().manual_seed(1)
            lens = torch.randint(1, L + 1, (S,), generator=g, dtype=torch.int32)
            for q in ps:
                q["lens"] = lens.to(DEV)
        lens = ps[0]["lens"].cpu().tolist()
        nbytes = sum(H * (2 * c * D * es + 2 * D * es) + ((c + B - 1) // B) * 4 + 4 for c in lens)
        out = torch.empty(ps[0]["q"].shape, dtype=ps[0]["dtype"], device=DEV)
        state = {"i": 0}
        row = dict(shape=name, seqs=S, heads=H, head_size=D, block_size=B, ctx=L, dtype=dt, mbytes=round(nbytes / 1e6, 1))
        ref_out = None
        for (v, w, s, x) in tunings:
            tun = _cabi.DecodeTuning(v, w, s, x) if v else None

            def call():
                state["i"] += 1
                p = ps[state["i"] % ncopies]
                attention_ops.single_query_cached_kv_attention(out, p["q"], p["kc"], p["vc"], p["scale"], p["tables"],
                                                               p["lens"], B, L, tuning=tun)
            try:
                ms = time_gpu_only(call, iters=20 if nbytes > 2e9 else 40)
            except RuntimeError as e:
                row[f"v{v}_w{w}_s{s}_x{x}"] = str(e)[:80]
                continue
            # same answer as the automatic grid kernel (on copy 0)
            state["i"] = -1
            call()
            if ref_out is None:
                ref_out = out.clone()
            else:
                err = float((out.float() - ref_out.float()).abs().max())
                assert err < 2e-2, (name, v, w, s, x, err)
            row[f"v{v}_w{w}_s{s}_x{x}"] = dict(us=round(ms * 1e3, 1), gbs=round(nbytes / ms / 1e6))
        best = min(((k, r["us"]) for k, r in row.items() if isinstance(r, dict)), key=lambda t: t[1])
        row["best"] = best[0]
        rows.append(row)
        print(json.dumps(row), flush=True)
        del ps
        torch.cuda.empty_cache()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "sweep_persistent.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
