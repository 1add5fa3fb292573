#!/usr/bin/env python
"""Launch-shape sweep of the decode kernel and timing of the cache movers on one B200.

    python tools/sweep.py decode [--shapes a,b,...]     -> gpurun_out/sweep_decode.json
    python tools/sweep.py movers                        -> gpurun_out/sweep_movers.json

Development tool: times explicit launch shapes (variant, warps, stages, splits) next to the
library's automatic choice so the heuristic in csrc/pa_api.cu can be checked against data.
"""
from __future__ import annotations

import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from vllm_b200 import _cabi, attention_ops, cache_ops  # noqa: E402

DEV = torch.device("cuda", 0)

# name -> (seqs, heads on this GPU, head_size, block_size, ctx, dtype)
SHAPES = {
    "cfg2_h32": (256, 32, 128, 16, 1024, "fp16"),
    "cfg3_g1_h40": (256, 40, 128, 16, 2048, "bf16"),
    "cfg3_g2_h20": (256, 20, 128, 16, 2048, "bf16"),
    "cfg3_g4_h10": (256, 10, 128, 16, 2048, "bf16"),
    "cfg3_g8_h5": (256, 5, 128, 16, 2048, "bf16"),
    "cfg5a_neox_h8_d96": (64, 8, 96, 16, 8192, "fp16"),
    "cfg5b_opt66_h9": (64, 9, 128, 16, 8192, "fp16"),
    "small_s8_h40": (8, 40, 128, 16, 2048, "fp16"),
    "small_s1_h40": (1, 40, 128, 16, 4096, "fp16"),
    "b32_h32": (256, 32, 128, 32, 1024, "fp16"),
    "b8_h32": (256, 32, 128, 8, 1024, "fp16"),
    "f32_h32": (128, 32, 128, 16, 1024, "fp32"),
    "d64_h32": (256, 32, 64, 16, 1024, "fp16"),
    "d80_h32": (256, 32, 80, 16, 1024, "fp16"),
}

TUNINGS = [(0, 0, 0, 0), (1, 4, 0, 0), (1, 8, 0, 0), (2, 4, 2, 0), (2, 4, 4, 0), (2, 2, 4, 0), (2, 2, 8, 0),
           (2, 8, 2, 0), (2, 1, 8, 0), (2, 4, 3, 0), (2, 4, 6, 0)]


def time_calls(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def time_gpu_only(fn, iters=30, warm=5):
    """Device time per call with the host running ahead: a spin kernel holds the stream while all
    calls are enqueued, so host-side launch cost is hidden and only GPU execution is between the events."""
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    torch.cuda._sleep(int(4e6) + iters * 100000)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def sweep_splits(names):
    res = []
    for name in names:
        S, H, D, B, L, dt = SHAPES[name]
        wl = dict(heads=H, head_size=D, block_size=B, seqs=S, ctx=L, dtype=dt)
        p = bench.make_problem(wl, H, DEV)
        out = torch.empty(p["q"].shape, dtype=p["dtype"], device=DEV)
        nbytes = bench.algorithmic_bytes(S, H, D, B, L, p["kc"].element_size())
        plan = attention_ops.decode_plan(S, H, D, B, L, p["dtype"])
        for w, st in ((4, 2), (8, 2), (4, 3), (8, 1), (2, 4)):
            for sp in (0, 1, 2, 3, 4, 6, 8, 12, 16, 24, 32):
                tun = _cabi.DecodeTuning(2, w, st, sp)
                fn = lambda: attention_ops.single_query_cached_kv_attention(out, p["q"], p["kc"], p["vc"], p["scale"],
                                                                            p["tables"], p["lens"], B, L, tuning=tun)
                try:
                    ms = time_gpu_only(fn, iters=40)
                except RuntimeError as e:
                    continue
                row = dict(shape=name, warps=w, stages=st, splits=sp, ms=round(ms, 4), gbs=round(nbytes / ms / 1e6, 1),
                           auto_plan=[plan.variant, plan.num_warps, plan.num_stages, plan.num_splits])
                res.append(row)
                print(json.dumps(row), flush=True)
        del p, out
        torch.cuda.empty_cache()
    return res


def sweep_decode(names, extra_splits=(0,)):
    res = []
    for name in names:
        S, H, D, B, L, dt = SHAPES[name]
        wl = dict(heads=H, head_size=D, block_size=B, seqs=S, ctx=L, dtype=dt)
        p = bench.make_problem(wl, H, DEV)
        out = torch.empty(p["q"].shape, dtype=p["dtype"], device=DEV)
        esize = p["kc"].element_size()
        nbytes = bench.algorithmic_bytes(S, H, D, B, L, esize)
        small = 2 * p["kc"].numel() * esize < 512e6
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV) if small else None
        plan = attention_ops.decode_plan(S, H, D, B, L, p["dtype"])
        for (v, w, st, sp) in TUNINGS:
            for sp2 in extra_splits:
                tun = _cabi.DecodeTuning(v, w, st, sp2 if sp2 else sp)

                def call():
                    if flush is not None:
                        flush.zero_()
                    attention_ops.single_query_cached_kv_attention(out, p["q"], p["kc"], p["vc"], p["scale"], p["tables"],
                                                                   p["lens"], B, L, tuning=tun)
                try:
                    ms = time_calls(call)
                    if flush is not None:
                        ms -= time_calls(lambda: flush.zero_())
                except RuntimeError as e:
                    res.append(dict(shape=name, tuning=[v, w, st, sp2 or sp], error=str(e)[:100]))
                    continue
                row = dict(shape=name, tuning=[v, w, st, sp2 or sp], ms=round(ms, 4), gbs=round(nbytes / ms / 1e6, 1),
                           auto_plan=[plan.variant, plan.num_warps, plan.num_stages, plan.num_splits])
                res.append(row)
                print(json.dumps(row), flush=True)
        del p, out
        torch.cuda.empty_cache()
    return res


def sweep_movers():
    res = []
    # copy_blocks, config 4: 40 layers, H=40, D=128, B=16 fp16, 192 pairs (64 groups x 3 forks)
    layers, H, D, B, NB = 40, 40, 128, 16, 1024
    kcs = [torch.randn((NB, H, D // 8, B, 8), device=DEV).half() for _ in range(layers)]
    vcs = [torch.randn((NB, H, D, B), device=DEV).half() for _ in range(layers)]
    for pairs in (1, 8, 64, 192):
        mapping = {}
        for i in range(pairs):
            mapping.setdefault(i // 3, []).append(512 + i)
        ms = time_calls(lambda: cache_ops.copy_blocks(kcs, vcs, mapping), iters=20)
        nbytes = 2 * pairs * layers * 2 * H * D * B * 2
        row = dict(op="copy_blocks", pairs=pairs, layers=layers, ms=round(ms, 4), gbs=round(nbytes / ms / 1e6, 1))
        res.append(row)
        print(json.dumps(row), flush=True)
    # reshape_and_cache: decode step (256 tokens) and prefill (2560 tokens)
    for N in (256, 2560):
        qkv = torch.randn((N, 3, H, D), device=DEV).half()
        if N <= 256:     # decode step: every token is the newest token of its own sequence -> its own block
            slots = (torch.randperm(NB, device=DEV)[:N] * B + torch.randint(0, B, (N,), device=DEV)).to(torch.int32)
        else:            # prefill: prompts of 512 tokens fill consecutive slots of scattered block runs
            starts = torch.randperm(NB // 32, device=DEV)[:N // 512] * 32 * B
            slots = (starts[:, None] + torch.arange(512, device=DEV)[None, :]).flatten().to(torch.int32)
        ms = time_calls(lambda: cache_ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kcs[0], vcs[0], slots), iters=50)
        gms = time_gpu_only(lambda: cache_ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kcs[0], vcs[0], slots), iters=50)
        nbytes = N * H * D * 2 * 2 * 2
        row = dict(op="reshape_and_cache", tokens=N, ms=round(ms, 4), gpu_only_ms=round(gms, 4),
                   gbs=round(nbytes / gms / 1e6, 1))
        res.append(row)
        print(json.dumps(row), flush=True)
        out = torch.empty_like(qkv)
        ms = time_calls(lambda: cache_ops.gather_cached_kv(out[:, 1], out[:, 2], kcs[0], vcs[0], slots), iters=50)
        row = dict(op="gather_cached_kv", tokens=N, ms=round(ms, 4), gbs=round(nbytes / ms / 1e6, 1))
        res.append(row)
        print(json.dumps(row), flush=True)
    # swap_blocks: one 4-sequence group of 2048-token contexts = 512 blocks, one layer K
    cpu = torch.empty((NB, H, D, B), dtype=torch.float16).pin_memory()
    for nblk, contiguous in ((128, True), (128, False), (512, False)):
        ids = list(range(nblk)) if contiguous else torch.randperm(NB)[:nblk].tolist()
        mapping = {int(s): int(d) for s, d in zip(ids, ids if contiguous else torch.randperm(NB)[:nblk].tolist())}
        for direction, (src, dst) in (("d2h", (vcs[0], cpu)), ("h2d", (cpu, vcs[0])), ("d2d", (vcs[0], vcs[1]))):
            for engine in (1, 2):
                ms = time_calls(lambda: cache_ops.swap_blocks(src, dst, mapping, engine=engine), iters=10, warm=2)
                nbytes = nblk * H * D * B * 2
                row = dict(op="swap_blocks", dir=direction, blocks=nblk, contiguous=contiguous,
                           engine="copy engines" if engine == 1 else "kernel", ms=round(ms, 4),
                           gbs=round(nbytes / ms / 1e6, 1))
                res.append(row)
                print(json.dumps(row), flush=True)
    return res


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "decode"
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    if what == "splits":
        names = sys.argv[2].split(",") if len(sys.argv) > 2 else ["cfg3_g8_h5", "cfg3_g4_h10", "cfg5a_neox_h8_d96",
                                                                    "cfg5b_opt66_h9", "small_s8_h40", "small_s1_h40"]
        out = sweep_splits(names)
        json.dump(out, open(os.path.join(ROOT, "gpurun_out", "sweep_splits.json"), "w"), indent=1)
    elif what == "decode":
        names = list(SHAPES)
        for a in sys.argv[2:]:
            if a.startswith("--shapes"):
                names = a.split("=", 1)[1].split(",")
        out = sweep_decode(names)
        json.dump(out, open(os.path.join(ROOT, "gpurun_out", "sweep_decode.json"), "w"), indent=1)
    else:
        out = sweep_movers()
        json.dump(out, open(os.path.join(ROOT, "gpurun_out", "sweep_movers.json"), "w"), indent=1)
