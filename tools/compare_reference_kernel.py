#!/usr/bin/env python
"""Times the reference's own CUDA kernels (oracle/_ref, compiled unmodified for sm_100) next to
ours on identical tensors on one B200 -- the like-for-like bar of BASELINE.md section 5.

    python tools/compare_reference_kernel.py   -> gpurun_out/compare_reference.json

Shapes are limited to NB*H*D*B < 2^31 elements: the reference indexes with 32-bit int
(attention_kernels.cu:157,243), so config 3 at one GPU is out of its reach.
"""
from __future__ import annotations

import importlib.util
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from tools.sweep import time_calls, time_gpu_only  # noqa: E402
from vllm_b200 import attention_ops, cache_ops  # noqa: E402

DEV = torch.device("cuda", 0)


def load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "oracle", "_ref", name + ".so"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


SHAPES = {
    "cfg2_llama7b_fp16_h32_L1024": (256, 32, 128, 16, 1024, "fp16"),
    "cfg3_llama13b_bf16_g2_h20_L2048": (256, 20, 128, 16, 2048, "bf16"),
    "cfg3_llama13b_bf16_g8_h5_L2048": (256, 5, 128, 16, 2048, "bf16"),
    "cfg5a_neox20b_fp16_h8_d96_L8192": (64, 8, 96, 16, 8192, "fp16"),
    "cfg5b_opt66b_fp16_h9_L8192": (64, 9, 128, 16, 8192, "fp16"),
}


def main():
    ref_att, ref_cache = load("ref_attention_ops"), load("ref_cache_ops")
    rows = []
    for name, (S, H, D, B, L, dt) in SHAPES.items():
        wl = dict(heads=H, head_size=D, block_size=B, seqs=S, ctx=L, dtype=dt)
        p = bench.make_problem(wl, H, DEV)
        assert p["kc"].numel() < 2 ** 31
        out_a = torch.empty(p["q"].shape, dtype=p["dtype"], device=DEV)
        out_b = torch.empty_like(out_a)
        nbytes = bench.algorithmic_bytes(S, H, D, B, L, p["kc"].element_size())
        ms_ours = time_calls(lambda: attention_ops.single_query_cached_kv_attention(
            out_a, p["q"], p["kc"], p["vc"], p["scale"], p["tables"], p["lens"], B, L))
        ms_ref = time_calls(lambda: ref_att.single_query_cached_kv_attention(
            out_b, p["q"], p["kc"], p["vc"], p["scale"], p["tables"], p["lens"], B, L), iters=10, warm=3)
        diff = float((out_a.float() - out_b.float()).abs().max() / out_b.float().abs().max())
        row = dict(op="decode_attention", shape=name, ours_ms=round(ms_ours, 4), reference_kernel_ms=round(ms_ref, 4),
                   speedup=round(ms_ref / ms_ours, 2), ours_gbs=round(nbytes / ms_ours / 1e6, 1),
                   reference_gbs=round(nbytes / ms_ref / 1e6, 1), max_rel_diff=diff)
        rows.append(row)
        print(json.dumps(row), flush=True)
        del p
        torch.cuda.empty_cache()

    # config 4 movers: OPT-13B, 40 layers, H=40, D=128, B=16, fp16
    layers, H, D, B, NB = 40, 40, 128, 16, 1024
    kcs = [torch.randn((NB, H, D // 8, B, 8), device=DEV).half() for _ in range(layers)]
    vcs = [torch.randn((NB, H, D, B), device=DEV).half() for _ in range(layers)]
    mapping = {}
    for i in range(192):
        mapping.setdefault(i // 3, []).append(512 + i)
    nbytes = 2 * 192 * layers * 2 * H * D * B * 2
    a = time_calls(lambda: cache_ops.copy_blocks(kcs, vcs, mapping), iters=20)
    b = time_calls(lambda: ref_cache.copy_blocks(kcs, vcs, mapping), iters=10, warm=2)
    rows.append(dict(op="copy_blocks", shape="40 layers x 192 pairs x 163840 B", ours_ms=round(a, 4),
                     reference_kernel_ms=round(b, 4), speedup=round(b / a, 2), ours_gbs=round(nbytes / a / 1e6, 1)))
    print(json.dumps(rows[-1]), flush=True)
    for N in (256, 2560):
        qkv = torch.randn((N, 3, H, D), device=DEV).half()
        if N <= 256:     # decode step: every token is the newest token of its own sequence -> its own block
            slots = (torch.randperm(NB, device=DEV)[:N] * B + torch.randint(0, B, (N,), device=DEV)).to(torch.int32)
        else:            # prefill: prompts of 512 tokens fill consecutive slots of scattered block runs
            starts = torch.randperm(NB // 32, device=DEV)[:N // 512] * 32 * B
            slots = (starts[:, None] + torch.arange(512, device=DEV)[None, :]).flatten().to(torch.int32)
        a = time_calls(lambda: cache_ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kcs[0], vcs[0], slots), iters=50)
        b = time_calls(lambda: ref_cache.reshape_and_cache(qkv[:, 1], qkv[:, 2], kcs[0], vcs[0], slots), iters=50)
        ag = time_gpu_only(lambda: cache_ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kcs[0], vcs[0], slots), iters=50)
        bg = time_gpu_only(lambda: ref_cache.reshape_and_cache(qkv[:, 1], qkv[:, 2], kcs[0], vcs[0], slots), iters=50)
        rows.append(dict(op="reshape_and_cache", shape=f"{N} tokens x 40 heads x 128", ours_ms=round(a, 4),
                         reference_kernel_ms=round(b, 4), speedup=round(b / a, 2), ours_gpu_only_ms=round(ag, 4),
                         reference_gpu_only_ms=round(bg, 4), gpu_only_speedup=round(bg / ag, 2)))
        print(json.dumps(rows[-1]), flush=True)
    # rotary embedding + cache write: the reference runs two kernels (pos_encoding_kernels.cu, cache_kernels.cu),
    # ours one fused launch
    try:
        ref_pos = load("ref_pos_encoding_ops")
    except Exception:
        ref_pos = None
    if ref_pos is not None:
        from oracle import paged_oracle
        from vllm_b200 import pos_encoding_ops
        cs = paged_oracle.rope_cos_sin_cache(D, 8192, torch.float16).to(DEV)
        for N, label in ((256, "decode step, 256 tokens"), (2560, "prefill, 2560 tokens")):
            qkv = torch.randn((N, 3, H * D), device=DEV).half()
            pos = torch.randint(0, 8192, (N,), device=DEV)
            if N <= 256:
                slots = (torch.randperm(NB, device=DEV)[:N] * B + torch.randint(0, B, (N,), device=DEV)).to(torch.int32)
            else:
                starts = torch.randperm(NB // 32, device=DEV)[:N // 512] * 32 * B
                slots = (starts[:, None] + torch.arange(512, device=DEV)[None, :]).flatten().to(torch.int32)
            k3, v3 = qkv[:, 1].view(N, H, D), qkv[:, 2].view(N, H, D)

            def ref_two():
                ref_pos.rotary_embedding_neox(pos, qkv[:, 0], qkv[:, 1], D, cs)
                ref_cache.reshape_and_cache(k3, v3, kcs[0], vcs[0], slots)

            def ours_two():
                pos_encoding_ops.rotary_embedding_neox(pos, qkv[:, 0], qkv[:, 1], D, cs)
                cache_ops.reshape_and_cache(k3, v3, kcs[0], vcs[0], slots)

            def ours_fused():
                pos_encoding_ops.rotary_embedding_neox_reshape_and_cache(pos, qkv[:, 0], qkv[:, 1], qkv[:, 2], kcs[0],
                                                                         vcs[0], slots, D, cs)
            r = time_gpu_only(ref_two, iters=50)
            o2 = time_gpu_only(ours_two, iters=50)
            of = time_gpu_only(ours_fused, iters=50)
            rows.append(dict(op="rotary_embedding_neox + reshape_and_cache", shape=f"{label} x 40 heads x 128",
                             reference_two_kernels_gpu_ms=round(r, 4), ours_two_kernels_gpu_ms=round(o2, 4),
                             ours_fused_gpu_ms=round(of, 4), speedup_fused_vs_reference=round(r / of, 2)))
            print(json.dumps(rows[-1]), flush=True)
    cpu = torch.empty((NB, H, D, B), dtype=torch.float16).pin_memory()
    perm_s, perm_d = torch.randperm(NB)[:512].tolist(), torch.randperm(NB)[:512].tolist()
    mp = {int(s): int(d) for s, d in zip(perm_s, perm_d)}
    for direction, (src, dst) in (("gpu->cpu", (vcs[0], cpu)), ("cpu->gpu", (cpu, vcs[0]))):
        a = time_calls(lambda: cache_ops.swap_blocks(src, dst, mp), iters=10, warm=2)
        b = time_calls(lambda: ref_cache.swap_blocks(src, dst, mp), iters=10, warm=2)
        rows.append(dict(op="swap_blocks", shape=f"512 scattered blocks x 163840 B {direction}", ours_ms=round(a, 4),
                         reference_kernel_ms=round(b, 4), speedup=round(b / a, 2),
                         ours_gbs=round(512 * H * D * B * 2 / a / 1e6, 1)))
        print(json.dumps(rows[-1]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "compare_reference.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
