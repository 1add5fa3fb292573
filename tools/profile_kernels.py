#!/usr/bin/env python
"""One launch of every kernel of the path at BASELINE shapes, for `ncu` (profiles/).
Order of launches (after warm-up launches of the same kernels, skipped with ncu -s):
  decode cfg2 (fp16 h32 L1024), decode cfg3 g8 shard (bf16 h5 L2048), decode cfg5a (fp16 h8 d96 L8192),
  decode short (fp16 h32 L64), copy_blocks cfg4 (40 layers x 192 pairs), reshape_and_cache decode (256 tokens),
  reshape_and_cache prefill (2560 tokens), gather_cached_kv prefill, swap_blocks kernel D2H + H2D (128 blocks)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from vllm_b200 import attention_ops, cache_ops  # noqa: E402

DEV = torch.device("cuda", 0)
REPS = int(sys.argv[1]) if len(sys.argv) > 1 else 2      # each kernel is launched REPS times; profile the last round


def decode(S, H, D, B, L, dt):
    wl = dict(heads=H, head_size=D, block_size=B, seqs=S, ctx=L, dtype=dt)
    p = bench.make_problem(wl, H, DEV)
    out = torch.empty(p["q"].shape, dtype=p["dtype"], device=DEV)
    return lambda: attention_ops.single_query_cached_kv_attention(out, p["q"], p["kc"], p["vc"], p["scale"], p["tables"],
                                                                 p["lens"], B, L)


def main():
    calls = [decode(256, 32, 128, 16, 1024, "fp16"), decode(256, 5, 128, 16, 2048, "bf16"),
             decode(64, 8, 96, 16, 8192, "fp16"), decode(256, 32, 128, 16, 64, "fp16")]
    layers, H, D, B, NB = 40, 40, 128, 16, 1024
    kcs = [torch.randn((NB, H, D // 8, B, 8), device=DEV).half() for _ in range(layers)]
    vcs = [torch.randn((NB, H, D, B), device=DEV).half() for _ in range(layers)]
    mapping = {}
    for i in range(192):
        mapping.setdefault(i // 3, []).append(512 + i)
    calls.append(lambda: cache_ops.copy_blocks(kcs, vcs, mapping))
    qkv_d = torch.randn((256, 3, H, D), device=DEV).half()
    slots_d = (torch.randperm(NB, device=DEV)[:256] * B + torch.randint(0, B, (256,), device=DEV)).to(torch.int32)
    calls.append(lambda: cache_ops.reshape_and_cache(qkv_d[:, 1], qkv_d[:, 2], kcs[0], vcs[0], slots_d))
    qkv_p = torch.randn((2560, 3, H, D), device=DEV).half()
    starts = torch.randperm(NB // 32, device=DEV)[:5] * 32 * B
    slots_p = (starts[:, None] + torch.arange(512, device=DEV)[None, :]).flatten().to(torch.int32)
    calls.append(lambda: cache_ops.reshape_and_cache(qkv_p[:, 1], qkv_p[:, 2], kcs[0], vcs[0], slots_p))
    back = torch.empty_like(qkv_p)
    calls.append(lambda: cache_ops.gather_cached_kv(back[:, 1], back[:, 2], kcs[0], vcs[0], slots_p))
    cpu = torch.empty((NB, H, D, B), dtype=torch.float16).pin_memory()
    sw = {int(s): int(d) for s, d in zip(torch.randperm(NB)[:128].tolist(), torch.randperm(NB)[:128].tolist())}
    calls.append(lambda: cache_ops.swap_blocks(vcs[0], cpu, sw, engine=2))
    calls.append(lambda: cache_ops.swap_blocks(cpu, vcs[0], sw, engine=2))
    for _ in range(REPS):
        for c in calls:
            c()
        torch.cuda.synchronize()
    print("launched", len(calls) * REPS)


if __name__ == "__main__":
    main()
