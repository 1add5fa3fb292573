"""Randomised soak of the prompt kernels: ragged batches of random shapes through multi_query_kv_attention (checked
against torch SDPA per prompt) and multi_query_cached_kv_attention (checked against the plain kernel over the gathered
rows), so that every mix of item types (whole / partial tiles, one / two groups, split parts) meets every grid order
many times.  A hang is the failure this looks for: run it under `timeout`.

    timeout 300 python tools/prefill_soak.py [iterations] [seed]
"""
from __future__ import annotations

import os
import random
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_b200 import attention_ops, cache_ops  # noqa: E402


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    rnd = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
    dev = "cuda"
    worst = 0.0
    for it in range(iters):
        dtype = rnd.choice([torch.bfloat16, torch.float16])
        d = rnd.choice([64, 80, 96, 128])
        h = rnd.choice([1, 2, 5, 8, 16, 32, 40])
        nseq = rnd.randint(1, 8)
        style = rnd.choice(["short", "mixed", "long", "aligned"])
        lens = []
        for _ in range(nseq):
            if style == "short":
                lens.append(rnd.randint(1, 400))
            elif style == "long":
                lens.append(rnd.randint(1500, 4200))
            elif style == "aligned":
                lens.append(256 * rnd.randint(1, 12))
            else:
                lens.append(rnd.choice([rnd.randint(1, 300), rnd.randint(300, 2000), rnd.randint(2000, 4200)]))
        while sum(lens) * h > 400_000:                       # keep one iteration in the milliseconds
            lens.pop()
            if not lens:
                lens = [256]
                break
        n = sum(lens)
        scale = d ** -0.5
        qkv = torch.empty((n, 3, h, d), dtype=dtype, device=dev).normal_()
        out = torch.empty((n, h, d), dtype=dtype, device=dev)
        os.environ["PA_PREFILL_SPLIT"] = rnd.choice(["0", "1", "1"])
        attention_ops.multi_query_kv_attention(out, qkv[:, 0], qkv[:, 1], qkv[:, 2], lens, scale)
        at = 0
        tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
        for L in lens:
            q, k, v = (qkv[at:at + L, i].transpose(0, 1).unsqueeze(0).float() for i in range(3))
            ref = F.scaled_dot_product_attention(q, k, v, is_causal=True, scale=scale)[0].transpose(0, 1)
            err = float((out[at:at + L].float() - ref).abs().max())
            worst = max(worst, err)
            assert err <= tol, f"iteration {it}: plain, dtype {dtype}, d {d}, h {h}, lens {lens}: max err {err}"
            at += L
        # the same prompts as chunks behind a cached prefix: the last `q` tokens of every prompt are the queries
        b = rnd.choice([8, 16, 32])
        ctx = lens
        qlens = [rnd.randint(1, L) for L in lens]
        nblocks = sum((c + b - 1) // b for c in ctx)
        kc = torch.empty((nblocks, h, d // 8, b, 8), dtype=dtype, device=dev)
        vc = torch.empty((nblocks, h, d, b), dtype=dtype, device=dev)
        kc.fill_(float("nan"))
        vc.fill_(float("nan"))
        mb = max((c + b - 1) // b for c in ctx)
        bt = torch.zeros((len(ctx), mb), dtype=torch.int32)
        perm = torch.randperm(nblocks).int()
        slots, pos, blk = [], 0, 0
        for i, c in enumerate(ctx):
            nb = (c + b - 1) // b
            bt[i, :nb] = perm[blk:blk + nb]
            t = torch.arange(c)
            slots.append(bt[i, t // b].long() * b + t % b)
            blk += nb
        slots = torch.cat(slots).int().to(dev)
        cache_ops.reshape_and_cache(qkv[:, 1].contiguous(), qkv[:, 2].contiguous(), kc, vc, slots)
        rows = torch.cat([torch.arange(at0 + L - ql, at0 + L) for at0, L, ql in
                          zip([sum(lens[:i]) for i in range(len(lens))], lens, qlens)]).to(dev)
        qc = qkv[rows, 0].contiguous()
        oc = torch.empty_like(qc)
        cu = [0]
        for ql in qlens:
            cu.append(cu[-1] + ql)
        os.environ["PA_PREFILL_SPLIT"] = rnd.choice(["0", "1", "1"])
        attention_ops.multi_query_cached_kv_attention(cu, oc, qc, kc, vc, scale, bt.to(dev), ctx, b)
        torch.cuda.synchronize()
        err = float((oc.float() - out[rows].float()).abs().max())
        worst = max(worst, err)
        assert torch.isfinite(oc.float()).all() and err <= tol, \
            f"iteration {it}: cached, dtype {dtype}, d {d}, b {b}, h {h}, pairs {list(zip(qlens, ctx))}: max err {err}"
        if it % 10 == 9:
            print(f"{it + 1} iterations ok, worst abs diff so far {worst:.2e}", flush=True)
    os.environ.pop("PA_PREFILL_SPLIT", None)
    print(f"soak ok: {iters} iterations, worst abs diff {worst:.2e}")


if __name__ == "__main__":
    main()
