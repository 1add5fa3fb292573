"""Bit-exact parity of the KV-cache movers (cache_ops -> C ABI -> sm_100a kernels) with the
reference's own test references (tests/kernels/test_cache.py:51-56,95-100,138-143 via the
committed golden hashes) and with the oracle at larger, ragged and benchmark sizes."""
import hashlib
import json
import os
import random

import pytest
import torch

import cases
import pa_testing as pt
from oracle import c_oracle, paged_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = os.path.join(os.path.dirname(__file__), "golden")
SHA_OUT = json.load(open(os.path.join(GOLD, "cache_outputs.sha256.json")))


def _sha(t):
    return hashlib.sha256(cases.to_numpy(t.cpu()).tobytes()).hexdigest()


def _same_bits(a, b):
    return torch.equal(pt.bits(a), pt.bits(b))


@pytest.mark.parametrize("cid", cases.cache_case_ids())
def test_golden_reference_outputs(cid):
    from vllm_b200 import cache_ops
    c = cases.cache_case(cid)
    qkv = c["key"]._base.to(DEV)
    key, value = qkv[:, 1], qkv[:, 2]                      # strided rows, as test_cache.py:79-81
    slots = c["slot_mapping"].to(DEV)
    kc, vc = c["key_caches"][0].to(DEV), c["value_caches"][0].to(DEV)
    cache_ops.reshape_and_cache(key, value, kc, vc, slots)
    assert _sha(kc) == SHA_OUT[f"{cid}/reshape_key_cache"]
    assert _sha(vc) == SHA_OUT[f"{cid}/reshape_value_cache"]

    out = torch.zeros_like(qkv)
    cache_ops.gather_cached_kv(out[:, 1], out[:, 2], c["key_caches"][1].to(DEV), c["value_caches"][1].to(DEV), slots)
    assert _sha(out[:, 1]) == SHA_OUT[f"{cid}/gather_key"]
    assert _sha(out[:, 2]) == SHA_OUT[f"{cid}/gather_value"]
    assert torch.count_nonzero(out[:, 0]) == 0            # the q third of the fused buffer is untouched

    kcs, vcs = [t.to(DEV) for t in c["key_caches"]], [t.to(DEV) for t in c["value_caches"]]
    cache_ops.copy_blocks(kcs, vcs, c["copy_map"])
    for l, (k, v) in enumerate(zip(kcs, vcs)):
        assert _sha(k) == SHA_OUT[f"{cid}/copy_key_cache{l}"]
        assert _sha(v) == SHA_OUT[f"{cid}/copy_value_cache{l}"]


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("tokens,heads,head_size,block_size", [
    (3, 2, 16, 8),            # the reference's shape (test_cache.py:157-162)
    (256, 40, 128, 16),       # config 4 decode step
    (2560, 40, 128, 16),      # config 4 prefill (max_num_batched_tokens)
    (77, 7, 80, 32), (1, 1, 64, 8), (513, 3, 96, 16), (97, 5, 64, 16), (512, 2, 128, 8), (96, 3, 80, 8),
])
def test_reshape_and_gather(dtype, tokens, heads, head_size, block_size):
    from vllm_b200 import cache_ops
    g = torch.Generator().manual_seed(tokens)
    x = pt.pack_factor(dtype)
    nb = max(4, (tokens + block_size - 1) // block_size + 3)
    qkv = torch.randn((tokens, 3, heads, head_size), generator=g).to(dtype)
    kc = torch.randn((nb, heads, head_size // x, block_size, x), generator=g).to(dtype)
    vc = torch.randn((nb, heads, head_size, block_size), generator=g).to(dtype)
    slots = torch.randperm(nb * block_size, generator=g)[:tokens].to(torch.int32)
    want_k, want_v = kc.clone(), vc.clone()
    c_oracle.reshape_and_cache(qkv[:, 1], qkv[:, 2], want_k, want_v, slots)
    dq, dk, dv, ds = qkv.to(DEV), kc.to(DEV), vc.to(DEV), slots.to(DEV)
    cache_ops.reshape_and_cache(dq[:, 1], dq[:, 2], dk, dv, ds)
    assert _same_bits(dk, want_k) and _same_bits(dv, want_v)      # untouched slots included

    # gather is the exact inverse
    back = torch.zeros_like(dq)
    cache_ops.gather_cached_kv(back[:, 1], back[:, 2], dk, dv, ds)
    assert _same_bits(back[:, 1], qkv[:, 1]) and _same_bits(back[:, 2], qkv[:, 2])
    # contiguous (non-fused) rows take the same path
    k2, v2 = dq[:, 1].contiguous(), dq[:, 2].contiguous()
    dk2, dv2 = kc.to(DEV), vc.to(DEV)
    cache_ops.reshape_and_cache(k2, v2, dk2, dv2, ds)
    assert torch.equal(dk2, dk) and torch.equal(dv2, dv)
    # the slot-pattern hint picks a kernel (97..512 "scattered" tokens: the per-token element sweep), never the result
    for hint in ("scattered", "runs"):
        dk3, dv3 = kc.to(DEV), vc.to(DEV)
        cache_ops.reshape_and_cache(dq[:, 1], dq[:, 2], dk3, dv3, ds, slots=hint)
        assert torch.equal(dk3, dk) and torch.equal(dv3, dv), hint


@pytest.mark.parametrize("tokens", [128, 512])
def test_reshape_hint_is_only_a_hint(tokens):
    """Consecutive slots declared "scattered" and scattered slots declared "runs": the wrong hint costs time, not bits."""
    from vllm_b200 import cache_ops
    h, d, b, dtype = 8, 128, 16, torch.float16
    g = torch.Generator().manual_seed(tokens)
    nb = tokens // b + 5
    qkv = torch.randn((tokens, 3, h, d), generator=g).to(dtype)
    for slots, hint in ((torch.arange(7, 7 + tokens, dtype=torch.int32), "scattered"),
                        ((torch.randperm(nb, generator=g)[:tokens // b + 1].repeat_interleave(b)[:tokens] * b
                          + torch.arange(tokens) % b).to(torch.int32), "scattered"),
                        (torch.randperm(nb * b, generator=g)[:tokens].to(torch.int32), "runs")):
        wk, wv = torch.zeros(nb, h, d // 8, b, 8, dtype=dtype), torch.zeros(nb, h, d, b, dtype=dtype)
        c_oracle.reshape_and_cache(qkv[:, 1], qkv[:, 2], wk, wv, slots)
        dk, dv = torch.zeros_like(wk, device=DEV), torch.zeros_like(wv, device=DEV)
        dq = qkv.to(DEV)
        cache_ops.reshape_and_cache(dq[:, 1], dq[:, 2], dk, dv, slots.to(DEV), slots=hint)
        assert _same_bits(dk, wk) and _same_bits(dv, wv)
    with pytest.raises(RuntimeError):
        cache_ops.reshape_and_cache(dq[:, 1], dq[:, 2], dk, dv, slots.to(DEV), slots="sorted")


def test_reshape_with_unaligned_rows():
    """key/value rows that are only element-aligned (odd offset into a larger buffer)."""
    from vllm_b200 import cache_ops
    n, h, d, b = 9, 3, 64, 16
    buf = torch.randn(n * 3 * h * d + 1).to(torch.float16).to(DEV)
    qkv = buf[1:].view(n, 3, h, d)
    kc = torch.zeros(4, h, d // 8, b, 8, dtype=torch.float16, device=DEV)
    vc = torch.zeros(4, h, d, b, dtype=torch.float16, device=DEV)
    slots = torch.arange(5, 5 + n, dtype=torch.int32, device=DEV)
    cache_ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, slots)
    wk, wv = torch.zeros_like(kc).cpu(), torch.zeros_like(vc).cpu()
    paged_oracle.reshape_and_cache(qkv[:, 1].cpu(), qkv[:, 2].cpu(), wk, wv, slots.cpu())
    assert _same_bits(kc, wk) and _same_bits(vc, wv)


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
@pytest.mark.parametrize("layers,pairs,heads,head_size,block_size,nb", [
    (7, 23, 17, 16, 8, 1024),        # the reference's shape (test_cache.py:149-154)
    (3, 200, 2, 64, 16, 700),        # more pairs than one launch carries
    (70, 5, 2, 64, 8, 32),           # more layers than one launch carries
    (2, 1, 5, 80, 32, 8),
])
def test_copy_blocks(dtype, layers, pairs, heads, head_size, block_size, nb):
    from vllm_b200 import cache_ops
    rnd = random.Random(pairs)
    x = pt.pack_factor(dtype)
    g = torch.Generator().manual_seed(layers)
    kcs = [torch.randn((nb, heads, head_size // x, block_size, x), generator=g).to(dtype) for _ in range(layers)]
    vcs = [torch.randn((nb, heads, head_size, block_size), generator=g).to(dtype) for _ in range(layers)]
    nsrc = (pairs + 1) // 2
    ids = rnd.sample(range(nb), nsrc + pairs)                 # sources and destinations never alias (H3)
    srcs, dsts = ids[:nsrc], ids[nsrc:]
    mapping = {}
    for i, d in enumerate(dsts):
        mapping.setdefault(srcs[i % nsrc], []).append(d)
    dk, dv = [t.to(DEV) for t in kcs], [t.to(DEV) for t in vcs]
    cache_ops.copy_blocks(dk, dv, mapping)
    paged_oracle.copy_blocks(kcs, vcs, mapping)
    for l in range(layers):
        assert _same_bits(dk[l], kcs[l]) and _same_bits(dv[l], vcs[l]), f"layer {l}"


@pytest.mark.parametrize("engine", [0, 1, 2])
def test_swap_blocks_round_trip_and_directions(engine):
    """swap_blocks is unpinned by the reference (no test there): pinned here on block assignment
    (and on the reference's implementation in test_gpu_vs_reference_kernels.py), for both engines."""
    from vllm_b200 import cache_ops as _ops
    import types
    cache_ops = types.SimpleNamespace(swap_blocks=lambda s, d, m: _ops.swap_blocks(s, d, m, engine=engine))
    nb, heads, d, b = 64, 5, 128, 16
    g = torch.Generator().manual_seed(1)
    gpu = torch.randn((nb, heads, d, b), generator=g).to(torch.float16).to(DEV)
    cpu = torch.randn((nb, heads, d, b), generator=g).to(torch.float16).pin_memory()
    gpu0, cpu0 = gpu.clone(), cpu.clone()
    out_map = {3: 7, 4: 8, 5: 9, 20: 0, 63: 62, 10: 11, 30: 33, 40: 35, 50: 37}   # (3,4,5)->(7,8,9) merges into one DMA
    cache_ops.swap_blocks(gpu, cpu, out_map)                      # GPU -> CPU
    torch.cuda.synchronize()
    want = cpu0.clone()
    paged_oracle.swap_blocks(gpu0.cpu(), want, out_map)
    assert _same_bits(cpu, want)
    in_map = {v: k for k, v in out_map.items()}
    gpu.zero_()
    cache_ops.swap_blocks(cpu, gpu, in_map)                       # CPU -> GPU
    for s_blk in out_map:
        assert torch.equal(gpu[s_blk], gpu0[s_blk])
    untouched = [i for i in range(nb) if i not in out_map]
    assert torch.count_nonzero(gpu[untouched]) == 0
    other = torch.zeros_like(gpu)
    cache_ops.swap_blocks(gpu0, other, {1: 2, 2: 1, 40: 41})      # GPU -> GPU
    assert torch.equal(other[2], gpu0[1]) and torch.equal(other[1], gpu0[2]) and torch.equal(other[41], gpu0[40])
    pageable = torch.zeros((nb, heads, d, b), dtype=torch.float16)
    cache_ops.swap_blocks(gpu0, pageable, {9: 1})                 # pageable host memory also works
    torch.cuda.synchronize()
    assert torch.equal(pageable[1], gpu0[9].cpu())
    many = {i: (i * 7) % nb for i in range(0, nb, 2)}             # > 6 runs: the automatic choice is the kernel
    pageable.zero_()
    cache_ops.swap_blocks(gpu0, pageable, many)                   # ... which falls back to DMA for pageable memory
    torch.cuda.synchronize()
    for s_blk, d_blk in many.items():
        assert torch.equal(pageable[d_blk], gpu0[s_blk].cpu())
    with pytest.raises(RuntimeError, match="Invalid device combination"):
        _ops.swap_blocks(pageable, pageable.clone(), {0: 1})


def test_reshape_and_cache_prefill_runs_and_padding_tokens():
    """Consecutive slots (a prompt filling its blocks in order), runs crossing block boundaries, a ragged
    last tile, and negative slots (skipped)."""
    from vllm_b200 import cache_ops
    # 203 tokens take the direct kernel (97..512 tokens), 1015 the tile mover with its run-aware lane order
    for dtype, rep in ((torch.float16, 1), (torch.float32, 1), (torch.float16, 5), (torch.float32, 5)):
        x = pt.pack_factor(dtype)
        n, h, d, b, nb = 203 * rep, 3, 128, 16, 40 * rep
        g = torch.Generator().manual_seed(7)
        qkv = torch.randn((n, 3, h, d), generator=g).to(dtype)
        kc = torch.randn((nb, h, d // x, b, x), generator=g).to(dtype)
        vc = torch.randn((nb, h, d, b), generator=g).to(dtype)
        # per repetition three prompts: 100 tokens from slot 37, 60 from slot 320, 43 scattered; a few padding tokens
        slots = torch.cat([torch.cat([torch.arange(37, 137), torch.arange(320, 380),
                                      torch.randperm(200, generator=g)[:43] + 400]) + 640 * r
                           for r in range(rep)]).to(torch.int32)
        slots[[5, 101, 202]] = -1
        wk, wv = kc.clone(), vc.clone()
        c_oracle.reshape_and_cache(qkv[:, 1], qkv[:, 2], wk, wv, slots)
        dq, dk, dv = qkv.to(DEV), kc.to(DEV), vc.to(DEV)
        cache_ops.reshape_and_cache(dq[:, 1], dq[:, 2], dk, dv, slots.to(DEV))
        assert _same_bits(dk, wk) and _same_bits(dv, wv)
        back = torch.zeros_like(dq)
        cache_ops.gather_cached_kv(back[:, 1], back[:, 2], dk, dv, slots.to(DEV))
        want = torch.zeros_like(qkv)
        c_oracle.gather_cached_kv(want[:, 1], want[:, 2], wk, wv, slots)
        assert _same_bits(back, want)
