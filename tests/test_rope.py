"""rotary_embedding_neox (the operator next to the path) and its fusion with
reshape_and_cache.  CPU: both oracle restatements reproduce, bit for bit, the outputs of the
reference's own RefRotaryEmbeddingNeox (tests/golden/rope_*.json, made by make_golden.py from
/root/reference/tests/kernels/test_pos_encoding.py).  GPU: the sm_100a kernels reproduce the same
bits, stand-alone and fused, and the fused launch leaves q, k and both caches exactly as the two
separate operators do."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

import cases
import pa_testing as pt
from oracle import c_oracle, paged_oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden")
SHA_IN = json.load(open(os.path.join(GOLD, "rope_inputs.sha256.json")))
SHA_OUT = json.load(open(os.path.join(GOLD, "rope_outputs.sha256.json")))
SMALL = np.load(os.path.join(GOLD, "rope_golden_small.npz"))
DEV = "cuda:0"


def _sha(t):
    return hashlib.sha256(cases.to_numpy(t.cpu()).tobytes()).hexdigest()


@pytest.mark.parametrize("impl", ["python", "c"])
@pytest.mark.parametrize("cid", cases.rope_case_ids())
def test_oracles_match_reference_output(cid, impl):
    c = cases.rope_case(cid)
    assert cases.digest(c["qkv"], c["positions"]) == SHA_IN[cid]
    qkv = c["qkv"].clone()
    cache = paged_oracle.rope_cos_sin_cache(c["rot_dim"], c["max_position"], c["dtype"])
    mod = paged_oracle if impl == "python" else c_oracle
    mod.rotary_embedding_neox(c["positions"], qkv[:, 0], qkv[:, 1], c["head_size"], cache)
    assert _sha(qkv[:, 0]) == SHA_OUT[f"{cid}/query"], "query differs from the reference output"
    assert _sha(qkv[:, 1]) == SHA_OUT[f"{cid}/key"], "key differs from the reference output"
    assert torch.equal(qkv[:, 2], c["qkv"][:, 2])                       # the value third is untouched
    if f"{cid}/query" in SMALL.files:
        assert np.array_equal(cases.to_numpy(qkv[:, 0]), SMALL[f"{cid}/query"])


@pytest.mark.gpu
@pytest.mark.parametrize("cid", cases.rope_case_ids())
def test_gpu_matches_reference_output_bit_for_bit(cid):
    from vllm_b200 import pos_encoding_ops
    c = cases.rope_case(cid)
    qkv = c["qkv"].to(DEV)
    cache = paged_oracle.rope_cos_sin_cache(c["rot_dim"], c["max_position"], c["dtype"]).to(DEV)
    pos_encoding_ops.rotary_embedding_neox(c["positions"].to(DEV), qkv[:, 0], qkv[:, 1], c["head_size"], cache)
    assert _sha(qkv[:, 0]) == SHA_OUT[f"{cid}/query"]
    assert _sha(qkv[:, 1]) == SHA_OUT[f"{cid}/key"]
    assert torch.equal(qkv[:, 2].cpu(), c["qkv"][:, 2])


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("tokens,cached,heads,head_size,rot,block_size,pattern", [
    (2145, 2145, 5, 128, 128, 16, "runs"),        # the reference test's token count (test_pos_encoding.py:141)
    (264, 256, 40, 128, 128, 16, "scattered"),    # decode step of 256 sequences, batch padded to a multiple of 8
    (48, 41, 8, 96, 24, 16, "runs"),              # GPT-NeoX: rotary_pct 0.25
    (16, 16, 3, 80, 80, 8, "scattered"), (7, 0, 2, 64, 32, 32, "runs"), (1, 1, 1, 256, 256, 16, "runs"),
])
def test_fused_rope_and_cache_write_equals_the_two_operators(dtype, tokens, cached, heads, head_size, rot, block_size,
                                                             pattern):
    from vllm_b200 import cache_ops, pos_encoding_ops
    x = pt.pack_factor(dtype)
    g = torch.Generator().manual_seed(tokens + heads)
    nb = max(4, (tokens + block_size - 1) // block_size + 3)
    qkv = torch.randn((tokens, 3, heads * head_size), generator=g).to(dtype)
    kc = torch.randn((nb, heads, head_size // x, block_size, x), generator=g).to(dtype)
    vc = torch.randn((nb, heads, head_size, block_size), generator=g).to(dtype)
    positions = torch.randint(0, 4096, (tokens,), generator=g)
    if pattern == "runs":
        slots = ((torch.arange(cached) + 3) % (nb * block_size)).to(torch.int32)
    else:
        slots = torch.randperm(nb * block_size, generator=g)[:cached].to(torch.int32)
    cache = paged_oracle.rope_cos_sin_cache(rot, 4096, dtype)

    # oracle: rotate, then scatter
    want = qkv.clone()
    wk, wv = kc.clone(), vc.clone()
    c_oracle.rotary_embedding_neox(positions, want[:, 0], want[:, 1], head_size, cache)
    if cached:
        c_oracle.reshape_and_cache(want[:cached, 1].view(cached, heads, head_size),
                                   want[:cached, 2].view(cached, heads, head_size), wk, wv, slots)

    # GPU, two operators
    a, ak, av = qkv.to(DEV), kc.to(DEV), vc.to(DEV)
    dpos, dslots, dcache = positions.to(DEV), slots.to(DEV), cache.to(DEV)
    pos_encoding_ops.rotary_embedding_neox(dpos, a[:, 0], a[:, 1], head_size, dcache)
    if cached:
        cache_ops.reshape_and_cache(a[:cached, 1].view(cached, heads, head_size),
                                    a[:cached, 2].view(cached, heads, head_size), ak, av, dslots)
    # GPU, fused
    b, bk, bv = qkv.to(DEV), kc.to(DEV), vc.to(DEV)
    pos_encoding_ops.rotary_embedding_neox_reshape_and_cache(dpos, b[:, 0], b[:, 1], b[:, 2], bk, bv, dslots, head_size,
                                                             dcache)
    for name, got in (("two operators", (a, ak, av)), ("fused", (b, bk, bv))):
        assert torch.equal(pt.bits(got[0]), pt.bits(want)), f"{name}: q/k/v rows"
        assert torch.equal(pt.bits(got[1]), pt.bits(wk)), f"{name}: key cache"
        assert torch.equal(pt.bits(got[2]), pt.bits(wv)), f"{name}: value cache"


@pytest.mark.gpu
def test_rope_errors():
    from vllm_b200 import pos_encoding_ops
    q = torch.zeros(4, 256, dtype=torch.float16, device=DEV)
    cache = torch.zeros(64, 128, dtype=torch.float16, device=DEV)
    pos = torch.zeros(4, dtype=torch.int64, device=DEV)
    with pytest.raises(RuntimeError):
        pos_encoding_ops.rotary_embedding_neox(pos.int(), q, q.clone(), 128, cache)          # int32 positions
    with pytest.raises(RuntimeError):
        pos_encoding_ops.rotary_embedding_neox(pos, q, q.clone(), 100, cache)                # 256 % 100
    with pytest.raises(RuntimeError):
        pos_encoding_ops.rotary_embedding_neox(pos, q, q.clone(), 64, cache)                 # rot_dim 128 > head 64
    with pytest.raises(RuntimeError):
        pos_encoding_ops.rotary_embedding_neox(pos, q.float(), q.clone(), 128, cache)        # mixed dtypes
