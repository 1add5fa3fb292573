"""Prompt (prefill) attention: oracle pinned on the reference's ref_multi_query_kv_attention
(tests/kernels/test_attention.py:69-101, goldens made by tests/golden/make_golden.py), and the
tcgen05 kernel (pa_prefill_attention) against the goldens and the oracle on the GPU."""
import json
import os
import random

import numpy as np
import pytest
import torch

import cases
import pa_testing as pt
from oracle import paged_oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden")
PF = np.load(os.path.join(GOLD, "prefill_golden.npz"))
PF_SHA = json.load(open(os.path.join(GOLD, "prefill_inputs.sha256.json")))

SAME_MATH_TOL = {torch.float16: 2e-3, torch.bfloat16: 1.6e-2, torch.float32: 1e-5}
# north star: 1e-3 fp16, 1e-2 bf16; fp32 path is exact fp32 arithmetic
TOL = {torch.float16: 1e-3, torch.bfloat16: 1e-2, torch.float32: 1e-4}


# ---------------------------------------------------------------------------- CPU: the oracle
@pytest.mark.parametrize("cid", cases.prefill_case_ids())
def test_prefill_inputs_are_the_golden_inputs(cid):
    assert cases.digest(cases.prefill_case(cid)["qkv"]) == PF_SHA[cid]


@pytest.mark.parametrize("cid", cases.prefill_case_ids())
def test_prefill_loop_oracle_matches_reference_output(cid):
    c = cases.prefill_case(cid)
    gold = cases.from_numpy(PF[cid], c["dtype"]).float()
    out = paged_oracle.prefill_attention_loop(c["seq_lens"], c["query"], c["key"], c["value"]).float()
    tol = SAME_MATH_TOL[c["dtype"]]
    torch.testing.assert_close(out, gold, atol=tol * float(gold.abs().max()), rtol=tol)


@pytest.mark.parametrize("cid", cases.prefill_case_ids())
def test_prefill_fp32_oracle_matches_reference_output(cid):
    c = cases.prefill_case(cid)
    gold = cases.from_numpy(PF[cid], c["dtype"]).float()
    truth = paged_oracle.prefill_attention_fp32(c["seq_lens"], c["query"], c["key"], c["value"], c["scale"])
    tol = {torch.float16: 2e-3, torch.bfloat16: 1.6e-2, torch.float32: 1e-5}[c["dtype"]]
    torch.testing.assert_close(truth, gold, atol=tol * float(gold.abs().max()), rtol=tol)


def test_prefill_symbols_exported():
    from vllm_b200 import _cabi
    lib = _cabi.lib()
    assert hasattr(lib, "pa_prefill_attention") and hasattr(lib, "pa_prefill_attention_debug")


# ---------------------------------------------------------------------------- GPU
def _run(c, dev="cuda"):
    from vllm_b200 import attention_ops
    qkv = c["qkv"].to(dev)
    q, k, v = qkv[:, 0], qkv[:, 1], qkv[:, 2]
    out = torch.full(q.shape, float("nan"), dtype=c["dtype"], device=dev)
    attention_ops.multi_query_kv_attention(out, q, k, v, c["seq_lens"], c["scale"])
    torch.cuda.synchronize()
    return out.cpu()


def _close(got, want, dtype, what=""):
    tol = TOL[dtype]
    got, want = got.float(), want.float()
    assert torch.isfinite(got).all(), f"{what}: non-finite output"
    err = (got - want).abs()
    bound = tol * float(want.abs().max()) + tol * want.abs()
    bad = err > bound
    assert not bad.any(), f"{what}: {int(bad.sum())} elements off, max err {float(err.max()):.3e} vs max |want| {float(want.abs().max()):.3e}"


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["bf16", "f16"])
@pytest.mark.parametrize("d", [128, 64])
def test_tensor_core_tiles_raw(dt, d):
    """First CTA's raw tiles: S = Q.K^T of its first K/V tile and the unnormalised O = P.V, against
    plain matmuls -- isolates descriptor / layout errors from softmax errors."""
    from vllm_b200 import attention_ops
    dtype = cases.DTYPES[dt]
    g = torch.Generator().manual_seed(5)
    n, h = 128, 2
    qkv = torch.randn((n, 3, h, d), generator=g).to(dtype)
    dev = "cuda"
    dq = qkv.to(dev)
    out = torch.empty((n, h, d), dtype=dtype, device=dev)
    dbg = torch.zeros(128 * 128 + 128 * d + 256, dtype=torch.float32, device=dev)
    scale = d ** -0.5
    attention_ops.multi_query_kv_attention(out, dq[:, 0], dq[:, 1], dq[:, 2], [n], scale, debug=dbg)
    torch.cuda.synchronize()
    dbg = dbg.cpu()
    q, k, v = qkv[:, 0, 0].float(), qkv[:, 1, 0].float(), qkv[:, 2, 0].float()
    s_want = q @ k.T
    s_got = dbg[:128 * 128].view(128, 128)
    torch.testing.assert_close(s_got, s_want, atol=2e-3 * float(s_want.abs().max()), rtol=1e-3)
    # unnormalised O: P = exp2((s - m) * scale * log2e) rounded to dtype, masked causally
    m = dbg[128 * 128 + 128 * d + 128: 128 * 128 + 128 * d + 256]
    l = dbg[128 * 128 + 128 * d: 128 * 128 + 128 * d + 128]
    sl = s_want * scale * 1.4426950408889634
    sl = sl.masked_fill(torch.triu(torch.ones(128, 128, dtype=torch.bool), diagonal=1), float("-inf"))
    torch.testing.assert_close(m, sl.max(dim=1).values, atol=1e-2, rtol=1e-3)
    p = torch.exp2(sl - m[:, None])
    torch.testing.assert_close(l, p.sum(dim=1), atol=1e-3, rtol=2e-3)
    o_want = p.to(dtype).float() @ v
    o_got = dbg[128 * 128: 128 * 128 + 128 * d].view(128, d)
    torch.testing.assert_close(o_got, o_want, atol=3e-3 * float(o_want.abs().max()), rtol=3e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("cid", cases.prefill_case_ids())
def test_prefill_matches_reference_golden(cid):
    c = cases.prefill_case(cid)
    gold = cases.from_numpy(PF[cid], c["dtype"])
    got = _run(c)
    if cid.endswith("-ref"):
        # the reference test's own criterion (test_attention.py:265): atol 1e-3, rtol 1e-5 -- and ours
        assert torch.allclose(got.float(), gold.float(), atol=1e-3, rtol=1e-5)
    _close(got, gold, c["dtype"], cid)


@pytest.mark.gpu
@pytest.mark.parametrize("cid", cases.prefill_case_ids())
def test_prefill_matches_fp32_oracle(cid):
    c = cases.prefill_case(cid)
    truth = paged_oracle.prefill_attention_fp32(c["seq_lens"], c["query"], c["key"], c["value"], c["scale"])
    _close(_run(c), truth, c["dtype"], cid)


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f16", "bf16", "f32"])
@pytest.mark.parametrize("d", [64, 80, 96, 128])
def test_prefill_reference_test_shape(dt, d):
    """The reference test's own matrix (test_attention.py:290-303): 5 prompts of random length
    below 4096 (here below 1200 to keep the CPU oracle quick), 3 heads, U(+-1e-3) inputs,
    checked at its atol 1e-3 / rtol 1e-5 and at the north-star tolerance on unit-scale inputs."""
    dtype = cases.DTYPES[dt]
    rnd = random.Random(d * 7 + len(dt))
    seq_lens = rnd.sample(range(1, 1200), 5)
    n, h = sum(seq_lens), 3
    for dist, gen_seed in (("ref", 1), ("unit", 2)):
        g = torch.Generator().manual_seed(gen_seed)
        qkv = torch.rand((n, 3, h, d), generator=g) * 2e-3 - 1e-3 if dist == "ref" else torch.randn((n, 3, h, d), generator=g)
        qkv = qkv.to(dtype)
        c = {"qkv": qkv, "seq_lens": seq_lens, "scale": d ** -0.5, "dtype": dtype}
        got = _run(c)
        truth = paged_oracle.prefill_attention_fp32(seq_lens, qkv[:, 0], qkv[:, 1], qkv[:, 2], c["scale"])
        if dist == "ref":
            assert torch.allclose(got.float(), truth, atol=1e-3, rtol=1e-5)
        _close(got, truth, dtype, f"{dt} d{d} {dist} lens {seq_lens}")


@pytest.mark.gpu
def test_prefill_padding_tokens_untouched_and_errors():
    from vllm_b200 import attention_ops
    dtype = torch.float16
    seq_lens = [5, 200]
    n_valid, n_pad = sum(seq_lens), 3                      # the model pads the token batch to a multiple of 8
    g = torch.Generator().manual_seed(3)
    qkv = torch.randn((n_valid + n_pad, 3, 4, 128), generator=g).to(dtype).cuda()
    out = torch.full((n_valid + n_pad, 4, 128), 7.0, dtype=dtype, device="cuda")
    attention_ops.multi_query_kv_attention(out, qkv[:, 0], qkv[:, 1], qkv[:, 2], seq_lens, 128 ** -0.5)
    torch.cuda.synchronize()
    assert (out[n_valid:] == 7.0).all()
    truth = paged_oracle.prefill_attention_fp32(seq_lens, qkv[:n_valid, 0].cpu(), qkv[:n_valid, 1].cpu(),
                                                qkv[:n_valid, 2].cpu(), 128 ** -0.5)
    _close(out[:n_valid].cpu(), truth, dtype)
    with pytest.raises(RuntimeError):                      # unsupported head size, as the reference's ValueError
        bad = torch.zeros((8, 3, 2, 48), dtype=dtype, device="cuda")
        attention_ops.multi_query_kv_attention(torch.empty((8, 2, 48), dtype=dtype, device="cuda"), bad[:, 0],
                                               bad[:, 1], bad[:, 2], [8], 1.0)
    # nothing to do
    attention_ops.multi_query_kv_attention(out, qkv[:, 0], qkv[:, 1], qkv[:, 2], [], 1.0)


@pytest.mark.gpu
def test_prefill_grid_order_does_not_change_results():
    """A batch whose K/V exceeds L2 runs with tiles fastest in the grid, a smaller one with heads fastest
    (prefill_attention.cu: tile_major).  Prompts are independent, so one call over 8 prompts (114 MB of K/V)
    must equal, bit for bit, two calls over 4 prompts each (57 MB) -- and a sample of rows must match the oracle."""
    from vllm_b200 import attention_ops
    dtype, h, d = torch.bfloat16, 8, 128
    lens = [3500, 3400, 3600, 3500, 3300, 3700, 3500, 3500]
    n = sum(lens)
    g = torch.Generator(device="cuda").manual_seed(11)
    qkv = torch.empty((n, 3, h, d), dtype=dtype, device="cuda").normal_(generator=g)
    scale = d ** -0.5
    whole = torch.empty((n, h, d), dtype=dtype, device="cuda")
    parts = torch.empty_like(whole)
    half = sum(lens[:4])
    # context splits are chosen per batch and change the order of the sums: off for the bit-for-bit comparison
    os.environ["PA_PREFILL_SPLIT"] = "0"
    try:
        attention_ops.multi_query_kv_attention(whole, qkv[:, 0], qkv[:, 1], qkv[:, 2], lens, scale)
        attention_ops.multi_query_kv_attention(parts[:half], qkv[:half, 0], qkv[:half, 1], qkv[:half, 2], lens[:4], scale)
        attention_ops.multi_query_kv_attention(parts[half:], qkv[half:, 0], qkv[half:, 1], qkv[half:, 2], lens[4:], scale)
        torch.cuda.synchronize()
    finally:
        del os.environ["PA_PREFILL_SPLIT"]
    assert torch.equal(whole.view(torch.int16), parts.view(torch.int16))
    # and with the planner free to split (it does on this batch: 8 heads leave most of the GPU idle)
    split = torch.empty_like(whole)
    attention_ops.multi_query_kv_attention(split, qkv[:, 0], qkv[:, 1], qkv[:, 2], lens, scale)
    torch.cuda.synchronize()
    assert (split.float() - whole.float()).abs().max() <= 1e-2 * whole.float().abs().max()
    # the last prompt against the fp32 oracle (one prompt keeps the CPU side quick)
    start = n - lens[-1]
    last = qkv[start:].cpu()
    truth = paged_oracle.prefill_attention_fp32([lens[-1]], last[:, 0], last[:, 1], last[:, 2], scale)
    _close(whole[start:].cpu(), truth, dtype, "last prompt of the large batch")
