"""Prompt chunks over the paged cache (multi_query_cached_kv_attention): oracle pinned on the reference's
ref_multi_query_cached_kv_attention (tests/kernels/test_attention.py:104-156, goldens made by
tests/golden/make_golden.py), and the CUDA path (page gather + tcgen05 prompt kernel with a shifted
diagonal) against the goldens and the oracle on the GPU."""
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
CP = np.load(os.path.join(GOLD, "cached_prefill_golden.npz"))
CP_SHA = json.load(open(os.path.join(GOLD, "cached_prefill_inputs.sha256.json")))

SAME_MATH_TOL = {torch.float16: 2e-3, torch.bfloat16: 1.6e-2, torch.float32: 1e-5}
# north star: 1e-3 fp16, 1e-2 bf16; the fp32 path is exact fp32 arithmetic
TOL = {torch.float16: 1e-3, torch.bfloat16: 1e-2, torch.float32: 1e-4}


# ---------------------------------------------------------------------------- CPU: the oracle
@pytest.mark.parametrize("cid", cases.cached_prefill_case_ids())
def test_inputs_are_the_golden_inputs(cid):
    c = cases.cached_prefill_case(cid)
    assert cases.digest(c["qkv"], c["key_cache"], c["value_cache"], c["block_tables"], c["context_lens"]) == CP_SHA[cid]


@pytest.mark.parametrize("cid", cases.cached_prefill_case_ids())
def test_loop_oracle_matches_reference_output(cid):
    c = cases.cached_prefill_case(cid)
    gold = cases.from_numpy(CP[cid], c["dtype"]).float()
    out = paged_oracle.cached_prefill_attention_loop(c["cu_query_lens"], c["query"], c["key_cache"], c["value_cache"],
                                                     c["block_tables"], c["context_lens"]).float()
    tol = SAME_MATH_TOL[c["dtype"]]
    torch.testing.assert_close(out, gold, atol=tol * float(gold.abs().max()), rtol=tol)


@pytest.mark.parametrize("cid", cases.cached_prefill_case_ids())
def test_fp32_oracle_matches_reference_output(cid):
    c = cases.cached_prefill_case(cid)
    gold = cases.from_numpy(CP[cid], c["dtype"]).float()
    truth = paged_oracle.cached_prefill_attention_fp32(c["cu_query_lens"], c["query"], c["key_cache"],
                                                       c["value_cache"], c["block_tables"], c["context_lens"], c["scale"])
    tol = SAME_MATH_TOL[c["dtype"]]
    torch.testing.assert_close(truth, gold, atol=tol * float(gold.abs().max()), rtol=tol)


def test_oracle_degenerates_to_its_neighbours():
    """query_len == context_len is plain prompt attention; query_len == 1 is decode attention."""
    c = cases.cached_prefill_case("f32-b16-d64-unit".replace("d64", "d128"))
    kc, vc, bt = c["key_cache"], c["value_cache"], c["block_tables"]
    ctx = [70, 33]
    rows = [paged_oracle._context_rows(kc, vc, bt[i], n) for i, n in enumerate(ctx)]
    keys = torch.cat([r[0] for r in rows]); values = torch.cat([r[1] for r in rows])
    q = torch.randn((sum(ctx), 3, 128), generator=torch.Generator().manual_seed(0))
    full = paged_oracle.cached_prefill_attention_fp32([0, 70, 103], q, kc, vc, bt[:2], ctx, c["scale"])
    plain = paged_oracle.prefill_attention_fp32(ctx, q, keys, values, c["scale"])
    torch.testing.assert_close(full, plain, atol=1e-6, rtol=1e-5)
    one = paged_oracle.cached_prefill_attention_fp32([0, 1, 2], q[:2], kc, vc, bt[:2], ctx, c["scale"])
    dec = paged_oracle.decode_attention_fp32(q[:2], kc, vc, bt[:2], torch.tensor(ctx, dtype=torch.int32), c["scale"])
    torch.testing.assert_close(one, dec, atol=1e-6, rtol=1e-5)


def test_symbols_exported():
    from vllm_b200 import _cabi
    lib = _cabi.lib()
    assert hasattr(lib, "pa_multi_query_cached_kv_attention")
    # 16-bit types read the pages directly; fp32 gathers: K rows + V rows (256-byte aligned each) + one int32 slot
    # per context token
    assert lib.pa_multi_query_cached_kv_workspace_bytes(1000, 8, 128, _cabi.PA_F16) == 0
    assert lib.pa_multi_query_cached_kv_workspace_bytes(1000, 8, 128, _cabi.PA_F32) == 2 * 4096000 + 4000
    assert lib.pa_multi_query_cached_kv_workspace_bytes(3, 1, 64, _cabi.PA_F32) == 2 * 768 + 12


# ---------------------------------------------------------------------------- GPU
def _run(c, dev="cuda", out_fill=float("nan")):
    from vllm_b200 import attention_ops
    qkv = c["qkv"].to(dev)
    q = qkv[:, 0]
    out = torch.full(q.shape, out_fill, dtype=c["dtype"], device=dev)
    attention_ops.multi_query_cached_kv_attention(c["cu_query_lens"], out, q, c["key_cache"].to(dev),
                                                  c["value_cache"].to(dev), c["scale"], c["block_tables"].to(dev),
                                                  c["context_lens"], c["block_size"])
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
@pytest.mark.parametrize("cid", cases.cached_prefill_case_ids())
def test_matches_reference_golden(cid):
    c = cases.cached_prefill_case(cid)
    gold = cases.from_numpy(CP[cid], c["dtype"])
    got = _run(c)
    if cid.endswith("-ref"):
        # the reference tests' own criterion (test_attention.py:224,265): atol 1e-3, rtol 1e-5 -- and ours
        assert torch.allclose(got.float(), gold.float(), atol=1e-3, rtol=1e-5)
    _close(got, gold, c["dtype"], cid)


@pytest.mark.gpu
@pytest.mark.parametrize("cid", cases.cached_prefill_case_ids())
def test_matches_fp32_oracle(cid):
    c = cases.cached_prefill_case(cid)
    truth = paged_oracle.cached_prefill_attention_fp32(c["cu_query_lens"], c["query"], c["key_cache"],
                                                       c["value_cache"], c["block_tables"], c["context_lens"], c["scale"])
    _close(_run(c), truth, c["dtype"], cid)


def _random_problem(dtype, d, b, pairs, h=4, seed=0, num_blocks=None):
    rs = np.random.RandomState(seed)
    ctx = [p[1] for p in pairs]
    cu = [0]
    for p in pairs:
        cu.append(cu[-1] + p[0])
    need = sum((c + b - 1) // b for c in ctx)
    num_blocks = num_blocks or need + 3
    x = 16 // torch.empty((), dtype=dtype).element_size()
    qkv = cases._tensor(rs, (cu[-1], 3, h, d), dtype, "unit")
    kc = cases._tensor(rs, (num_blocks, h, d // x, b, x), dtype, "unit")
    vc = cases._tensor(rs, (num_blocks, h, d, b), dtype, "unit")
    mb = max((c + b - 1) // b for c in ctx) + 2
    tables = np.zeros((len(pairs), mb), dtype=np.int32)
    perm = rs.permutation(num_blocks)
    at = 0
    for i, c in enumerate(ctx):
        n = (c + b - 1) // b
        tables[i, :n] = perm[at:at + n]
        at += n
    return {"dtype": dtype, "block_size": b, "head_size": d, "num_heads": h, "scale": d ** -0.5, "qkv": qkv,
            "query": qkv[:, 0], "cu_query_lens": cu, "key_cache": kc, "value_cache": vc,
            "block_tables": torch.from_numpy(tables), "context_lens": torch.tensor(ctx, dtype=torch.int32)}


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f16", "bf16", "f32"])
@pytest.mark.parametrize("d,b", [(128, 16), (64, 8), (80, 32), (96, 16)])
def test_chunked_prompts_of_many_shapes(dt, d, b):
    """Chunks that start and end anywhere relative to the 128-token tiles and 256-row CTAs: the shifted
    diagonal crosses one or two K/V tiles, tile B has as many tiles as tile A or one more."""
    dtype = cases.DTYPES[dt]
    rnd = random.Random(d * 31 + b + len(dt))
    pairs = [(256, 256 + 128), (300, 300 + 1), (129, 129 + 127), (1, 700), (255, 511), (257, 1100), (7, 7)]
    for _ in range(4):
        q = rnd.randint(1, 600)
        pairs.append((q, q + rnd.randint(0, 700)))
    c = _random_problem(dtype, d, b, pairs, seed=d + b)
    truth = paged_oracle.cached_prefill_attention_fp32(c["cu_query_lens"], c["query"], c["key_cache"], c["value_cache"],
                                                       c["block_tables"], c["context_lens"], c["scale"])
    _close(_run(c), truth, dtype, f"{dt} d{d} b{b} {pairs}")


@pytest.mark.gpu
def test_equals_plain_prompt_attention_and_decode():
    """Size-independent properties: with no cached prefix the operator agrees with the prompt kernel over
    the gathered rows, and with one query row per sequence with the decode kernel."""
    from vllm_b200 import attention_ops, cache_ops
    dtype = torch.bfloat16
    pairs = [(513, 513), (64, 64), (1000, 1000)]
    c = _random_problem(dtype, 128, 16, pairs, h=5, seed=9)
    got = _run(c)
    dev = "cuda"
    kc, vc, bt = c["key_cache"].to(dev), c["value_cache"].to(dev), c["block_tables"]
    slots = torch.cat([bt[i, torch.arange(n) // 16].int() * 16 + (torch.arange(n) % 16).int()
                       for i, (n, _) in enumerate(pairs)]).int().to(dev)
    n = slots.numel()
    k = torch.empty((n, 5, 128), dtype=dtype, device=dev)
    v = torch.empty_like(k)
    cache_ops.gather_cached_kv(k, v, kc, vc, slots)
    out = torch.empty_like(k)
    attention_ops.multi_query_kv_attention(out, c["qkv"].to(dev)[:, 0], k, v, [p[0] for p in pairs], c["scale"])
    torch.cuda.synchronize()
    _close(got, out.cpu(), dtype, "no cached prefix vs the prompt kernel over gathered rows")

    ctx = [1, 16, 17, 300, 1025]
    c = _random_problem(dtype, 128, 16, [(1, n) for n in ctx], h=5, seed=10)
    got = _run(c)
    dec = torch.empty((len(ctx), 5, 128), dtype=dtype, device=dev)
    attention_ops.single_query_cached_kv_attention(dec, c["qkv"].to(dev)[:, 0], c["key_cache"].to(dev),
                                                   c["value_cache"].to(dev), c["scale"], c["block_tables"].to(dev),
                                                   c["context_lens"].to(dev), 16, max(ctx))
    torch.cuda.synchronize()
    _close(got, dec.cpu(), dtype, "one-row chunks vs decode kernel")


@pytest.mark.gpu
def test_untouched_rows_unwritten_slots_and_errors():
    from vllm_b200 import attention_ops
    dtype = torch.float16
    pairs = [(5, 40), (200, 333)]
    c = _random_problem(dtype, 128, 16, pairs, seed=3)
    dev = "cuda"
    # slots past each context (never written by the model) hold NaN: they must not reach the output
    kc, vc = c["key_cache"].clone(), c["value_cache"].clone()
    for i, (_, n) in enumerate(pairs):
        last = int(c["block_tables"][i, (n - 1) // 16])
        kc[last, :, :, n % 16 if n % 16 else 16:, :] = float("nan")
        vc[last, :, :, n % 16 if n % 16 else 16:] = float("nan")
    c["key_cache"], c["value_cache"] = kc, vc
    n_valid, n_pad = c["cu_query_lens"][-1], 3
    qkv = torch.cat([c["qkv"], torch.zeros((n_pad, 3, 4, 128), dtype=dtype)]).to(dev)
    out = torch.full((n_valid + n_pad, 4, 128), 7.0, dtype=dtype, device=dev)
    attention_ops.multi_query_cached_kv_attention(c["cu_query_lens"], out, qkv[:, 0], kc.to(dev), vc.to(dev), c["scale"],
                                                  c["block_tables"].to(dev), c["context_lens"], 16)
    torch.cuda.synchronize()
    assert (out[n_valid:] == 7.0).all()
    truth = paged_oracle.cached_prefill_attention_fp32(c["cu_query_lens"], c["query"], kc, vc, c["block_tables"],
                                                       c["context_lens"], c["scale"])
    _close(out[:n_valid].cpu(), truth, dtype)
    with pytest.raises(RuntimeError):      # fewer cached tokens than query tokens
        attention_ops.multi_query_cached_kv_attention([0, 5, 205], out, qkv[:, 0], kc.to(dev), vc.to(dev), c["scale"],
                                                      c["block_tables"].to(dev), [4, 333], 16)
    with pytest.raises(RuntimeError):      # context longer than the block table row
        attention_ops.multi_query_cached_kv_attention([0, 5, 205], out, qkv[:, 0], kc.to(dev), vc.to(dev), c["scale"],
                                                      c["block_tables"].to(dev), [40, 10 ** 6], 16)
    # nothing to do
    attention_ops.multi_query_cached_kv_attention([0], out, qkv[:, 0], kc.to(dev), vc.to(dev), c["scale"],
                                                  c["block_tables"][:0].to(dev), [], 16)


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f16", "bf16"])
@pytest.mark.parametrize("d,b", [(128, 8), (128, 16), (128, 32), (64, 16), (64, 32), (80, 8), (80, 16), (96, 32)])
def test_every_tail_of_every_page_layout(dt, d, b):
    """The page producer hands cache pages to the tensor core as they lie (K: no-swizzle core matrices, V: rows of
    8 / 16 / 32 tokens without / with 32-byte / 64-byte swizzle) and repairs the context's last k-step in shared
    memory.  Contexts ending at every position of a 16-token k-step, in the first and in later tiles, with NaN in
    every slot past each context and in every block no table names: nothing of it may reach the output."""
    dtype = cases.DTYPES[dt]
    ctxs = list(range(1, 34)) + [120 + r for r in range(0, 18)] + [256 + r for r in (1, 7, 8, 9, 15, 16, 17, 24, 25, 31)]
    pairs = [(min(c, 1 + i % 4), c) for i, c in enumerate(ctxs)]
    c = _random_problem(dtype, d, b, pairs, h=2, seed=d * 7 + b)
    kc, vc = c["key_cache"].clone(), c["value_cache"].clone()
    named = torch.zeros(kc.shape[0], dtype=torch.bool)
    for i, (_, n) in enumerate(pairs):
        nb = (n + b - 1) // b
        named[c["block_tables"][i, :nb].long()] = True
        if n % b:
            last = int(c["block_tables"][i, nb - 1])
            kc[last, :, :, n % b:, :] = float("nan")
            vc[last, :, :, n % b:] = float("nan")
    kc[~named] = float("nan")
    vc[~named] = float("nan")
    # padding entries of the tables name block 0: make sure it is somebody's real block or poisoned
    c["key_cache"], c["value_cache"] = kc, vc
    got = _run(c)
    assert torch.isfinite(got.float()).all(), "NaN from an unwritten slot reached the output"
    truth = paged_oracle.cached_prefill_attention_fp32(c["cu_query_lens"], c["query"], kc, vc, c["block_tables"],
                                                       c["context_lens"], c["scale"])
    _close(got, truth, dtype, f"{dt} d{d} b{b}")


@pytest.mark.gpu
def test_grid_order_does_not_change_results():
    """Same property for the paged kernel: 6 chunks over 4000-token contexts, 10 heads (123 MB of K/V: tiles fastest)
    against the same sequences in two calls of 3 (heads fastest), bit for bit."""
    from vllm_b200 import attention_ops
    dtype, h, d, b = torch.float16, 10, 128, 16
    pairs = [(700, 4000), (1, 4000), (4000, 4000), (333, 4096), (2048, 4000), (129, 3900)]
    assert sum(p[1] for p in pairs) * h * d * 4 > 96 << 20
    c = _random_problem(dtype, d, b, pairs, h=h, seed=21)
    dev = "cuda"
    q = c["qkv"].to(dev)[:, 0]
    kc, vc, bt = c["key_cache"].to(dev), c["value_cache"].to(dev), c["block_tables"].to(dev)
    cu, ctx = c["cu_query_lens"], [p[1] for p in pairs]
    whole = torch.empty(q.shape, dtype=dtype, device=dev)
    parts = torch.empty_like(whole)
    # context splits are chosen per batch and change the order of the sums: off for the bit-for-bit comparison
    os.environ["PA_PREFILL_SPLIT"] = "0"
    try:
        attention_ops.multi_query_cached_kv_attention(cu, whole, q, kc, vc, c["scale"], bt, ctx, b)
        for lo, hi in ((0, 3), (3, 6)):
            sub_cu = [v - cu[lo] for v in cu[lo:hi + 1]]
            attention_ops.multi_query_cached_kv_attention(sub_cu, parts[cu[lo]:cu[hi]], q[cu[lo]:cu[hi]], kc, vc, c["scale"],
                                                          bt[lo:hi].contiguous(), ctx[lo:hi], b)
        torch.cuda.synchronize()
    finally:
        del os.environ["PA_PREFILL_SPLIT"]
    assert torch.equal(whole.view(torch.int16), parts.view(torch.int16))
    # the one-token and the 129-token chunks against the fp32 oracle
    for i in (1, 5):
        truth = paged_oracle.cached_prefill_attention_fp32([0, pairs[i][0]], c["query"][cu[i]:cu[i + 1]], c["key_cache"],
                                                           c["value_cache"], c["block_tables"][i:i + 1], [ctx[i]], c["scale"])
        _close(whole[cu[i]:cu[i + 1]].cpu(), truth, dtype, f"chunk {i}")


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f16", "bf16"])
@pytest.mark.parametrize("d,b,h", [(128, 16, 8), (96, 32, 6), (64, 8, 10)])
def test_context_splits_match_the_oracle_and_the_whole_items(dt, d, b, h):
    """A few long items on an otherwise idle GPU are cut over their K/V tiles (plan_splits) and merged by
    prefill_merge_kernel: the result must match the fp32 oracle and, within rounding, the same call with every item
    kept whole; the launch counter shows that the merge kernel really ran."""
    from vllm_b200 import _cabi
    dtype = cases.DTYPES[dt]
    # one-group and two-group items, contexts ending inside a page, a short sequence beside them
    pairs = [(100, 3000 + 7), (256, 2500), (1, 2049), (40, 40)]
    c = _random_problem(dtype, d, b, pairs, h=h, seed=d + b + h)
    truth = paged_oracle.cached_prefill_attention_fp32(c["cu_query_lens"], c["query"], c["key_cache"], c["value_cache"],
                                                       c["block_tables"], c["context_lens"], c["scale"])
    before = _cabi.launch_count()
    got = _run(c)
    launches = _cabi.launch_count() - before
    assert launches == 2, f"expected the prompt kernel and the merge kernel, counted {launches} launches"
    _close(got, truth, dtype, f"split {dt} d{d} b{b}")
    os.environ["PA_PREFILL_SPLIT"] = "0"
    try:
        before = _cabi.launch_count()
        whole = _run(c)
        assert _cabi.launch_count() - before == 1
    finally:
        del os.environ["PA_PREFILL_SPLIT"]
    _close(got, whole, dtype, "split vs whole items")
