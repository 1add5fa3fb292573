"""Parity of the persistent work-queue decode kernel (variant 4, csrc/decode_persistent.cuh) with the
oracle and with the grid kernel: the committed goldens of the reference's own function, the reference's shape
matrix, every (warps, stages, chunks) shape including chunks that cut contexts mid-way, ragged batches whose
units are shorter than the ring, empty contexts, garbage tails (H2), offsets beyond 2^31 (H1), over-long
table rows (H4), back-to-back calls (the unit counter and the tickets re-arm themselves)."""
import os
import random

import numpy as np
import pytest
import torch

import cases
import pa_testing as pt

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
DEV = "cuda:0"


@pytest.mark.parametrize("cid", cases.attention_case_ids())
def test_golden_reference_outputs(cid):
    att = np.load(os.path.join(GOLD, "attention_golden.npz"))
    c = cases.attention_case(cid)
    qkv = c["query"]._base.to(DEV)
    p = pt.DecodeProblem(qkv[:, 0], c["key_cache"].to(DEV), c["value_cache"].to(DEV), c["block_tables"].to(DEV),
                         c["context_lens"].to(DEV), c["scale"], c["block_size"], c["max_context_len"])
    gold = cases.from_numpy(att[cid], c["dtype"])
    for chunks in (0, 1, 3):
        got = pt.run_decode(p, tuning=pt.tuning(4, 0, 0, chunks))
        pt.assert_decode_close(got, gold, c["dtype"], f"{cid} persistent x{chunks} vs reference output")
        pt.assert_flat_north_star(got, gold, c["dtype"], f"{cid} persistent x{chunks}")
        if cid.endswith("-ref"):
            assert torch.allclose(got.float().cpu(), gold.float(), atol=1e-3, rtol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("head_size", [64, 80, 96, 128])
def test_reference_matrix(dtype, block_size, head_size):
    rnd = random.Random(hash((str(dtype), block_size, head_size)) & 0xffff)
    ctx = [rnd.randint(1, 4096) for _ in range(37)]
    for dist in ("ref", "unit"):
        p = pt.make_decode_problem(37, 3, head_size, block_size, 1024, dtype, ctx, dist=dist, seed=block_size + head_size)
        got = pt.run_decode(p.to(DEV), tuning=pt.tuning(4))
        want = pt.oracle_decode(p)
        pt.assert_decode_close(got, want, dtype, f"{dtype} B{block_size} D{head_size} {dist}")
        if dist == "ref":
            assert torch.allclose(got.float().cpu(), want.float(), atol=1e-3, rtol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("warps,stages,chunks", [
    (1, 2, 1), (1, 1, 1), (4, 2, 1), (12, 2, 1), (0, 0, 2), (3, 3, 5), (6, 4, 16), (2, 8, 128), (16, 1, 7),
    (5, 2, 64), (0, 0, 0),
])
def test_launch_shapes(dtype, warps, stages, chunks):
    ctx = [1, 15, 16, 17, 31, 32, 33, 255, 256, 257, 777, 1500, 2048, 513, 529]
    p = pt.make_decode_problem(len(ctx), 5, 128, 16, 512, dtype, ctx, dist="unit", seed=3, extra_table_cols=3)
    want = pt.oracle_decode(p)
    got = pt.run_decode(p.to(DEV), tuning=pt.tuning(4, warps, stages, chunks))
    pt.assert_decode_close(got, want, dtype, f"persistent w{warps} s{stages} x{chunks}")


@pytest.mark.parametrize("block_size,head_size", [(8, 64), (8, 80), (32, 96), (32, 128), (16, 80), (16, 96), (8, 128)])
@pytest.mark.parametrize("chunks", [1, 3, 0])
def test_other_geometries(block_size, head_size, chunks):
    for dtype in (torch.float16, torch.float32):
        ctx = [1, block_size - 1, block_size, block_size + 1, 5 * block_size, 333, 1025]
        p = pt.make_decode_problem(len(ctx), 4, head_size, block_size, 300, dtype, ctx, dist="unit", seed=11)
        got = pt.run_decode(p.to(DEV), tuning=pt.tuning(4, 0, 0, chunks))
        pt.assert_decode_close(got, pt.oracle_decode(p), dtype, f"B{block_size} D{head_size} x{chunks}")


def test_more_units_than_warps_and_long_single_chunks():
    """Many units per warp through the dynamic counter; chunks longer than one 32-page id batch."""
    rnd = random.Random(5)
    ctx = [rnd.randint(1, 3000) for _ in range(300)]
    p = pt.make_decode_problem(300, 24, 64, 8, 2000, torch.bfloat16, ctx, dist="unit", seed=21)
    want = pt.oracle_decode(p)
    pd = p.to(DEV)
    for chunks in (1, 2, 9):
        got = pt.run_decode(pd, tuning=pt.tuning(4, 0, 0, chunks))
        pt.assert_decode_close(got, want, torch.bfloat16, f"many units x{chunks}")
    # same bits as the grid kernel when neither splits the context
    a = pt.run_decode(pd, tuning=pt.tuning(4, 0, 0, 1))
    b = pt.run_decode(pd, tuning=pt.tuning(3, 1, 2, 1))
    assert torch.equal(a, b)


def test_repeated_calls_rearm_counters():
    p = pt.make_decode_problem(6, 4, 128, 16, 900, torch.float16, [2048, 100, 4000, 16, 3000, 1], seed=5)
    want = pt.oracle_decode(p)
    pd = p.to(DEV)
    for i in range(6):
        got = pt.run_decode(pd, tuning=pt.tuning(4, 0, 0, 1 + (i % 3) * 3))
        pt.assert_decode_close(got, want, torch.float16, f"repeat {i}")
    p2 = pt.make_decode_problem(40, 9, 128, 16, 900, torch.float16, [700] * 40, seed=6)
    pt.assert_decode_close(pt.run_decode(p2.to(DEV), tuning=pt.tuning(4, 0, 0, 4)), pt.oracle_decode(p2),
                           torch.float16, "grown")
    # interleaved with the grid kernel's splits on the same stream (they share the tickets)
    got = pt.run_decode(pd, tuning=pt.tuning(2, 4, 2, 6))
    pt.assert_decode_close(got, want, torch.float16, "grid after persistent")
    got = pt.run_decode(pd, tuning=pt.tuning(4, 0, 0, 6))
    pt.assert_decode_close(got, want, torch.float16, "persistent after grid")
    s = torch.cuda.Stream(device=DEV)
    with torch.cuda.stream(s):
        got = pt.run_decode(pd, tuning=pt.tuning(4, 0, 0, 5))
    s.synchronize()
    pt.assert_decode_close(got, want, torch.float16, "side stream")


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("chunks", [1, 4])
def test_garbage_in_masked_tail_is_harmless(dtype, chunks):
    B = 16
    ctx = [1, 5, 16, 17, 100, 250]
    p = pt.make_decode_problem(len(ctx), 3, 128, B, 64, dtype, ctx, dist="unit", seed=9, unique_blocks=True)
    want = pt.oracle_decode(p)
    kc, vc = p.key_cache.clone(), p.value_cache.clone()
    for i, c in enumerate(ctx):
        if c % B:
            blk = int(p.block_tables[i, c // B])
            kc[blk, :, :, c % B:, :] = float("nan")
            vc[blk, :, :, c % B:] = float("inf")
    used = set(int(b) for i, c in enumerate(ctx) for b in p.block_tables[i, :(c + B - 1) // B])
    for blk in range(64):
        if blk not in used:
            kc[blk] = float("nan")
            vc[blk] = float("nan")
    pd = pt.DecodeProblem(p.query._base.to(DEV)[:, 0], kc.to(DEV), vc.to(DEV), p.block_tables.to(DEV),
                          p.context_lens.to(DEV), p.scale, B, p.max_context_len)
    got = pt.run_decode(pd, tuning=pt.tuning(4, 0, 0, chunks))
    pt.assert_decode_close(got, want, dtype, "garbage tail")


@pytest.mark.parametrize("chunks", [1, 3])
def test_zero_context_rows_are_zero(chunks):
    ctx = [40, 0, 300, 0]
    p = pt.make_decode_problem(4, 2, 64, 8, 64, torch.float16, ctx, seed=2)
    pd = p.to(DEV)
    out = torch.full(p.query.shape, 7.0, dtype=torch.float16, device=DEV)
    from vllm_b200 import attention_ops
    attention_ops.single_query_cached_kv_attention(out, pd.query, pd.key_cache, pd.value_cache, pd.scale,
                                                   pd.block_tables, pd.context_lens, 8, 300,
                                                   tuning=pt.tuning(4, 0, 0, chunks))
    assert torch.count_nonzero(out[1]) == 0 and torch.count_nonzero(out[3]) == 0
    pt.assert_decode_close(out, pt.oracle_decode(p), torch.float16, "zero ctx")
    # a batch of nothing but empty contexts
    p0 = pt.make_decode_problem(3, 2, 64, 8, 8, torch.float16, [0, 0, 0], seed=2)
    out0 = pt.run_decode(p0.to(DEV), tuning=pt.tuning(4))
    assert torch.count_nonzero(out0) == 0


def test_element_aligned_query_rows():
    p = pt.make_decode_problem(9, 4, 96, 16, 64, torch.bfloat16, [77] * 9, seed=8)
    pd = p.to(DEV)
    a = pt.run_decode(pd, tuning=pt.tuning(4))
    buf = torch.zeros(9 * 3 * 4 * 96 + 1, dtype=torch.bfloat16, device=DEV)
    q_odd = buf[1:].view(9, 3, 4, 96)[:, 0]
    q_odd.copy_(pd.query)
    po = pt.DecodeProblem(q_odd, pd.key_cache, pd.value_cache, pd.block_tables, pd.context_lens, pd.scale,
                          pd.block_size, pd.max_context_len)
    assert torch.equal(a, pt.run_decode(po, tuning=pt.tuning(4)))
    pt.assert_decode_close(a, pt.oracle_decode(p), torch.bfloat16, "odd q")


def test_offsets_beyond_2_31_elements():
    H, D, B = 40, 128, 16
    nb_big = 27000
    ctx = [300, 1000, 17]
    p = pt.make_decode_problem(3, H, D, B, 96, torch.float16, ctx, seed=12, unique_blocks=True)
    want = pt.oracle_decode(p)
    kc = torch.zeros((nb_big, H, D // 8, B, 8), dtype=torch.float16, device=DEV)
    vc = torch.zeros((nb_big, H, D, B), dtype=torch.float16, device=DEV)
    remap = torch.arange(96, dtype=torch.int64) * 7 + (nb_big - 96 * 7)
    kc[remap.to(DEV)] = p.key_cache.to(DEV)
    vc[remap.to(DEV)] = p.value_cache.to(DEV)
    tables = remap[p.block_tables.long()].to(torch.int32)
    pd = pt.DecodeProblem(p.query._base.to(DEV)[:, 0], kc, vc, tables.to(DEV), p.context_lens.to(DEV), p.scale, B,
                          p.max_context_len)
    pt.assert_decode_close(pt.run_decode(pd, tuning=pt.tuning(4)), want, torch.float16, "H1 persistent")
    del kc, vc
    torch.cuda.empty_cache()


def test_too_many_sequences_is_an_error_not_a_fallback():
    p = pt.make_decode_problem(4097, 1, 64, 16, 8, torch.float16, [5] * 4097).to(DEV)
    with pytest.raises(RuntimeError, match="sequences"):
        pt.run_decode(p, tuning=pt.tuning(4))
