"""Parity of the sm_100a decode-attention kernel with the oracle, through the operator shim
(attention_ops -> C ABI).  Mirrors the reference's tests/kernels/test_attention.py:159-287
(shape matrix, strided query, repeated blocks) and adds what that test leaves out: unit-scale
inputs, block-boundary / single-token / empty contexts, garbage in the masked tail (hazard H2),
over-long block-table rows (H4), element offsets beyond 2^31 (H1), every launch shape, and the
BASELINE.json benchmark sizes through row samples and size-independent properties."""
import os
import random

import numpy as np
import pytest
import torch

import cases
import pa_testing as pt

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

GOLD = os.path.join(os.path.dirname(__file__), "golden")
DEV = "cuda:0"


def _gpu_case(cid):
    c = cases.attention_case(cid)
    qkv = c["query"]._base.to(DEV)
    return c, pt.DecodeProblem(qkv[:, 0], c["key_cache"].to(DEV), c["value_cache"].to(DEV),
                               c["block_tables"].to(DEV), c["context_lens"].to(DEV), c["scale"],
                               c["block_size"], c["max_context_len"])


# ---------------------------------------------------------------------------
# 1. the committed golden vectors: outputs of the reference's own PyTorch reference
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("variant", [1, 2])
@pytest.mark.parametrize("cid", cases.attention_case_ids())
def test_golden_reference_outputs(cid, variant):
    att = np.load(os.path.join(GOLD, "attention_golden.npz"))
    c, p = _gpu_case(cid)
    got = pt.run_decode(p, tuning=pt.tuning(variant=variant))
    gold = cases.from_numpy(att[cid], c["dtype"])
    # tolerance of the north star (1e-3 fp16, 1e-2 bf16), relative to the output scale
    pt.assert_decode_close(got, gold, c["dtype"], f"{cid} v{variant} vs reference output")
    # the north star's tolerance flat (atol = rtol = 1e-3 fp16 / 1e-2 bf16), not scaled by the output magnitude
    pt.assert_flat_north_star(got, gold, c["dtype"], f"{cid} v{variant}")
    # and by the letter of the reference test (atol=1e-3, rtol=1e-5; test_attention.py:224)
    if cid.endswith("-ref"):
        assert torch.allclose(got.float().cpu(), gold.float(), atol=1e-3, rtol=1e-5)
    # against the double-precision C oracle, same tolerance
    cq = pt.DecodeProblem(c["query"], c["key_cache"], c["value_cache"], c["block_tables"], c["context_lens"],
                          c["scale"], c["block_size"], c["max_context_len"])
    pt.assert_decode_close(got, pt.oracle_decode(cq), c["dtype"], f"{cid} v{variant} vs C oracle")


@pytest.mark.parametrize("variant", [0, 1, 2, 3])
@pytest.mark.parametrize("cid", cases.attention_large_case_ids())
def test_large_goldens_config1_and_long_contexts(cid, variant):
    """BASELINE.json configs[0] as the reference test builds it (7 x 40 x 128, block 16, fp32, 1024 aliasing
    blocks, contexts <= 4096 from random.seed(0)) and 4096-token contexts per dtype, against outputs of the
    reference's own function at those sizes (tests/golden/attention_golden_large.npz)."""
    att = np.load(os.path.join(GOLD, "attention_golden_large.npz"))
    c = cases.attention_large_case(cid)
    qkv = c["query"]._base.to(DEV)
    p = pt.DecodeProblem(qkv[:, 0], c["key_cache"].to(DEV), c["value_cache"].to(DEV), c["block_tables"].to(DEV),
                         c["context_lens"].to(DEV), c["scale"], c["block_size"], c["max_context_len"])
    got = pt.run_decode(p, tuning=pt.tuning(variant=variant))
    gold = cases.from_numpy(att[cid], c["dtype"])
    pt.assert_decode_close(got, gold, c["dtype"], f"{cid} v{variant} vs reference output")
    pt.assert_flat_north_star(got, gold, c["dtype"], f"{cid} v{variant}")
    if cid.endswith("-ref"):
        # the reference test's own tolerance (test_attention.py:224)
        assert torch.allclose(got.float().cpu(), gold.float(), atol=1e-3, rtol=1e-5)
    if c["dtype"] == torch.float32:
        # fp32 is held to 1e-4 relative to the output scale on top of the letter
        assert float((got.float().cpu() - gold.float()).abs().max()) <= 1e-4 * float(gold.float().abs().max())


# ---------------------------------------------------------------------------
# 2. the reference's shape matrix (37 sequences, 3 heads... test_attention.py:270-287)
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("head_size", [64, 80, 96, 128])
@pytest.mark.parametrize("dist", ["ref", "unit"])
def test_reference_matrix(dtype, block_size, head_size, dist):
    rnd = random.Random(hash((str(dtype), block_size, head_size)) & 0xffff)
    ctx = [rnd.randint(1, 4096) for _ in range(37)]
    p = pt.make_decode_problem(37, 3, head_size, block_size, 1024, dtype, ctx, dist=dist, seed=block_size + head_size)
    got = pt.run_decode(p.to(DEV))
    want = pt.oracle_decode(p)
    pt.assert_decode_close(got, want, dtype, f"{dtype} B{block_size} D{head_size} {dist}")
    if dist == "ref":
        assert torch.allclose(got.float().cpu(), want.float(), atol=1e-3, rtol=1e-5)


# ---------------------------------------------------------------------------
# 3. every launch shape gives the same answer
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("variant,warps,stages,splits", [
    (1, 1, 0, 1), (1, 4, 0, 1), (1, 8, 0, 3), (2, 1, 1, 1), (2, 2, 2, 1), (2, 4, 3, 1), (2, 4, 2, 2),
    (2, 8, 2, 5), (2, 4, 4, 16), (1, 2, 0, 7), (2, 1, 8, 1),
])
def test_launch_shapes(dtype, variant, warps, stages, splits):
    ctx = [1, 15, 16, 17, 31, 32, 33, 255, 256, 257, 777, 1500, 2048]
    p = pt.make_decode_problem(len(ctx), 5, 128, 16, 512, dtype, ctx, dist="unit", seed=3, extra_table_cols=3)
    want = pt.oracle_decode(p)
    got = pt.run_decode(p.to(DEV), tuning=pt.tuning(variant, warps, stages, splits))
    pt.assert_decode_close(got, want, dtype, f"v{variant} w{warps} s{stages} split{splits}")


@pytest.mark.parametrize("block_size,head_size", [(8, 64), (8, 80), (32, 96), (32, 128), (16, 80), (16, 96)])
@pytest.mark.parametrize("variant,splits", [(1, 1), (2, 1), (2, 3)])
def test_launch_shapes_other_geometries(block_size, head_size, variant, splits):
    for dtype in (torch.float16, torch.float32):
        ctx = [1, block_size - 1, block_size, block_size + 1, 5 * block_size, 333, 1025]
        p = pt.make_decode_problem(len(ctx), 4, head_size, block_size, 300, dtype, ctx, dist="unit", seed=11)
        got = pt.run_decode(p.to(DEV), tuning=pt.tuning(variant, 0, 0, splits))
        pt.assert_decode_close(got, pt.oracle_decode(p), dtype, f"B{block_size} D{head_size} v{variant} x{splits}")


def test_repeated_calls_reuse_split_workspace():
    """The split-context tickets re-arm themselves: back-to-back calls on one stream agree."""
    p = pt.make_decode_problem(6, 4, 128, 16, 900, torch.float16, [2048, 100, 4000, 16, 3000, 1], seed=5)
    want = pt.oracle_decode(p)
    pd = p.to(DEV)
    for _ in range(4):
        got = pt.run_decode(pd, tuning=pt.tuning(2, 4, 2, 6))
        pt.assert_decode_close(got, want, torch.float16, "repeat")
    # a different item count right after (workspace grows)
    p2 = pt.make_decode_problem(40, 9, 128, 16, 900, torch.float16, [700] * 40, seed=6)
    pt.assert_decode_close(pt.run_decode(p2.to(DEV), tuning=pt.tuning(2, 4, 2, 4)), pt.oracle_decode(p2),
                           torch.float16, "grown")
    # and on a side stream
    s = torch.cuda.Stream(device=DEV)
    with torch.cuda.stream(s):
        got = pt.run_decode(pd, tuning=pt.tuning(2, 4, 2, 6))
    s.synchronize()
    pt.assert_decode_close(got, want, torch.float16, "side stream")


# ---------------------------------------------------------------------------
# 4. edge cases the reference tests never reach
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("variant,splits", [(1, 1), (2, 1), (2, 4)])
def test_garbage_in_masked_tail_is_harmless(dtype, variant, splits):
    """H2: slots past context_len hold NaN/Inf (never-written cache memory); the output must not see them."""
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
    got = pt.run_decode(pd, tuning=pt.tuning(variant, 0, 0, splits))
    pt.assert_decode_close(got, want, dtype, "garbage tail")


@pytest.mark.parametrize("variant,splits", [(1, 1), (2, 1), (2, 3)])
def test_zero_context_rows_are_zero(variant, splits):
    ctx = [40, 0, 300, 0]
    p = pt.make_decode_problem(4, 2, 64, 8, 64, torch.float16, ctx, seed=2)
    pd = p.to(DEV)
    out = torch.full(p.query.shape, 7.0, dtype=torch.float16, device=DEV)
    from vllm_b200 import attention_ops
    attention_ops.single_query_cached_kv_attention(out, pd.query, pd.key_cache, pd.value_cache, pd.scale,
                                                   pd.block_tables, pd.context_lens, 8, 300,
                                                   tuning=pt.tuning(variant, 0, 0, splits))
    assert torch.count_nonzero(out[1]) == 0 and torch.count_nonzero(out[3]) == 0
    pt.assert_decode_close(out, pt.oracle_decode(p), torch.float16, "zero ctx")


def test_large_logits_do_not_overflow():
    """Scores of +-60 before softmax: the running maximum keeps exp() in range."""
    p = pt.make_decode_problem(4, 2, 128, 16, 64, torch.float16, [500, 64, 33, 1000], seed=4)
    p.query.mul_(6.0)
    p.key_cache.mul_(6.0)
    got = pt.run_decode(p.to(DEV))
    pt.assert_decode_close(got, pt.oracle_decode(p), torch.float16, "large logits")


def test_contiguous_and_strided_query_agree():
    p = pt.make_decode_problem(9, 4, 96, 16, 64, torch.bfloat16, [77] * 9, seed=8)
    pd = p.to(DEV)
    a = pt.run_decode(pd)
    pc = pt.DecodeProblem(pd.query.contiguous(), pd.key_cache, pd.value_cache, pd.block_tables, pd.context_lens,
                          pd.scale, pd.block_size, pd.max_context_len)
    assert torch.equal(a, pt.run_decode(pc))
    # a query whose rows are only element-aligned (odd element offset into a larger buffer)
    buf = torch.zeros(9 * 3 * 4 * 96 + 1, dtype=torch.bfloat16, device=DEV)
    q_odd = buf[1:].view(9, 3, 4, 96)[:, 0]
    q_odd.copy_(pd.query)
    po = pt.DecodeProblem(q_odd, pd.key_cache, pd.value_cache, pd.block_tables, pd.context_lens, pd.scale,
                          pd.block_size, pd.max_context_len)
    assert torch.equal(a, pt.run_decode(po))


def test_offsets_beyond_2_31_elements():
    """H1: NB*H*D*B > 2^31 elements (the reference's int offsets overflow there)."""
    H, D, B = 40, 128, 16
    nb_big = 27000                                  # 27000*40*128*16 = 2.2e9 elements per cache (4.4 GB fp16)
    ctx = [300, 1000, 17]
    p = pt.make_decode_problem(3, H, D, B, 96, torch.float16, ctx, seed=12, unique_blocks=True)
    want = pt.oracle_decode(p)
    kc = torch.zeros((nb_big, H, D // 8, B, 8), dtype=torch.float16, device=DEV)
    vc = torch.zeros((nb_big, H, D, B), dtype=torch.float16, device=DEV)
    remap = torch.arange(96, dtype=torch.int64) * 7 + (nb_big - 96 * 7)      # all beyond element 2^31
    assert int(remap.min()) * H * D * B > 2 ** 31
    kc[remap.to(DEV)] = p.key_cache.to(DEV)
    vc[remap.to(DEV)] = p.value_cache.to(DEV)
    tables = remap[p.block_tables.long()].to(torch.int32)
    pd = pt.DecodeProblem(p.query._base.to(DEV)[:, 0], kc, vc, tables.to(DEV), p.context_lens.to(DEV), p.scale, B,
                          p.max_context_len)
    for variant in (1, 2):
        pt.assert_decode_close(pt.run_decode(pd, tuning=pt.tuning(variant)), want, torch.float16, f"H1 v{variant}")
    del kc, vc
    torch.cuda.empty_cache()


def test_shape_and_dtype_errors_surface_as_runtime_error():
    from vllm_b200 import attention_ops
    p = pt.make_decode_problem(2, 2, 64, 16, 8, torch.float16, [20, 20]).to(DEV)
    out = torch.empty_like(p.query.contiguous())
    with pytest.raises(RuntimeError):       # int64 tables: data_ptr<int>() throws in the reference
        attention_ops.single_query_cached_kv_attention(out, p.query, p.key_cache, p.value_cache, p.scale,
                                                       p.block_tables.long(), p.context_lens, 16, 20)
    q112 = torch.zeros(2, 2, 112, dtype=torch.float16, device=DEV)
    kc = torch.zeros(8, 2, 14, 16, 8, dtype=torch.float16, device=DEV)
    vc = torch.zeros(8, 2, 112, 16, dtype=torch.float16, device=DEV)
    with pytest.raises(RuntimeError, match="head size"):        # attention_kernels.cu:398
        attention_ops.single_query_cached_kv_attention(torch.empty_like(q112), q112, kc, vc, 0.1, p.block_tables,
                                                       p.context_lens, 16, 20)
    kc = torch.zeros(8, 2, 8, 4, 8, dtype=torch.float16, device=DEV)
    vc = torch.zeros(8, 2, 64, 4, dtype=torch.float16, device=DEV)
    with pytest.raises(RuntimeError, match="block size"):       # :446
        attention_ops.single_query_cached_kv_attention(out, p.query, kc, vc, 0.1, p.block_tables, p.context_lens,
                                                       4, 20)
    with pytest.raises(RuntimeError):                           # :467
        attention_ops.single_query_cached_kv_attention(out.double(), p.query.double(), p.key_cache.double(),
                                                       p.value_cache.double(), 0.1, p.block_tables, p.context_lens,
                                                       16, 20)


def test_debug_checks_catch_a_context_longer_than_max_context_len():
    """PA_DEBUG_CHECKS=1 turns the documented silent case (max_context_len < max(context_lens): tail tokens dropped) into
    an error.  The switch is read once per process, so the check runs in a child."""
    import os
    import subprocess
    import sys
    code = (
        "import sys, torch\n"
        f"sys.path[:0] = [{ROOT!r}, {os.path.join(ROOT, 'tests')!r}]\n"
        "import pa_testing as pt\n"
        "p = pt.make_decode_problem(3, 2, 64, 16, 64, torch.float16, [40, 100, 7], seed=1).to('cuda')\n"
        "pt.run_decode(p)\n"                       # max_context_len = 100: fine
        "p.max_context_len = 64\n"
        "try:\n"
        "    pt.run_decode(p)\n"
        "except RuntimeError as e:\n"
        "    assert 'context_lens[1] = 100' in str(e), e\n"
        "    print('CAUGHT')\n")
    env = dict(os.environ, PA_DEBUG_CHECKS="1")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0 and "CAUGHT" in r.stdout, r.stdout + r.stderr
