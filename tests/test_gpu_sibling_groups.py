"""Sibling groups: sequences of one request (parallel sampling n, beam search) share the physical
pages of their prefix; the decode kernel detects aligned groups of 4 on the device and reads each
shared page from HBM once.  Results must match the oracle exactly as in the ungrouped kernel,
whatever the alignment, the shared length, the split or the raggedness."""
import pytest
import torch

import pa_testing as pt

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _shared_problem(dtype, groups, ctx, shared_pages, H=6, D=128, B=16, seed=0, nb=None):
    """groups: list of group sizes (1..4+); every group shares its first `shared_pages[g]` pages."""
    S = sum(groups)
    p = pt.make_decode_problem(S, H, D, B, nb or (S * ((max(ctx) + B - 1) // B) + 8), dtype, ctx, dist="unit",
                               seed=seed, unique_blocks=True)
    row = 0
    for g, sp in zip(groups, shared_pages):
        for r in range(1, g):
            p.block_tables[row + r, :sp] = p.block_tables[row, :sp]
        row += g
    return p


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("splits", [0, 1, 2, 5])
def test_parallel_sampling_groups(dtype, splits):
    """8 aligned groups of 4 siblings: 40 shared prompt pages, each sibling with its own last pages."""
    groups = [4] * 8
    ctx = [40 * 16 + 1 + (i * 7) % 50 for i in range(32)]
    p = _shared_problem(dtype, groups, ctx, [40] * 8)
    want = pt.oracle_decode(p)
    pd = p.to(DEV)
    for variant in (2, 3):                     # 2: sibling groups served together, 3: every sequence on its own
        got = pt.run_decode(pd, tuning=pt.tuning(variant, 4, 2, splits))
        pt.assert_decode_close(got, want, dtype, f"variant {variant} splits {splits}")


@pytest.mark.parametrize("groups,shared", [
    ([4, 4, 2], [30, 3, 30]),                 # last group smaller than 4; one group below the sharing threshold
    ([1, 4, 4, 1], [0, 25, 25, 0]),           # groups not aligned to multiples of 4: partial overlap with the quads
    ([2, 2, 2, 2], [20, 20, 20, 20]),         # quads made of two different pairs: first pages differ inside the quad
    ([4], [64]), ([3], [10]), ([8], [33]),    # one group only; a beam of 8 = two quads sharing the same pages
    ([4, 4], [1, 200]),
])
def test_group_shapes_and_alignment(groups, shared):
    S = sum(groups)
    base = max(shared) * 16
    ctx = [base + 5 + 3 * i for i in range(S)]
    p = _shared_problem(torch.float16, groups, ctx, shared, seed=len(groups))
    want = pt.oracle_decode(p)
    for splits in (0, 3):
        got = pt.run_decode(p.to(DEV), tuning=pt.tuning(2, 4, 2, splits))
        pt.assert_decode_close(got, want, torch.float16, f"groups {groups} shared {shared} splits {splits}")


def test_beam_partial_sharing_and_ragged_lengths():
    """Beam search: candidates share a prefix of different length with each other; lengths differ inside a
    group, one sibling ends exactly on the shared prefix, one inside it (its own context is shorter)."""
    B = 16
    ctx = [50 * B, 50 * B + 9, 47 * B + 3, 60 * B + 1,      # quad 0: third sibling is shorter than the shared run
           33 * B + 1, 33 * B + 2, 33 * B + 3, 33 * B + 4]
    p = pt.make_decode_problem(8, 5, 128, B, 700, torch.bfloat16, ctx, dist="unit", seed=3, unique_blocks=True)
    p.block_tables[1, :50] = p.block_tables[0, :50]
    p.block_tables[2, :47] = p.block_tables[0, :47]
    p.block_tables[3, :45] = p.block_tables[0, :45]          # diverges earlier
    for r in (5, 6, 7):
        p.block_tables[r, :33] = p.block_tables[4, :33]
    p.block_tables[6, 20] = p.block_tables[4, 21]             # a hole in the middle of the shared run
    want = pt.oracle_decode(p)
    for splits in (0, 2, 4):
        got = pt.run_decode(p.to(DEV), tuning=pt.tuning(2, 4, 2, splits))
        pt.assert_decode_close(got, want, torch.bfloat16, f"beam splits {splits}")


def test_groups_with_garbage_tails_and_repeat_calls():
    B = 16
    groups, ctx = [4, 4], [20 * B + 3, 20 * B + 7, 20 * B + 1, 21 * B, 9 * B + 5, 9 * B + 6, 9 * B + 7, 9 * B + 8]
    p = _shared_problem(torch.float16, groups, ctx, [20, 9], seed=9)
    want = pt.oracle_decode(p)
    kc, vc = p.key_cache.clone(), p.value_cache.clone()
    for i, c in enumerate(ctx):
        if c % B:
            blk = int(p.block_tables[i, c // B])
            kc[blk, :, :, c % B:, :] = float("nan")
            vc[blk, :, :, c % B:] = float("inf")
    pd = pt.DecodeProblem(p.query._base.to(DEV)[:, 0], kc.to(DEV), vc.to(DEV), p.block_tables.to(DEV),
                          p.context_lens.to(DEV), p.scale, B, p.max_context_len)
    for _ in range(3):
        for splits in (0, 2):
            pt.assert_decode_close(pt.run_decode(pd, tuning=pt.tuning(2, 4, 2, splits)), want, torch.float16,
                                   f"garbage tail, splits {splits}")
