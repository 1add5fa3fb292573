"""Generates tests/golden/cfg4_scenario.json: the cache-op trace of BASELINE configs[3]
(OPT-13B-style parallel sampling n=4 and beam width 4) produced by the REFERENCE's own
BlockSpaceManager / Sequence classes.

Run in the build container, where /root/reference exists:

    python tests/golden/make_cfg4_scenario.py

vllm/core/block_manager.py, vllm/block.py, vllm/sequence.py, vllm/sampling_params.py and
vllm/utils.py are imported unmodified from /root/reference (the package __init__, which pulls in
ray / the compiled extensions, is bypassed with an empty stand-in package object).  This script
plays the role of the scheduler + worker around them, restating only their glue:

  prompt step   Scheduler._allocate (core/scheduler.py:318-321) -> BlockSpaceManager.allocate;
                slot mapping of the prompt tokens as Worker._prepare_inputs (worker/worker.py:150-167)
  decode step   Scheduler.update (core/scheduler.py:284-309): beam forks = free(child), parent.fork(child),
                BlockSpaceManager.fork; then append_token_id;
                Scheduler._append_slot (:326-338) -> blocks_to_copy; decode slot / block tables /
                context lens as worker/worker.py:170-199
  preemption    BlockSpaceManager.swap_out / swap_in (core/block_manager.py:167-223) of one group

Only the TRACE (block numbers, slots, tables) is stored -- no reference code.  Tests replay it
through cache_ops / attention_ops on the GPU and through the oracle on the CPU.
"""
from __future__ import annotations

import json
import os
import random
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("PA_REFERENCE_ROOT", "/root/reference")


def _reference_modules():
    pkg = types.ModuleType("vllm")
    pkg.__path__ = [os.path.join(REF, "vllm")]          # submodules resolve here; vllm/__init__.py is not run
    sys.modules["vllm"] = pkg
    import vllm.core.block_manager as bm                  # noqa: E402  (the reference's file)
    import vllm.sampling_params as sp                     # noqa: E402
    import vllm.sequence as sq                            # noqa: E402
    assert bm.__file__.startswith(REF) and sq.__file__.startswith(REF)
    return bm, sq, sp


def make(seed: int = 0, num_groups: int = 6, width: int = 4, block_size: int = 16, num_gpu_blocks: int = 160,
         num_cpu_blocks: int = 64, steps: int = 40) -> dict:
    bm, sq, sp = _reference_modules()
    rnd = random.Random(seed)
    manager = bm.BlockSpaceManager(block_size, num_gpu_blocks, num_cpu_blocks, watermark=0.0)
    ops = []
    groups = []
    next_id = 0
    # half of the groups sample in parallel (n = width), half run beam search (best_of = width)
    for g in range(num_groups):
        prompt_len = rnd.choice([5, 16, 17, 31, 32, 47, 64, 90])
        beam = g % 2 == 1
        params = sp.SamplingParams(n=width, best_of=width, use_beam_search=beam, temperature=0.0 if beam else 1.0)
        seqs = [sq.Sequence(next_id + i, "", list(range(prompt_len)), block_size) for i in range(width)]
        next_id += width
        group = sq.SequenceGroup(f"req{g}", seqs, params, arrival_time=float(g))
        groups.append((group, beam))

    # ---- prompt step ---------------------------------------------------------------
    slots, owners = [], []
    for group, _ in groups:
        manager.allocate(group)
        for s in group.get_seqs():
            s.status = sq.SequenceStatus.RUNNING
        first = group.get_seqs()[0]
        table = manager.get_block_table(first)
        for i in range(first.get_len()):
            slots.append(table[i // block_size] * block_size + i % block_size)
            owners.append(first.seq_id)
    ops.append({"op": "reshape_and_cache", "phase": "prompt", "slots": slots, "seq_ids": owners,
                "sharers": {str(grp.get_seqs()[0].seq_id): [s.seq_id for s in grp.get_seqs()] for grp, _ in groups}})

    swapped = None           # (group index, step it left)
    for step in range(steps):
        running = [(gi, grp, beam) for gi, (grp, beam) in enumerate(groups)
                   if all(s.status == sq.SequenceStatus.RUNNING for s in grp.get_seqs())]
        # ---- Scheduler.update of the previous model step: beam forks, then the new token ----
        for gi, group, beam in running:
            seqs = group.get_seqs(status=sq.SequenceStatus.RUNNING)
            if beam and step > 0:
                ids = [s.seq_id for s in seqs]
                parents = {sid: rnd.choice(ids) for sid in ids}
                keep = rnd.choice(ids)
                parents[keep] = keep                      # a beam always survives as its own parent
                # the reference forks only from sequences that are not themselves replaced this step
                survivors = [sid for sid in ids if parents[sid] == sid]
                for sid in ids:
                    if parents[sid] != sid and parents[sid] not in survivors:
                        parents[sid] = rnd.choice(survivors)
                for s in seqs:
                    if parents[s.seq_id] != s.seq_id:
                        manager.free(s)
                        parent = group.find(parents[s.seq_id])
                        parent.fork(s)
                        manager.fork(parent, s)
                        ops.append({"op": "fork", "parent": parent.seq_id, "child": s.seq_id})
            for s in seqs:
                s.append_token_id(7, {7: 0.0})
        # ---- Scheduler._schedule: preempt / resume one group, then append_slot for the rest ----
        if step == steps // 3:
            gi, group, _ = running[-1]
            mapping = manager.swap_out(group)
            for s in group.get_seqs():
                s.status = sq.SequenceStatus.SWAPPED
            ops.append({"op": "swap_out", "map": {str(k): v for k, v in mapping.items()}})
            swapped = gi
            running = running[:-1]
        elif swapped is not None and step == 2 * steps // 3:
            group = groups[swapped][0]
            mapping = manager.swap_in(group)
            for s in group.get_seqs():
                s.status = sq.SequenceStatus.RUNNING
            ops.append({"op": "swap_in", "map": {str(k): v for k, v in mapping.items()}})
            running.append((swapped, group, groups[swapped][1]))
            swapped = None
        blocks_to_copy = {}
        for gi, group, beam in running:
            for s in group.get_seqs(status=sq.SequenceStatus.RUNNING):
                ret = manager.append_slot(s)
                if ret is not None:
                    blocks_to_copy.setdefault(ret[0], []).append(ret[1])
        if blocks_to_copy:
            ops.append({"op": "copy_blocks", "map": {str(k): v for k, v in blocks_to_copy.items()}})
        # ---- Worker._prepare_inputs for the decode tokens ----
        slots, tables, lens, seq_ids = [], [], [], []
        for gi, group, beam in running:
            for s in group.get_seqs(status=sq.SequenceStatus.RUNNING):
                table = manager.get_block_table(s)
                ctx = s.get_len()
                pos = ctx - 1
                slots.append(table[pos // block_size] * block_size + pos % block_size)
                tables.append(table)
                lens.append(ctx)
                seq_ids.append(s.seq_id)
        ops.append({"op": "reshape_and_cache", "phase": "decode", "slots": slots, "seq_ids": seq_ids})
        ops.append({"op": "attention", "block_tables": tables, "context_lens": lens, "seq_ids": seq_ids})
    return {"block_size": block_size, "num_gpu_blocks": num_gpu_blocks, "num_cpu_blocks": num_cpu_blocks,
            "width": width, "num_groups": num_groups, "steps": steps, "seed": seed, "ops": ops,
            "generator": "tests/golden/make_cfg4_scenario.py (reference BlockSpaceManager, core/block_manager.py)"}


if __name__ == "__main__":
    trace = make()
    counts = {}
    for op in trace["ops"]:
        counts[op["op"]] = counts.get(op["op"], 0) + 1
    pairs = sum(len(v) for op in trace["ops"] if op["op"] == "copy_blocks" for v in op["map"].values())
    print(counts, "copy pairs:", pairs)
    with open(os.path.join(HERE, "cfg4_scenario.json"), "w") as f:
        json.dump(trace, f, separators=(",", ":"))
    print(os.path.getsize(os.path.join(HERE, "cfg4_scenario.json")), "bytes")
