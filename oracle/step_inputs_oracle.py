"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's per-step input preparation.

Follows `Worker._prepare_inputs` (vllm/worker/worker.py:128-237) with its helpers `_pad_to_alignment` /
`_pad_to_max` (:304-309): the flat token batch is |prompt tokens ... |generation tokens ... |padding to 8|,
slot = block_table[pos // B] * B + pos % B, generation block tables zero-padded to the widest row.

Pinned: tests/golden/step_inputs_trace.json.gz holds, per step of a trace driven by the reference's own
Scheduler, the SHA-256 of what the reference's `_prepare_inputs` (executed unmodified, see
tests/golden/make_step_inputs_golden.py) returned; tests/test_step_inputs.py requires this restatement
to reproduce every hash.

Only tests/, __graft_entry__.smoke() and bench/tools CPU legs may import this module.
"""
from __future__ import annotations

import hashlib
from typing import Any, Dict, List, Sequence

import numpy as np


class SeqData:
    """Stand-in for vllm.sequence.SequenceData (sequence.py:36-66): only what _prepare_inputs reads."""

    def __init__(self, token_ids: Sequence[int] = (), length: int = None, last: int = None) -> None:
        self._tokens = list(token_ids) if token_ids is not None else None
        self._len = len(self._tokens) if length is None else length
        self._last = (self._tokens[-1] if self._tokens else None) if last is None else last

    def get_len(self) -> int:
        return self._len

    def get_token_ids(self) -> List[int]:
        return self._tokens

    def get_last_token_id(self) -> int:
        return self._last


class SamplingStub:
    def __init__(self, use_beam_search: bool = False) -> None:
        self.use_beam_search = use_beam_search


class GroupMetadata:
    """Stand-in for vllm.sequence.SequenceGroupMetadata (sequence.py:201-215)."""

    def __init__(self, request_id, is_prompt, seq_data, sampling_params, block_tables) -> None:
        self.request_id = request_id
        self.is_prompt = is_prompt
        self.seq_data = seq_data
        self.sampling_params = sampling_params
        self.block_tables = block_tables


def groups_from_trace(step: Dict[str, Any]) -> List[GroupMetadata]:
    """One step of step_inputs_trace.json.gz -> the scheduler's metadata list."""
    out = []
    for g in step["groups"]:
        data = {s["id"]: SeqData(s["tokens"], s["len"], s["last"]) for s in g["seqs"]}
        tables = None if g["block_tables"] is None else {int(k): v for k, v in g["block_tables"].items()}
        out.append(GroupMetadata(g["request_id"], g["is_prompt"], data, SamplingStub(g["beam"]), tables))
    return out


def prepare_inputs(groups: Sequence[Any], block_size: int) -> Dict[str, Any]:
    """worker.py:128-237.  Returns numpy arrays with the reference's dtypes (tokens / positions int64,
    the rest int32) and its host-side values."""
    tokens: List[int] = []
    positions: List[int] = []
    slots: List[int] = []
    prompt_lens: List[int] = []
    for g in groups:                                   # :138-167
        if not g.is_prompt:
            continue
        seq_id = next(iter(g.seq_data))                # "use any sequence in the group"
        ids = g.seq_data[seq_id].get_token_ids()
        n = len(ids)
        prompt_lens.append(n)
        tokens.extend(ids)
        positions.extend(range(n))
        if g.block_tables is None:                     # memory profiling: dummy slots (:160-164)
            slots.extend([0] * n)
            continue
        table = g.block_tables[seq_id]
        slots.extend(table[i // block_size] * block_size + i % block_size for i in range(n))
    max_context_len = 0
    width = 0
    context_lens: List[int] = []
    tables: List[List[int]] = []
    for g in groups:                                   # :170-199
        if g.is_prompt:
            continue
        for seq_id, data in g.seq_data.items():
            tokens.append(data.get_last_token_id())
            ctx = data.get_len()
            pos = ctx - 1
            positions.append(pos)
            table = g.block_tables[seq_id]
            tables.append(table)
            max_context_len = max(max_context_len, ctx)
            width = max(width, len(table))
            context_lens.append(ctx)
            slots.append(table[pos // block_size] * block_size + pos % block_size)
    pad = (-len(tokens)) % 8                           # :201-204
    tokens += [0] * pad
    positions += [0] * pad
    bt = np.zeros((len(tables), width), dtype=np.int32) if tables else np.zeros((0,), dtype=np.int32)
    for i, t in enumerate(tables):                     # :211-214 (_pad_to_max pads with 0)
        bt[i, :len(t)] = t
    return {"tokens": np.asarray(tokens, dtype=np.int64), "positions": np.asarray(positions, dtype=np.int64),
            "slot_mapping": np.asarray(slots, dtype=np.int32), "context_lens": np.asarray(context_lens, dtype=np.int32),
            "block_tables": bt, "max_context_len": max_context_len, "prompt_lens": prompt_lens}


def digest(r: Dict[str, Any]) -> Dict[str, Any]:
    """Same summary as tests/golden/make_step_inputs_golden.py::digest."""
    h = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:32]
    return {"tokens": h(r["tokens"]), "positions": h(r["positions"]), "slot_mapping": h(r["slot_mapping"]),
            "context_lens": h(r["context_lens"]), "block_tables": h(r["block_tables"]),
            "block_tables_shape": list(r["block_tables"].shape), "max_context_len": r["max_context_len"],
            "prompt_lens": list(r["prompt_lens"]), "num_tokens": int(r["tokens"].shape[0])}
