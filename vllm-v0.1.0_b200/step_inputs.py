"""Per-step model inputs from device-resident block tables: the drop-in for the reference's
`Worker._prepare_inputs` (vllm/worker/worker.py:128-237).

    prep = StepInputs(block_size, max_num_seqs, max_blocks_per_seq, device)
    tokens, positions, input_metadata = prep.prepare_inputs(seq_group_metadata_list)

Same argument (the scheduler's list of SequenceGroupMetadata: `is_prompt`, `seq_data`, `sampling_params`,
`block_tables`), same three results, the five tensors bit-identical to the reference's.  What differs is
the traffic: the reference re-uploads every block table every step; here the tables live on the GPU inside
a `pa_step_state` (csrc/step_tables.cu) and a step describes each generation sequence in O(1) -- id, last
token, length, number of blocks, last block.  A full table travels only for a sequence that was not in the
previous step (first decode step after its prompt, or swapped back in) and for beam-search groups, whose
tables are rewritten by forks (core/block_manager.py:137-143) -- those are diffed against the resident copy
in C++ and only the changed span is sent.  One pinned copy and one kernel launch per step, no host
synchronisation.  There is no CPU path: without libpaged_b200.so or a CUDA device this raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Any, List, Sequence, Tuple

import numpy as np
import torch

from . import _cabi
from .input_metadata import InputMetadata


def _ptr(a: np.ndarray) -> int:
    return a.ctypes.data


class StepInputs:

    def __init__(self, block_size: int, max_num_seqs: int, max_blocks_per_seq: int,
                 device: torch.device | str = "cuda", incremental: bool = True) -> None:
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("StepInputs keeps its block tables on a CUDA device; there is no CPU path")
        self.block_size = int(block_size)
        # incremental=True relies on the reference scheduler's discipline: between two consecutive steps a
        # non-beam sequence's table only grows by one block or has its last block replaced
        # (BlockSpaceManager.append_slot).  False sends every table to the library, which diffs it.
        self.incremental = bool(incremental)
        self._handle = C.c_void_p()
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().pa_step_state_create(int(max_num_seqs), int(max_blocks_per_seq), self.block_size,
                                                         C.byref(self._handle)))
        self._prev_ids: set = set()
        self.last_info = _cabi.StepInfo()

    def close(self) -> None:
        if self._handle:
            _cabi.lib().pa_step_state_destroy(self._handle)
            self._handle = C.c_void_p()

    def __del__(self) -> None:   # pragma: no cover - interpreter shutdown order
        try:
            self.close()
        except Exception:
            pass

    def reset(self) -> None:
        _cabi.check(_cabi.lib().pa_step_state_reset(self._handle))
        self._prev_ids = set()

    # ------------------------------------------------------------------
    def prepare_inputs(self, seq_group_metadata_list: Sequence[Any]) -> Tuple[torch.Tensor, torch.Tensor, InputMetadata]:
        seq_groups: List[Tuple[List[int], Any]] = []
        seq_data = {}
        # prompts first (worker.py:138-167)
        prompt_lens: List[int] = []
        prompt_tokens: List[int] = []
        prompt_tables: List[int] = []
        prompt_table_offsets = [0]
        for g in seq_group_metadata_list:
            if not g.is_prompt:
                continue
            ids = list(g.seq_data.keys())
            seq_groups.append((ids, g.sampling_params))
            toks = g.seq_data[ids[0]].get_token_ids()
            prompt_lens.append(len(toks))
            prompt_tokens.extend(toks)
            if g.block_tables is not None:              # None during memory profiling -> dummy slots
                prompt_tables.extend(g.block_tables[ids[0]])
            prompt_table_offsets.append(len(prompt_tables))
        # generation sequences (worker.py:170-199): O(1) per sequence unless its table must travel
        prev = self._prev_ids
        ids_l: List[int] = []
        tok_l: List[int] = []
        len_l: List[int] = []
        nb_l: List[int] = []
        last_l: List[int] = []
        full: List[int] = []
        offs = [0]
        for g in seq_group_metadata_list:
            if g.is_prompt:
                continue
            gids = list(g.seq_data.keys())
            seq_groups.append((gids, g.sampling_params))
            send_all = not self.incremental or bool(getattr(g.sampling_params, "use_beam_search", False))
            tables = g.block_tables
            for sid in gids:
                d = g.seq_data[sid]
                t = tables[sid]
                ids_l.append(sid)
                tok_l.append(d.get_last_token_id())
                len_l.append(d.get_len())
                nb_l.append(len(t))
                last_l.append(t[-1])
                if send_all or sid not in prev:
                    full.extend(t)
                offs.append(len(full))
        for g in seq_group_metadata_list:
            seq_data.update(g.seq_data)
        out = self._run(ids_l, tok_l, len_l, nb_l, last_l, offs, full, prompt_lens, prompt_tokens,
                        prompt_table_offsets, prompt_tables, seq_group_metadata_list)
        tokens, positions, slots, ctx, tables_t, max_context_len = out
        meta = InputMetadata(seq_groups=seq_groups, seq_data=seq_data, prompt_lens=prompt_lens, slot_mapping=slots,
                             context_lens=ctx, max_context_len=max_context_len, block_tables=tables_t)
        return tokens, positions, meta

    # ------------------------------------------------------------------
    def _run(self, ids_l, tok_l, len_l, nb_l, last_l, offs, full, prompt_lens, prompt_tokens, prompt_table_offsets,
             prompt_tables, groups):
        S, P = len(ids_l), len(prompt_lens)
        npt = len(prompt_tokens)
        total = npt + S
        if total == 0:
            raise RuntimeError("prepare_inputs: empty step")
        width = max(nb_l) if S else 0
        dev = self.device
        padded = total + (-total) % 8
        tokens = torch.empty(padded, dtype=torch.int64, device=dev)
        positions = torch.empty(padded, dtype=torch.int64, device=dev)
        slots = torch.empty(total, dtype=torch.int32, device=dev)
        ctx = torch.empty(S, dtype=torch.int32, device=dev)
        # the reference's IntTensor([]) of an empty generation batch is 1-D with 0 elements
        tables_t = torch.empty((S, width), dtype=torch.int32, device=dev) if S else torch.empty(0, dtype=torch.int32, device=dev)

        i64, i32 = np.int64, np.int32
        a_ids, a_tok = np.asarray(ids_l, dtype=i64), np.asarray(tok_l, dtype=i64)
        a_len, a_nb, a_last = np.asarray(len_l, dtype=i32), np.asarray(nb_l, dtype=i32), np.asarray(last_l, dtype=i32)
        a_offs, a_full = np.asarray(offs, dtype=i64), np.asarray(full, dtype=i32)
        p_off = np.zeros(P + 1, dtype=i64)
        if P:
            np.cumsum(np.asarray(prompt_lens, dtype=i64), out=p_off[1:])
        p_tok = np.asarray(prompt_tokens, dtype=i64)
        p_toff, p_tab = np.asarray(prompt_table_offsets, dtype=i64), np.asarray(prompt_tables, dtype=i32)

        b = _cabi.StepBatch(S, _ptr(a_ids), _ptr(a_tok), _ptr(a_len), _ptr(a_nb), _ptr(a_last), _ptr(a_offs), _ptr(a_full),
                            P, _ptr(p_off), _ptr(p_tok), _ptr(p_toff), _ptr(p_tab))
        o = _cabi.StepOutputs(tokens.data_ptr(), positions.data_ptr(), slots.data_ptr(), ctx.data_ptr(),
                              tables_t.data_ptr(), width)
        need = np.zeros(max(S, 1), dtype=np.uint8)
        stream = torch.cuda.current_stream(dev).cuda_stream
        lib = _cabi.lib()
        with torch.cuda.device(dev):
            rc = lib.pa_step_prepare_inputs(self._handle, C.byref(b), C.byref(o), C.byref(self.last_info), _ptr(need), stream)
            if rc == _cabi.PA_ERR_NEED_TABLE:
                # a sequence changed by more than its last block without leaving the batch (not something the
                # reference's scheduler does, but harmless): send the flagged tables and go again
                flat = [t for g in groups if not g.is_prompt for sid in g.seq_data for t in (g.block_tables[sid],)]
                full2: List[int] = []
                offs2 = [0]
                for i in range(S):
                    lo, hi = offs[i], offs[i + 1]
                    if hi > lo:
                        full2.extend(full[lo:hi])
                    elif need[i]:
                        full2.extend(flat[i])
                    offs2.append(len(full2))
                a_offs, a_full = np.asarray(offs2, dtype=i64), np.asarray(full2, dtype=i32)
                b.table_offsets, b.table_blocks = _ptr(a_offs), _ptr(a_full)
                rc = lib.pa_step_prepare_inputs(self._handle, C.byref(b), C.byref(o), C.byref(self.last_info), _ptr(need), stream)
            _cabi.check(rc)
        self._prev_ids = set(ids_l)
        return tokens, positions, slots, ctx, tables_t, max(len_l) if S else 0
