"""Generates tests/golden/step_inputs_trace.json.gz: a multi-step serving trace and, for every step,
the SHA-256 of what the REFERENCE's own `Worker._prepare_inputs` returns for it.

Run in the build container, where /root/reference exists:

    python tests/golden/make_step_inputs_golden.py

What runs unmodified from /root/reference (the package __init__, which pulls in ray and the compiled
extensions, is bypassed with an empty stand-in package object, as in make_cfg4_scenario.py):

  vllm/core/scheduler.py (Scheduler.schedule / update / free_seq / free_finished_seq_groups -- the real
  scheduling, preemption by swap and by recompute, beam forks), vllm/core/block_manager.py,
  vllm/core/policy.py, vllm/sequence.py, vllm/sampling_params.py, vllm/config.py, and -- cut out of
  vllm/worker/worker.py with `ast` because the module itself imports the compiled ops --
  `Worker._prepare_inputs` (:128-237), `_pad_to_alignment`, `_pad_to_max` (:304-309), executed with
  `torch.cuda.LongTensor / IntTensor` pointed at their CPU twins.

This script plays the engine around them (llm_engine.py:196-235): add requests over time, invent the
sampler's outputs (random tokens; beam parents drawn at random), stop sequences at random lengths.

Only the TRACE is stored -- per step the scheduler's SequenceGroupMetadata contents (ids, token ids,
block tables) and the hashes of the five tensors + three host values the reference produced -- no
reference code.  tests/test_step_inputs.py replays it through oracle/step_inputs_oracle.py (CPU; must
hash-match, which pins the oracle) and through vllm_b200.step_inputs on the GPU (must equal the oracle).
"""
from __future__ import annotations

import ast
import gzip
import hashlib
import json
import os
import random
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("PA_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "step_inputs_trace.json.gz")


def _reference():
    pkg = types.ModuleType("vllm")
    pkg.__path__ = [os.path.join(REF, "vllm")]
    sys.modules["vllm"] = pkg
    import vllm.config as cfg
    import vllm.core.scheduler as sch
    import vllm.sampling_params as sp
    import vllm.sequence as sq
    assert sch.__file__.startswith(REF) and sq.__file__.startswith(REF)

    # Worker._prepare_inputs and its two helpers, cut out of the reference file and executed as they are
    path = os.path.join(REF, "vllm", "worker", "worker.py")
    tree = ast.parse(open(path).read())
    wanted = []
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in ("_pad_to_alignment", "_pad_to_max"):
            wanted.append(node)
        if isinstance(node, ast.ClassDef) and node.name == "Worker":
            wanted += [n for n in node.body if isinstance(n, ast.FunctionDef) and n.name == "_prepare_inputs"]
    assert len(wanted) == 3

    class InputMetadata:                      # records the constructor arguments (input_metadata.py:12-29)
        def __init__(self, **kw):
            self.__dict__.update(kw)

    cuda = types.SimpleNamespace(LongTensor=torch.LongTensor, IntTensor=torch.IntTensor)
    fake_torch = types.SimpleNamespace(cuda=cuda, Tensor=torch.Tensor)
    from typing import Dict, List, Tuple
    ns = {"torch": fake_torch, "List": List, "Dict": Dict, "Tuple": Tuple, "InputMetadata": InputMetadata,
          "SequenceGroupMetadata": sq.SequenceGroupMetadata, "SamplingParams": sp.SamplingParams,
          "SequenceData": sq.SequenceData}
    exec(compile(ast.Module(body=wanted, type_ignores=[]), path, "exec"), ns)
    return cfg, sch, sp, sq, ns["_prepare_inputs"]


def digest(tokens, positions, meta) -> dict:
    def h(t, dtype):
        assert t.dtype == dtype, (t.dtype, dtype)
        return hashlib.sha256(t.contiguous().numpy().tobytes()).hexdigest()[:32]
    return {"tokens": h(tokens, torch.int64), "positions": h(positions, torch.int64),
            "slot_mapping": h(meta.slot_mapping, torch.int32), "context_lens": h(meta.context_lens, torch.int32),
            "block_tables": h(meta.block_tables, torch.int32), "block_tables_shape": list(meta.block_tables.shape),
            "max_context_len": meta.max_context_len, "prompt_lens": list(meta.prompt_lens),
            "num_tokens": int(tokens.shape[0])}


def make(seed=0, block_size=16, num_gpu_blocks=72, num_cpu_blocks=128, max_num_seqs=24, max_batched=512,
         num_requests=48, max_steps=1500) -> dict:
    cfg, sch, sp, sq, prepare_inputs = _reference()
    rnd = random.Random(seed)
    cache_config = cfg.CacheConfig(block_size, 0.9, 1)
    cache_config.num_gpu_blocks, cache_config.num_cpu_blocks = num_gpu_blocks, num_cpu_blocks
    scheduler = sch.Scheduler(cfg.SchedulerConfig(max_batched, max_num_seqs), cache_config, log_stats=False)
    worker = types.SimpleNamespace(block_size=block_size)

    pending = []
    next_seq = 0
    for r in range(num_requests):
        kind = rnd.choice(["single", "single", "sample4", "beam4", "sample2"])
        width = {"single": 1, "sample2": 2, "sample4": 4, "beam4": 4}[kind]
        beam = kind == "beam4"
        prompt_len = rnd.choice([1, 5, 15, 16, 17, 31, 32, 33, 48, 70, 100])
        max_tokens = rnd.choice([1, 3, 17, 30, 45, 60, 90])
        params = sp.SamplingParams(n=width, best_of=width, use_beam_search=beam, temperature=0.0 if beam else 1.0,
                                   max_tokens=max_tokens)
        prompt = [rnd.randrange(1, 32000) for _ in range(prompt_len)]
        seqs = [sq.Sequence(next_seq + i, "", list(prompt), block_size) for i in range(width)]
        next_seq += width
        group = sq.SequenceGroup(f"req{r}", seqs, params, arrival_time=float(r))
        # per-sequence stop lengths: sampling siblings stop at different times, beams together
        stop_at = {s.seq_id: (max_tokens if beam else rnd.randint(1, max_tokens)) for s in seqs}
        pending.append((rnd.randint(0, 40), group, stop_at))
    pending.sort(key=lambda t: t[0])
    stops = {}

    steps = []
    step = 0
    while (pending or scheduler.has_unfinished_seqs()) and step < max_steps:
        while pending and pending[0][0] <= step:
            _, group, stop_at = pending.pop(0)
            stops.update(stop_at)
            scheduler.add_seq_group(group)
        metadata_list, outputs = scheduler.schedule()
        step += 1
        if not metadata_list:
            continue
        tokens, positions, meta = prepare_inputs(worker, metadata_list)
        groups = []
        for m in metadata_list:
            seqs = []
            for sid, data in m.seq_data.items():
                seqs.append({"id": sid, "len": data.get_len(), "last": data.get_last_token_id(),
                             "tokens": list(data.get_token_ids()) if m.is_prompt else None})
            groups.append({"request_id": m.request_id, "is_prompt": m.is_prompt,
                           "beam": bool(m.sampling_params.use_beam_search), "seqs": seqs,
                           "block_tables": {str(k): list(v) for k, v in m.block_tables.items()}})
        steps.append({"groups": groups, "want": digest(tokens, positions, meta),
                      "swap_in": len(outputs.blocks_to_swap_in), "swap_out": len(outputs.blocks_to_swap_out),
                      "copies": sum(len(v) for v in outputs.blocks_to_copy.values())})
        # ---- the sampler (invented) and the engine's stop logic (llm_engine.py:237-262) ----
        seq_outputs = {}
        for group in scheduler.running:
            seqs = group.get_seqs(status=sq.SequenceStatus.RUNNING)
            ids = [s.seq_id for s in seqs]
            parents = {sid: sid for sid in ids}
            if group.sampling_params.use_beam_search and len(ids) > 1 and rnd.random() < 0.7:
                parents = {sid: rnd.choice(ids) for sid in ids}
                keep = rnd.choice(ids)
                parents[keep] = keep
                survivors = [sid for sid in ids if parents[sid] == sid]
                for sid in ids:                      # forks only from beams that survive as themselves
                    if parents[sid] != sid and parents[sid] not in survivors:
                        parents[sid] = rnd.choice(survivors)
            for sid in ids:
                tok = rnd.randrange(1, 32000)
                seq_outputs[sid] = sq.SequenceOutputs(sid, parents[sid], tok, {tok: 0.0})
        for group in scheduler.update(seq_outputs):
            for s in group.get_seqs(status=sq.SequenceStatus.RUNNING):
                if s.get_output_len() >= stops[s.seq_id]:
                    scheduler.free_seq(s, sq.SequenceStatus.FINISHED_LENGTH_CAPPED)
        scheduler.free_finished_seq_groups()
    assert not pending and not scheduler.has_unfinished_seqs(), "trace did not drain; raise max_steps"
    return {"block_size": block_size, "num_gpu_blocks": num_gpu_blocks, "max_num_seqs": max_num_seqs, "seed": seed,
            "steps": steps,
            "generator": "tests/golden/make_step_inputs_golden.py (reference Scheduler + Worker._prepare_inputs)"}


if __name__ == "__main__":
    trace = make()
    s = trace["steps"]
    print(len(s), "steps;", sum(1 for x in s if any(g["is_prompt"] for g in x["groups"])), "with prompts;",
          sum(1 for x in s if x["swap_out"]), "with swap-out;", sum(1 for x in s if x["swap_in"]), "with swap-in;",
          sum(x["copies"] for x in s), "copy pairs; max decode batch",
          max(x["want"]["block_tables_shape"][0] for x in s), "max table width",
          max(x["want"]["block_tables_shape"][1] if len(x["want"]["block_tables_shape"]) > 1 else 0 for x in s))
    with gzip.GzipFile(OUT, "wb", mtime=0) as f:
        f.write(json.dumps(trace, separators=(",", ":")).encode())
    print(os.path.getsize(OUT), "bytes ->", OUT)
