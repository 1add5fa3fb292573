"""Times one decode step's input preparation at BASELINE shapes: the reference's way (restated
`Worker._prepare_inputs`, worker.py:128-237: nested lists -> five torch.cuda tensors) against
`StepInputs.prepare_inputs` (resident tables, one pinned copy + one kernel).

    python tools/step_inputs_bench.py [--json out.json]

Both produce the same tensors (checked).  Reported per step: host wall time until the inputs are ready on the
device (synchronised), host->device bytes, and for ours the device time of copy + kernel (CUDA events).
"""
from __future__ import annotations

import argparse
import json
import os
import random
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import step_inputs_oracle as so           # noqa: E402  (the CPU leg: the reference's algorithm)
from vllm_b200 import StepInputs                       # noqa: E402


def reference_way(groups, block_size, dev):
    r = so.prepare_inputs(groups, block_size)
    # the reference builds the tensors straight from Python lists (torch.cuda.IntTensor(list));
    # from numpy is the cheaper variant, so this leg is a lower bound for it
    return {k: torch.from_numpy(r[k]).to(dev) for k in ("tokens", "positions", "slot_mapping", "context_lens", "block_tables")}


def reference_literal(groups, block_size, dev):
    """The reference's own construction (worker.py:206-214): tensors straight from (nested) Python lists."""
    r = so.prepare_inputs(groups, block_size)
    lists = {k: r[k].tolist() for k in r if hasattr(r[k], "tolist")}
    t0 = time.perf_counter()
    out = [torch.tensor(lists["tokens"], dtype=torch.int64, device=dev), torch.tensor(lists["positions"], dtype=torch.int64, device=dev),
           torch.tensor(lists["slot_mapping"], dtype=torch.int32, device=dev),
           torch.tensor(lists["context_lens"], dtype=torch.int32, device=dev),
           torch.tensor(lists["block_tables"], dtype=torch.int32, device=dev)]
    torch.cuda.synchronize()
    return time.perf_counter() - t0, out


def run(name, num_seqs, ctx, block_size, steps, dev):
    rnd = random.Random(0)
    per = ctx // block_size + steps // block_size + 2
    free = list(range(num_seqs * per))
    rnd.shuffle(free)
    lens = {i: ctx for i in range(num_seqs)}
    tables = {i: [free.pop() for _ in range((ctx + block_size - 1) // block_size)] for i in range(num_seqs)}
    prep = StepInputs(block_size, num_seqs, per, dev)
    t_ref, t_ours, dev_ms, ref_bytes, our_bytes, t_lit = [], [], [], [], [], []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for step in range(steps):
        data = {i: so.SeqData(None, lens[i], rnd.randrange(1, 32000)) for i in tables}
        groups = [so.GroupMetadata(f"g{i}", False, {i: data[i]}, so.SamplingStub(False), {i: tables[i]}) for i in tables]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        want = reference_way(groups, block_size, dev)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        if step >= 3:
            t_lit.append(reference_literal(groups, block_size, dev)[0])
        t1b = time.perf_counter()
        e0.record()
        tokens, positions, meta = prep.prepare_inputs(groups)
        e1.record()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        got = {"tokens": tokens, "positions": positions, "slot_mapping": meta.slot_mapping,
               "context_lens": meta.context_lens, "block_tables": meta.block_tables}
        for k in want:
            assert torch.equal(want[k], got[k]), (step, k)
        if step >= 3:
            t_ref.append(t1 - t0)
            t_ours.append(t2 - t1b)
            dev_ms.append(e0.elapsed_time(e1))
            ref_bytes.append(sum(v.numel() * v.element_size() for v in want.values()))
            our_bytes.append(prep.last_info.h2d_bytes)
        for i in tables:
            lens[i] += 1
            if (lens[i] + block_size - 1) // block_size > len(tables[i]):
                tables[i] = tables[i] + [free.pop()]
    prep.close()
    med = lambda v: float(np.median(v))
    return {"shape": name, "num_seqs": num_seqs, "context_len": ctx, "block_size": block_size, "steps_timed": len(t_ref),
            "reference_way_host_us": round(med(t_ref) * 1e6, 1), "ours_host_us": round(med(t_ours) * 1e6, 1),
            "reference_tensor_construction_from_lists_us": round(med(t_lit) * 1e6, 1),
            "speedup": round(med(t_ref) / med(t_ours), 2), "ours_enqueue_to_done_device_us": round(med(dev_ms) * 1e3, 1),
            "reference_h2d_bytes_per_step": int(med(ref_bytes)), "ours_h2d_bytes_per_step": int(med(our_bytes)),
            "ours_h2d_bytes_max": int(max(our_bytes)), "identical_tensors": True}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--json")
    ap.add_argument("--steps", type=int, default=60)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    rows = [run("llama13b cfg3: 256 seqs x 2048", 256, 2048, 16, a.steps, dev),
            run("llama7b cfg2: 256 seqs x 1024", 256, 1024, 16, a.steps, dev),
            run("cfg5: 64 seqs x 8192", 64, 8192, 16, a.steps, dev)]
    for r in rows:
        print(json.dumps(r))
    if a.json:
        with open(a.json, "w") as f:
            json.dump({"host_cores": os.cpu_count(), "rows": rows}, f, indent=1)
