"""Real multi-GPU check of GraphAllReduce (pa_allreduce) under torchrun: correctness against the fp32 sum of the
gathered inputs, both paths (multimem / peer), and timing next to NCCL's all_reduce on the same buffer.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/allreduce_check.py [--json gpurun_out/allreduce_check.json]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def time_fn(fn, iters=200, warmup=20):
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def graph_time_us(fn, per_graph=20, replays=10):
    """Device time per call: `per_graph` calls captured in one CUDA graph (no host launch cost in the
    number; also shows the call is capturable), replayed `replays` times; max over ranks."""
    fn()
    torch.cuda.synchronize()
    dist.barrier()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(per_graph):
            fn()
    torch.cuda.synchronize()
    dist.barrier()
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    dist.barrier()
    e0.record()
    for _ in range(replays):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (per_graph * replays)], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_leg(rank, world, dtypes=(torch.bfloat16,), hidden=5120, max_tokens=2560, token_counts=(8, 256, 2560),
            modes=(True, False), emit=None):
    """Correctness (fp32 sum of the gathered inputs; identical bits on every rank; three launches in a row) and
    graph-timed device time per call of GraphAllReduce for every (dtype, path, token count), next to NCCL's
    all_reduce on the same buffer.  Needs an initialised NCCL process group.  Rows are returned on rank 0."""
    from vllm_b200.all_reduce import GraphAllReduce
    rows = []
    for dtype in dtypes:
        for mm in modes:
            try:
                ar = GraphAllReduce(max_tokens, hidden, dtype, group=None, use_multimem=mm)
            except RuntimeError as e:
                if rank == 0:
                    rows.append({"dtype": str(dtype).split(".")[-1], "multimem": mm, "skipped": str(e)[:200]})
                continue
            for tokens in token_counts:
                g = torch.Generator(device="cuda").manual_seed(1234 + rank)
                x = ar.buffer[:tokens]
                src = torch.empty((tokens, hidden), dtype=dtype, device="cuda").normal_(generator=g)
                gathered = [torch.empty_like(src) for _ in range(world)]
                dist.all_gather(gathered, src)
                want = torch.zeros((tokens, hidden), dtype=torch.float32, device="cuda")
                for t in gathered:
                    want += t.float()
                ok_all = True
                for rep in range(3):                       # repeated launches reuse the flags
                    x.copy_(src)
                    torch.cuda.synchronize()
                    dist.barrier()
                    ar.launch(x)
                    torch.cuda.synchronize()
                    tol = {torch.bfloat16: 1.6e-2, torch.float16: 2e-3, torch.float32: 1e-5}[dtype]
                    ok = bool(((x.float() - want).abs() <= tol * want.abs() + tol).all())
                    # identical bits on every rank
                    mine = x.clone()
                    ref0 = mine.clone()
                    dist.broadcast(ref0, src=0)
                    same = bool(torch.equal(mine.view(torch.uint8), ref0.view(torch.uint8)))
                    flag = torch.tensor([int(ok and same)], device="cuda")
                    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                    ok_all = ok_all and bool(flag.item())
                us_ours = graph_time_us(lambda: ar.launch(x))
                us_nccl = graph_time_us(lambda: dist.all_reduce(x))
                if rank == 0:
                    nbytes = tokens * hidden * x.element_size()
                    row = {"dtype": str(dtype).split(".")[-1], "path": "multimem (NVLS)" if ar.path_for(nbytes) == "multimem" else "peer loads/stores",
                           "multimem": mm, "tokens": tokens, "bytes": nbytes,
                           "world": world, "correct_and_identical_on_all_ranks": ok_all, "ours_us": round(us_ours, 2),
                           "nccl_us": round(us_nccl, 2), "speedup_vs_nccl": round(us_nccl / us_ours, 2),
                           "ours_busbw_gbs": round(2 * (world - 1) / world * nbytes / us_ours / 1e3, 1),
                           "timing": "device time per call, 20 calls captured in one CUDA graph, 10 replays, max over ranks"}
                    if emit:
                        emit(row)
                    rows.append(row)
            del ar
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=os.path.join(ROOT, "gpurun_out", "allreduce_check.json"))
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rows = run_leg(rank, world, dtypes=(torch.bfloat16, torch.float16, torch.float32),
                   emit=lambda r: print(json.dumps(r), flush=True))
    if rank == 0:
        os.makedirs(os.path.dirname(args.json), exist_ok=True)
        json.dump(rows, open(args.json, "w"), indent=1)
        bad = [r for r in rows if r.get("correct_and_identical_on_all_ranks") is False]
        print("ALLREDUCE_CHECK", "FAIL" if bad else "OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
