"""The N>1 path on CPU (gloo, world_size 2): head sharding needs no exchange -- each rank
computes its own heads from its own cache shard and replicated block tables, and the
concatenation over ranks equals the unsharded result; bench.py's reference arm prints one
line from rank 0 only.  The compute here is the oracle (there is no GPU in this container):
what is under test is the sharding arithmetic and the launch contract."""
import json
import os
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import pa_testing as pt
from vllm_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_head_ranges():
    assert sharding.heads_per_gpu(40, 8) == 5
    assert [sharding.head_range(r, 4, 40) for r in range(4)] == [(0, 10), (10, 20), (20, 30), (30, 40)]
    for total in (40, 64, 72):                       # LLaMA-13B, GPT-NeoX-20B, OPT-66B
        for g in (1, 2, 4, 8):
            rs = [sharding.head_range(r, g, total) for r in range(g)]
            assert rs[0][0] == 0 and rs[-1][1] == total and all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
    with pytest.raises(ValueError, match="divisible"):
        sharding.heads_per_gpu(40, 3)
    with pytest.raises(ValueError):
        sharding.head_range(2, 2, 40)


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        H = 8
        # every rank builds the same problem from the same seed (the scheduler output is replicated)
        p = pt.make_decode_problem(6, H, 64, 16, 64, torch.float32, [1, 16, 40, 300, 77, 512], seed=21)
        lo, hi = sharding.head_range(rank, world, H)
        shard = pt.DecodeProblem(sharding.shard_heads(p.query, rank, world).contiguous(),
                                 sharding.shard_heads(p.key_cache, rank, world).contiguous(),
                                 sharding.shard_heads(p.value_cache, rank, world).contiguous(),
                                 p.block_tables, p.context_lens, p.scale, p.block_size, p.max_context_len)
        mine = pt.oracle_decode(shard)                      # [S, H/world, D], no communication
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)                        # test-only: bring the shards together to compare
        tables = [torch.empty_like(p.block_tables) for _ in range(world)]
        dist.all_gather(tables, p.block_tables)
        # max-over-ranks of a per-rank time, as bench.py reports it
        t = torch.tensor([1.0 + rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            full = pt.oracle_decode(p)
            ok = torch.equal(torch.cat(parts, dim=1), full) and all(torch.equal(tb, tables[0]) for tb in tables) \
                and float(t) == float(world)
            open(os.path.join(tmp, "result"), "w").write("ok" if ok else "mismatch")
    finally:
        dist.destroy_process_group()


def test_sharded_heads_concatenate_to_the_full_result(tmp_path):
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert open(tmp_path / "result").read() == "ok"


def test_reference_arm_prints_one_line_from_rank0_under_torchrun():
    port = 31500 + os.getpid() % 2000
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"),
           "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "3", "--workload", "llama7b"]
    env = dict(os.environ, OMP_NUM_THREADS="4")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["unit"] == "seqs/s"
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0
    assert line["value"] > 0 and line["config"]["workload"] == "llama7b"
