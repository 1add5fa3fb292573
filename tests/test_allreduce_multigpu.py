"""pa_allreduce / GraphAllReduce on REAL ranks: two processes, two GPUs, NCCL rendezvous + symmetric memory.

Skipped unless the box has at least two GPUs (the single-GPU tests in test_allreduce.py cover the peer path
in an emulated world).  The ranks run tools/allreduce_check.py: for bf16 / fp16 / fp32 and both paths
(multimem.ld_reduce through the NVSwitch when the buffer has a multicast mapping, peer loads / stores otherwise)
the result must equal the fp32 sum of the gathered inputs within the dtype's rounding and be bit-identical on
every rank, three launches in a row; the launch must be capturable in a CUDA graph (the reference's contract,
vllm/model_executor/parallel_utils/parallel_state.py:531-556)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_allreduce_both_paths(tmp_path):
    out = tmp_path / "allreduce.json"
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "allreduce_check.py"), "--json", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "ALLREDUCE_CHECK OK" in r.stdout, r.stdout[-2000:]
    rows = json.load(open(out))
    ran = [x for x in rows if "ours_us" in x]
    assert ran and all(x["correct_and_identical_on_all_ranks"] for x in ran)
    assert any(not x["multimem"] for x in ran), "the peer path did not run"
    # on an NVSwitch box the multicast path must have run too; elsewhere it reports why it was skipped
    mm = [x for x in rows if x.get("multimem")]
    assert mm and all(("ours_us" in x) or ("skipped" in x) for x in mm)
