"""BASELINE configs[3]: parallel sampling n=4 / beam width 4 -- reshape_and_cache + copy_blocks
(copy-on-write) + swap_blocks driven by a trace that the reference's own BlockSpaceManager
produced (tests/golden/cfg4_scenario.json, generator: tests/golden/make_cfg4_scenario.py).

CPU (not gpu): both oracles replay the trace to identical bits, and after every step each
sequence sees, through its block table, exactly the tokens of its own lineage (prompt shared by
the group, forks copy the parent's history) -- the trace and the oracle are semantically right.
GPU: cache_ops replays the trace on OPT-13B-shaped caches (40 heads x 128, block 16, fp16) and
must match the oracle replay bit for bit after every op, GPU and CPU caches both; decode
attention over the copy-on-written tables must match the oracle."""
import json
import os

import pytest
import torch

import pa_testing as pt
from oracle import c_oracle, paged_oracle

TRACE = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "cfg4_scenario.json")))


class Replay:
    """Applies the trace to `layers` (key, value) cache pairs through an ops module with the
    reference's cache_ops signatures."""

    def __init__(self, ops, layers, heads, head_size, dtype, device, pin=False):
        self.ops, self.layers, self.B = ops, layers, TRACE["block_size"]
        x = pt.pack_factor(dtype)
        g = torch.Generator().manual_seed(5)
        nb, ncpu = TRACE["num_gpu_blocks"], TRACE["num_cpu_blocks"]
        # caches start as garbage (never-written memory), identical on every replay
        mk = lambda n, shape: [torch.randn((n, *shape), generator=g).to(dtype) for _ in range(layers)]
        kshape, vshape = (heads, head_size // x, self.B, x), (heads, head_size, self.B)
        self.gpu_k = [t.to(device) for t in mk(nb, kshape)]
        self.gpu_v = [t.to(device) for t in mk(nb, vshape)]
        self.cpu_k, self.cpu_v = mk(ncpu, kshape), mk(ncpu, vshape)
        if pin:
            self.cpu_k, self.cpu_v = [t.pin_memory() for t in self.cpu_k], [t.pin_memory() for t in self.cpu_v]
        ntok = sum(len(op["slots"]) for op in TRACE["ops"] if op["op"] == "reshape_and_cache")
        self.rows = torch.randn((ntok, 2, heads, head_size), generator=g).to(dtype)     # token uid -> (k row, v row)
        self.device, self.next_uid = device, 0
        self.history = {}                # seq id -> list of token uids (lineage)

    def step(self, op):
        kind = op["op"]
        if kind == "reshape_and_cache":
            n = len(op["slots"])
            uids = list(range(self.next_uid, self.next_uid + n))
            self.next_uid += n
            if op["phase"] == "prompt":
                for uid, owner in zip(uids, op["seq_ids"]):
                    for sid in op["sharers"][str(owner)]:
                        self.history.setdefault(sid, []).append(uid)
            else:
                for uid, sid in zip(uids, op["seq_ids"]):
                    self.history[sid].append(uid)
            rows = self.rows[uids[0]:uids[0] + n].to(self.device)
            slots = torch.tensor(op["slots"], dtype=torch.int32, device=self.device)
            for l in range(self.layers):
                # each layer stores a different rotation of the row so layers cannot be mixed up
                r = torch.roll(rows, l, dims=-1)
                self.ops.reshape_and_cache(r[:, 0], r[:, 1], self.gpu_k[l], self.gpu_v[l], slots)
        elif kind == "fork":
            self.history[op["child"]] = list(self.history[op["parent"]])
        elif kind == "copy_blocks":
            self.ops.copy_blocks(self.gpu_k, self.gpu_v, {int(k): v for k, v in op["map"].items()})
        elif kind == "swap_out":
            m = {int(k): v for k, v in op["map"].items()}
            for l in range(self.layers):               # worker/cache_engine.py:108-118
                self.ops.swap_blocks(self.gpu_k[l], self.cpu_k[l], m)
                self.ops.swap_blocks(self.gpu_v[l], self.cpu_v[l], m)
        elif kind == "swap_in":
            m = {int(k): v for k, v in op["map"].items()}
            for l in range(self.layers):
                self.ops.swap_blocks(self.cpu_k[l], self.gpu_k[l], m)
                self.ops.swap_blocks(self.cpu_v[l], self.gpu_v[l], m)

    def tables(self, op):
        mb = max(len(t) for t in op["block_tables"])
        bt = torch.zeros((len(op["block_tables"]), mb), dtype=torch.int32)
        for i, t in enumerate(op["block_tables"]):
            bt[i, :len(t)] = torch.tensor(t, dtype=torch.int32)
        return bt, torch.tensor(op["context_lens"], dtype=torch.int32)


def _lineage_ok(rp: Replay, op) -> bool:
    """Every sequence reads back, through its block table, the K/V rows of its own history."""
    B = rp.B
    for table, ctx, sid in zip(op["block_tables"], op["context_lens"], op["seq_ids"]):
        hist = rp.history[sid]
        if len(hist) != ctx:
            return False
        slots = torch.tensor([table[i // B] * B + i % B for i in range(ctx)], dtype=torch.int32)
        for l in (0, rp.layers - 1):
            got = torch.zeros((ctx, 2, *rp.rows.shape[2:]), dtype=rp.rows.dtype)
            paged_oracle.gather_cached_kv(got[:, 0], got[:, 1], rp.gpu_k[l], rp.gpu_v[l], slots)
            if not torch.equal(got, torch.roll(rp.rows[hist], l, dims=-1)):
                return False
    return True


def test_trace_comes_from_the_reference_block_manager_and_is_nontrivial():
    kinds = [op["op"] for op in TRACE["ops"]]
    assert "reference BlockSpaceManager" in TRACE["generator"]
    assert kinds.count("copy_blocks") >= 20 and kinds.count("fork") >= 50
    assert kinds.count("swap_out") == 1 and kinds.count("swap_in") == 1
    pairs = sum(len(v) for op in TRACE["ops"] if op["op"] == "copy_blocks" for v in op["map"].values())
    assert pairs >= 100
    # one source forked to several destinations in one call (parallel sampling's first decode step)
    assert any(len(v) >= 3 for op in TRACE["ops"] if op["op"] == "copy_blocks" for v in op["map"].values())


def test_oracles_replay_to_the_same_bits_and_preserve_every_lineage():
    a = Replay(paged_oracle, 2, 3, 32, torch.float16, "cpu")
    b = Replay(c_oracle, 2, 3, 32, torch.float16, "cpu")
    checked = 0
    for i, op in enumerate(TRACE["ops"]):
        a.step(op)
        b.step(op)
        if op["op"] == "attention" and (checked < 3 or i % 7 == 0 or i == len(TRACE["ops"]) - 1):
            assert _lineage_ok(a, op), f"lineage broken at op {i}"
            checked += 1
    for x, y in zip(a.gpu_k + a.gpu_v + a.cpu_k + a.cpu_v, b.gpu_k + b.gpu_v + b.cpu_k + b.cpu_v):
        assert torch.equal(pt.bits(x), pt.bits(y))
    assert checked >= 8


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_gpu_replay_is_bit_exact_and_attention_matches(dtype):
    from vllm_b200 import cache_ops
    heads, head_size, layers = 40, 128, 3              # OPT-13B head shape; 3 of its 40 layers
    dev = "cuda:0"
    want = Replay(c_oracle, layers, heads, head_size, dtype, "cpu")
    got = Replay(cache_ops, layers, heads, head_size, dtype, dev, pin=True)
    n_att = 0
    for i, op in enumerate(TRACE["ops"]):
        want.step(op)
        got.step(op)
        if op["op"] in ("copy_blocks", "swap_out", "swap_in") or i % 11 == 0:
            torch.cuda.synchronize()
            for l in range(layers):
                assert torch.equal(pt.bits(got.gpu_k[l]), pt.bits(want.gpu_k[l])), f"op {i} {op['op']} key layer {l}"
                assert torch.equal(pt.bits(got.gpu_v[l]), pt.bits(want.gpu_v[l])), f"op {i} {op['op']} value layer {l}"
                assert torch.equal(pt.bits(got.cpu_k[l]), pt.bits(want.cpu_k[l])), f"op {i} host key layer {l}"
                assert torch.equal(pt.bits(got.cpu_v[l]), pt.bits(want.cpu_v[l])), f"op {i} host value layer {l}"
        if op["op"] == "attention" and (i % 5 == 0 or i == len(TRACE["ops"]) - 1):
            bt, cl = want.tables(op)
            g = torch.Generator().manual_seed(i)
            q = torch.randn((len(cl), 3, heads, head_size), generator=g).to(dtype)[:, 0]
            host = pt.DecodeProblem(q, want.gpu_k[1], want.gpu_v[1], bt, cl, head_size ** -0.5, want.B, int(cl.max()))
            devp = pt.DecodeProblem(q._base.to(dev)[:, 0], got.gpu_k[1], got.gpu_v[1], bt.to(dev), cl.to(dev),
                                    host.scale, want.B, host.max_context_len)
            pt.assert_decode_close(pt.run_decode(devp), pt.oracle_decode(host), dtype, f"attention at op {i}")
            n_att += 1
    assert n_att >= 5
