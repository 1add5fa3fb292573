#!/usr/bin/env python
"""Benchmark of the PagedAttention decode path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]

A "step" is one decode-attention pass over one batch of synthetic paged KV (random K/V pages,
every page scattered through a random block table).  The default workload for every N is
BASELINE.json configs[2], the one the metric's "1/2/4/8 B200" refers to: LLaMA-13B decode,
40 heads x 128, 256 sequences x 2048-token context, bf16, heads sharded over the N GPUs
(40/N per GPU, replicated block tables, no collective on the path) -> "scaling": "strong".
At N=1 the line also carries configs[1] (LLaMA-7B, fp16) under "also".

Besides the headline numbers the ONE JSON line (rank 0) carries, all measured in this run:
  check                 rank 0's output rows against the CPU oracle (a sample of sequences, all local heads)
  e2e                   host-buffer round trip per step, serial and pipelined
  gpu_reference_kernel  the reference's own CUDA kernel (oracle/_ref, built unmodified for sm_100) on the same
                        tensors, timed in a SUBPROCESS so that this process never maps it
  also (N = 1)          BASELINE.json configs[1] and configs[4] decode shapes, the configs[3] cache movers, the
                        prompt kernel next to the library kernel
  allreduce (N > 1)     the row-parallel all-reduce that follows attention: multimem / peer / NCCL
See DESIGN.md "Measurement" for how each field is obtained.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "PagedAttention decode HBM GB/s (% of ~8 TB/s) and seqs/s at 1/2/4/8 B200"
NOMINAL_HBM_GBS = 8000.0
FALLBACK_HBM_GBS = 6650.0      # /opt/skills/guides/B200_PROFILING.md, used only without MEASURED_PEAKS.json

# name -> total heads, head size, block size, sequences, context, dtype, sharded over GPUs?
WORKLOADS = {
    "llama13b_tp": dict(heads=40, head_size=128, block_size=16, seqs=256, ctx=2048, dtype="bf16", shard=True,
                        note="BASELINE.json configs[2]: LLaMA-13B paged decode head-sharded at 1/2/4/8 B200"),
    "llama7b": dict(heads=32, head_size=128, block_size=16, seqs=256, ctx=1024, dtype="fp16", shard=True,
                    note="BASELINE.json configs[1]: LLaMA-7B decode attention on 1 B200"),
    "neox20b_tp8": dict(heads=8, head_size=96, block_size=16, seqs=64, ctx=8192, dtype="fp16", shard=False,
                        note="BASELINE.json configs[4]: GPT-NeoX-20B per-GPU shape at TP=8 (8 heads x 96)"),
    "opt66b_tp8": dict(heads=9, head_size=128, block_size=16, seqs=64, ctx=8192, dtype="fp16", shard=False,
                       note="BASELINE.json configs[4]: OPT-66B per-GPU shape at TP=8 (9 heads x 128)"),
}


def _torch_dtype(name):
    import torch
    return {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[name]


def algorithmic_bytes(seqs, heads, head_size, block_size, ctx, esize):
    """K and V of exactly ctx tokens per head, q read + out write per head,
    the used block-table entries and the context length once per sequence."""
    per_seq = heads * (2 * ctx * head_size * esize + 2 * head_size * esize) + ((ctx + block_size - 1) // block_size) * 4 + 4
    return seqs * per_seq


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------
# clocks and throttle reasons during the timed region (NVML, the source nvidia-smi reads)
# ---------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, device_index: int, period_s: float = 0.01):
        super().__init__(daemon=True)
        self.period = period_s
        self.samples = []          # (wall time, sm MHz, reasons bitmask, power W)
        self.stop_flag = threading.Event()
        self.handle = None
        self.max_mhz = None
        try:
            import pynvml
            import torch
            self.nv = pynvml
            pynvml.nvmlInit()
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(device_index).uuid)
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid)
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
                idx = int(vis.split(",")[device_index]) if vis and vis.split(",")[device_index].isdigit() else device_index
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        except Exception as e:     # no NVML: the clocks object says so
            self.error = repr(e)
            self.handle = None

    def run(self):
        if self.handle is None:
            return
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                mhz = int(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
                try:
                    rs = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                except Exception:
                    rs = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(self.handle) / 1000.0
                except Exception:
                    pw = None
                self.samples.append((time.perf_counter(), mhz, rs, pw))
            except Exception:
                pass
            time.sleep(self.period)

    def summary(self, windows):
        """Median SM clock and the union of throttle reasons over samples inside the timed windows."""
        if self.handle is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": getattr(self, "error", "no NVML")}
        inside = [s for s in self.samples if any(a <= s[0] <= b for a, b in windows)]
        use = inside if inside else self.samples
        clocks = sorted(s[1] for s in use)
        bits = 0
        for s in use:
            bits |= s[2]
        reasons = [name for bit, name in self.REASONS.items() if bits & bit]
        powers = [s[3] for s in use if s[3] is not None]
        return {"sm_mhz": clocks[len(clocks) // 2] if clocks else None, "sm_max_mhz": self.max_mhz,
                "reasons": reasons, "samples": len(inside), "power_w_max": max(powers) if powers else None}


# ---------------------------------------------------------------------------
# synthetic problem on the device
# ---------------------------------------------------------------------------
def make_problem(wl, heads_local, device, seed=0):
    """Random KV pages for `heads_local` heads; block tables are a random permutation of all
    blocks (identical on every rank: same seed), so every page is a scattered chunk."""
    import torch
    dtype = _torch_dtype(wl["dtype"])
    S, D, B, L = wl["seqs"], wl["head_size"], wl["block_size"], wl["ctx"]
    x = 16 // torch.empty((), dtype=dtype).element_size()
    nblk = (L + B - 1) // B
    NB = S * nblk
    g = torch.Generator(device=device).manual_seed(seed)
    kc = torch.empty((NB, heads_local, D // x, B, x), dtype=dtype, device=device).normal_(generator=g)
    vc = torch.empty((NB, heads_local, D, B), dtype=dtype, device=device).normal_(generator=g)
    qkv = torch.empty((S, 3, heads_local, D), dtype=dtype, device=device).normal_(generator=g)
    gt = torch.Generator().manual_seed(seed)            # host generator: same tables on every rank
    tables = torch.randperm(NB, generator=gt).to(torch.int32).view(S, nblk).contiguous()
    lens = torch.full((S,), L, dtype=torch.int32)
    return dict(q=qkv[:, 0], kc=kc, vc=vc, tables=tables.to(device), lens=lens.to(device), tables_host=tables,
                lens_host=lens, scale=float(D) ** -0.5, B=B, L=L, dtype=dtype)


def run_kernel_steps(p, steps, warmup, barrier):
    """W untimed + K timed calls of the operator on device-resident inputs; returns (ms total, wall window, launches)."""
    import torch
    from vllm_b200 import _cabi, attention_ops
    out = torch.empty(p["q"].shape, dtype=p["dtype"], device=p["q"].device)

    def step():
        attention_ops.single_query_cached_kv_attention(out, p["q"], p["kc"], p["vc"], p["scale"], p["tables"],
                                                       p["lens"], p["B"], p["L"])
    for _ in range(warmup):
        step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    n0 = _cabi.launch_count()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    barrier()
    return e0.elapsed_time(e1), (t0, t1), _cabi.launch_count() - n0, out


def run_e2e_steps(p, steps, warmup, barrier):
    """The same operator called the way a worker calls it each decode step, from HOST buffers:
    every step copies that step's inputs (query, block tables, context lengths; one pinned staging
    arena, one copy) from pinned host memory to the device, runs the operator and reads the result back to pinned host memory.
    The KV cache is the worker's resident state (vllm/worker/cache_engine.py allocates it on the
    GPU once); it is not a per-step input."""
    import torch
    from vllm_b200 import attention_ops
    dev = p["q"].device
    # one pinned staging arena for the step's inputs [query | block tables | context lengths] and one device arena of
    # the same layout: the three inputs cross PCIe as ONE copy per step (three small copies cost ~8 us of latency each)
    q_src, t_src, l_src = p["q"].contiguous().cpu(), p["tables_host"], p["lens_host"]
    sizes = [t.numel() * t.element_size() for t in (q_src, t_src, l_src)]
    offs = [0, (sizes[0] + 255) // 256 * 256]
    offs.append(offs[1] + (sizes[1] + 255) // 256 * 256)
    total = offs[2] + sizes[2]
    arena_host = torch.empty(total, dtype=torch.uint8).pin_memory()
    arena_dev = torch.empty(total, dtype=torch.uint8, device=dev)

    def view(arena, i, like):
        return arena[offs[i]:offs[i] + sizes[i]].view(like.dtype).view(like.shape)
    q_host, t_host, l_host = view(arena_host, 0, q_src), view(arena_host, 1, t_src), view(arena_host, 2, l_src)
    q_host.copy_(q_src); t_host.copy_(t_src); l_host.copy_(l_src)
    q_dev, t_dev, l_dev = view(arena_dev, 0, q_src), view(arena_dev, 1, t_src), view(arena_dev, 2, l_src)
    out_host = torch.empty(q_src.shape, dtype=p["dtype"]).pin_memory()
    out_dev = torch.empty_like(q_dev)
    h2d = total
    d2h = out_host.numel() * out_host.element_size()

    def step():
        arena_dev.copy_(arena_host, non_blocking=True)
        attention_ops.single_query_cached_kv_attention(out_dev, q_dev, p["kc"], p["vc"], p["scale"], t_dev, l_dev,
                                                       p["B"], p["L"])
        out_host.copy_(out_dev, non_blocking=True)
        torch.cuda.current_stream().synchronize()          # the host has this step's result

    for _ in range(warmup):
        step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    barrier()
    return e0.elapsed_time(e1), (t0, t1), h2d, d2h


def run_e2e_graph(p, steps, warmup, barrier):
    """run_e2e_steps with the three stream operations of a step -- staging arena host -> device, the operator, output
    device -> pinned host -- captured once into a CUDA graph and replayed: one launch per step instead of three API
    calls (the reference's own answer to fixed launch cost, parallel_state.py:531-556).  Same contract as the serial
    loop: the same bytes cross PCIe inside every timed step, the host synchronises on every step's result before it
    issues the next one.  pa_decode_attention_reserve sizes the library's scratch before the capture."""
    import torch
    from vllm_b200 import attention_ops
    dev = p["q"].device
    q_src, t_src, l_src = p["q"].contiguous().cpu(), p["tables_host"], p["lens_host"]
    sizes = [t.numel() * t.element_size() for t in (q_src, t_src, l_src)]
    offs = [0, (sizes[0] + 255) // 256 * 256]
    offs.append(offs[1] + (sizes[1] + 255) // 256 * 256)
    total = offs[2] + sizes[2]
    arena_host = torch.empty(total, dtype=torch.uint8).pin_memory()
    arena_dev = torch.empty(total, dtype=torch.uint8, device=dev)

    def view(arena, i, like):
        return arena[offs[i]:offs[i] + sizes[i]].view(like.dtype).view(like.shape)
    for i, src in enumerate((q_src, t_src, l_src)):
        view(arena_host, i, src).copy_(src)
    q_dev, t_dev, l_dev = view(arena_dev, 0, q_src), view(arena_dev, 1, t_src), view(arena_dev, 2, l_src)
    out_host = torch.empty(q_src.shape, dtype=p["dtype"]).pin_memory()
    out_dev = torch.empty_like(q_dev)

    def body():
        arena_dev.copy_(arena_host, non_blocking=True)
        attention_ops.single_query_cached_kv_attention(out_dev, q_dev, p["kc"], p["vc"], p["scale"], t_dev, l_dev,
                                                       p["B"], p["L"])
        out_host.copy_(out_dev, non_blocking=True)

    side = torch.cuda.Stream(device=dev)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        body()                                     # eager once: sizes every scratch buffer before the capture
        side.synchronize()
        with torch.cuda.graph(graph, stream=side):
            body()
    torch.cuda.synchronize()
    out_host.zero_()

    def step():
        graph.replay()
        torch.cuda.current_stream().synchronize()          # the host has this step's result

    for _ in range(warmup):
        step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    barrier()
    ok = bool(torch.equal(out_host, out_dev.cpu()) and out_host.float().abs().sum() > 0)
    del graph
    return e0.elapsed_time(e1), (t0, t1), ok


def time_gpu_config1(dev):
    """BASELINE.json configs[0] (7 x 40 x 128, block 16, fp32, 1024 blocks, contexts <= 4096 from random.seed(0),
    aliasing block tables) through the fp32 decode kernel: seconds per call, device-timed -- the GPU figure beside
    cpu_baseline.config1_seconds."""
    import random
    import torch
    from vllm_b200 import attention_ops
    rnd = random.Random(0)
    S, H, D, B, NB = 7, 40, 128, 16, 1024
    ctx = [rnd.randint(1, 4096) for _ in range(S)]
    mb = (max(ctx) + B - 1) // B
    tables = torch.tensor([[rnd.randint(0, NB - 1) for _ in range(mb)] for _ in range(S)], dtype=torch.int32, device=dev)
    g = torch.Generator(device=dev).manual_seed(0)
    q = torch.empty((S, H, D), device=dev).uniform_(-1e-3, 1e-3, generator=g)
    kc = torch.empty((NB, H, D // 4, B, 4), device=dev).uniform_(-1e-3, 1e-3, generator=g)
    vc = torch.empty((NB, H, D, B), device=dev).uniform_(-1e-3, 1e-3, generator=g)
    out = torch.empty_like(q)
    lens = torch.tensor(ctx, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # the caches (84 MB) fit the L2: flush between calls

    def call():
        flush.zero_()
        attention_ops.single_query_cached_kv_attention(out, q, kc, vc, D ** -0.5, tables, lens, B, max(ctx))
    ms = _time_events(call, iters=20) - _time_events(lambda: flush.zero_(), iters=20)
    return ms / 1e3


def run_e2e_pipelined(p, steps, warmup, barrier):
    """The same host-buffer round trip with the fixed copy latencies overlapped: two staging arenas and two
    output buffers; step i's inputs go up on an upload stream while step i-1's kernel runs, step i's result
    comes down on a download stream while step i+1 runs.  Contract: the result of step i is in host memory
    before step i+2 is issued (the host blocks on it then), and every step's bytes still cross PCIe inside
    the timed region."""
    import torch
    from vllm_b200 import attention_ops
    dev = p["q"].device
    q_src, t_src, l_src = p["q"].contiguous().cpu(), p["tables_host"], p["lens_host"]
    sizes = [t.numel() * t.element_size() for t in (q_src, t_src, l_src)]
    offs = [0, (sizes[0] + 255) // 256 * 256]
    offs.append(offs[1] + (sizes[1] + 255) // 256 * 256)
    total = offs[2] + sizes[2]

    def view(arena, i, like):
        return arena[offs[i]:offs[i] + sizes[i]].view(like.dtype).view(like.shape)
    slots = []
    for _ in range(2):
        ah = torch.empty(total, dtype=torch.uint8).pin_memory()
        ad = torch.empty(total, dtype=torch.uint8, device=dev)
        for i, src in enumerate((q_src, t_src, l_src)):
            view(ah, i, src).copy_(src)
        slots.append(dict(ah=ah, ad=ad, q=view(ad, 0, q_src), t=view(ad, 1, t_src), l=view(ad, 2, l_src),
                          out_d=torch.empty(q_src.shape, dtype=p["dtype"], device=dev),
                          out_h=torch.empty(q_src.shape, dtype=p["dtype"]).pin_memory(),
                          up=torch.cuda.Event(), done=torch.cuda.Event(), down=torch.cuda.Event(), used=False))
    main_s = torch.cuda.current_stream()
    up_s, down_s = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def step(i):
        s = slots[i & 1]
        if s["used"]:
            s["down"].synchronize()                    # the host holds step i-2's result before its buffers are reused
        with torch.cuda.stream(up_s):
            if s["used"]:
                up_s.wait_event(s["done"])             # step i-2's kernel has read this arena
            s["ad"].copy_(s["ah"], non_blocking=True)
            s["up"].record(up_s)
        main_s.wait_event(s["up"])
        attention_ops.single_query_cached_kv_attention(s["out_d"], s["q"], p["kc"], p["vc"], p["scale"], s["t"], s["l"],
                                                       p["B"], p["L"])
        s["done"].record(main_s)
        with torch.cuda.stream(down_s):
            down_s.wait_event(s["done"])
            s["out_h"].copy_(s["out_d"], non_blocking=True)
            s["down"].record(down_s)
        s["used"] = True

    def drain():
        for s in slots:
            if s["used"]:
                main_s.wait_event(s["down"])
                s["down"].synchronize()

    for i in range(warmup):
        step(i)
    drain()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    for i in range(steps):
        step(warmup + i)
    drain()                                            # the last results are on the host inside the timed region
    e1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    barrier()
    return e0.elapsed_time(e1), (t0, t1)


# ---------------------------------------------------------------------------
# CPU side: the reference's PyTorch path (restated in oracle/paged_oracle.py) on host cores
# ---------------------------------------------------------------------------
def make_host_sample(wl, nseq, seed=0):
    """`nseq` sequences of the workload with ALL heads, on the host, compact cache (only their pages)."""
    import torch
    dtype = _torch_dtype(wl["dtype"])
    H, D, B, L = wl["heads"], wl["head_size"], wl["block_size"], wl["ctx"]
    x = 16 // torch.empty((), dtype=dtype).element_size()
    nblk = (L + B - 1) // B
    NB = nseq * nblk
    g = torch.Generator().manual_seed(seed)
    # uniform values are as good as normal ones for timing and much cheaper to draw
    kc = (torch.rand((NB, H, D // x, B, x), generator=g) - 0.5).to(dtype)
    vc = (torch.rand((NB, H, D, B), generator=g) - 0.5).to(dtype)
    q = (torch.rand((nseq, H, D), generator=g) - 0.5).to(dtype)
    tables = torch.randperm(NB, generator=g).to(torch.int32).view(nseq, nblk).contiguous()
    lens = torch.full((nseq,), L, dtype=torch.int32)
    return q, kc, vc, tables, lens


def use_all_host_threads():
    """The CPU legs use every host core this process may run on (torchrun exports OMP_NUM_THREADS=1)."""
    import torch
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def time_cpu_reference(wl, nseq, steps=1, warmup=0):
    """seconds per pass of the PyTorch loop oracle over `nseq` sequences (all heads, full context)."""
    import torch
    from oracle import paged_oracle
    use_all_host_threads()
    q, kc, vc, tables, lens = make_host_sample(wl, nseq)
    out = torch.empty_like(q)
    for _ in range(warmup):
        paged_oracle.decode_attention_loop(out, q, kc, vc, tables, lens)
    t0 = time.perf_counter()
    for _ in range(steps):
        paged_oracle.decode_attention_loop(out, q, kc, vc, tables, lens)
    return (time.perf_counter() - t0) / steps


def time_cpu_config1():
    """BASELINE.json configs[0] as the reference test builds it (tests/kernels/test_attention.py:159-224,270-287):
    7 sequences x 40 heads x 128, block 16, fp32, 1024 blocks, context lengths and aliasing block tables from
    random.seed(0); one pass of the PyTorch loop reference on all host threads.  Values are uniform like the
    test's (their bits do not matter for the timing)."""
    import random
    import torch
    from oracle import paged_oracle
    threads = use_all_host_threads()
    rnd = random.Random(0)
    S, H, D, B, NB = 7, 40, 128, 16, 1024
    ctx = [rnd.randint(1, 4096) for _ in range(S)]
    mb = (max(ctx) + B - 1) // B
    tables = torch.tensor([[rnd.randint(0, NB - 1) for _ in range(mb)] for _ in range(S)], dtype=torch.int32)
    g = torch.Generator().manual_seed(0)
    q = torch.rand((S, H, D), generator=g) * 2e-3 - 1e-3
    kc = torch.rand((NB, H, D // 4, B, 4), generator=g) * 2e-3 - 1e-3
    vc = torch.rand((NB, H, D, B), generator=g) * 2e-3 - 1e-3
    out = torch.empty_like(q)
    lens = torch.tensor(ctx, dtype=torch.int32)
    t0 = time.perf_counter()
    paged_oracle.decode_attention_loop(out, q, kc, vc, tables, lens)
    sec = time.perf_counter() - t0
    return {"seconds": sec, "seqs_per_s": S / sec, "cores": threads,
            "shape": "7 seqs x 40 heads x 128, block 16, fp32, 1024 blocks, contexts <= 4096 (random.seed(0))"}


def cpu_sample_size(wl, budget_s, cap):
    """Sequences per CPU pass so that one pass takes about `budget_s` (calibrated on one sequence)."""
    one = time_cpu_reference(wl, 1)
    return max(1, min(cap, int(budget_s / max(one, 1e-3)))), one


def workload_config(name, wl, n_gpus, heads_local):
    return {"workload": name, "note": wl["note"], "num_seqs": wl["seqs"], "total_heads": wl["heads"],
            "heads_per_gpu": heads_local, "head_size": wl["head_size"], "block_size": wl["block_size"],
            "context_len": wl["ctx"], "kv_dtype": wl["dtype"],
            "parallelism": f"heads sharded over {n_gpus} GPU(s), no collective" if wl["shard"] else f"{n_gpus} replica(s)"}


def run_reference(args, wl):
    """--impl reference: the reference's own CPU implementation of the path (its PyTorch test
    reference, restated in oracle/paged_oracle.py because /root/reference is absent on the GPU box),
    all host threads, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    nseq, one = cpu_sample_size(wl, budget_s=120.0 / max(1, args.steps + args.warmup), cap=16)
    sec = time_cpu_reference(wl, nseq, steps=args.steps, warmup=args.warmup)
    value = nseq / sec
    esize = 2 if wl["dtype"] != "fp32" else 4
    gbs = algorithmic_bytes(nseq, wl["heads"], wl["head_size"], wl["block_size"], wl["ctx"], esize) / sec / 1e9
    sample = f"{nseq} of {wl['seqs']} sequences per step (all {wl['heads']} heads, {wl['ctx']}-token contexts)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "seqs/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "strong" if wl["shard"] else "weak", "vs_baseline": None, "dtype": wl["dtype"],
            "data": "synthetic", "hbm_gbs": gbs,
            "config": workload_config(args.workload, wl, args.gpus, wl["heads"]),
            "cpu_baseline": {"value": value, "unit": "seqs/s", "cores": torch.get_num_threads(), "kind": "port",
                             "sample": sample, "host_cpus": os.cpu_count()},
            "e2e": {"value": value, "unit": "seqs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def traffic_from_profiles(workload, heads_local):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture of this
    workload's decode kernel at this many heads per GPU (profiles/roofline_traffic.json), or None when no
    capture of exactly that shape is on file."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f).get(workload, {}).get(str(heads_local), {}).get("dram_bytes_per_launch")
    except Exception:
        return None


# ---------------------------------------------------------------------------
# checker: rank 0's output against the CPU oracle on a sample of sequences (outside every timed region)
# ---------------------------------------------------------------------------
def dispatch_matrix_summary():
    """Worst cells of the committed sweep over the reference's 36-kernel dispatch set (tools/dispatch_matrix.py ->
    profiles/r3_dispatch_matrix.json, builder-run on a B200): not measured in this run, and keyed as such."""
    path = os.path.join(ROOT, "profiles", "r3_dispatch_matrix.json")
    try:
        d = json.load(open(path))
    except Exception:
        return {}
    out = {}
    cell = lambda r: f"{r['dtype']}/B{r['block_size']}/D{r['head_size']}"
    w = d.get("worst_frac_cell")
    if w:
        out["profiled_worst_dispatch_cell"] = cell(w)
        out["profiled_worst_dispatch_cell_gbs"] = w["ours_gbs"]
        out["profiled_worst_dispatch_cell_frac_of_8tbs"] = w["frac_of_8tbs"]
    v = d.get("worst_speedup_cell")
    if v:
        out["profiled_worst_speedup_cell"] = cell(v)
        out["profiled_worst_speedup_vs_reference_cuda_kernel"] = v["speedup"]
    return out


def oracle_row_check(p, out, nrows=3):
    """Pulls the pages of `nrows` sequences back from the GPU cache, runs the C oracle (double precision) on
    them for all local heads and compares with the kernel's rows at the north-star tolerance of the dtype."""
    import torch
    from oracle import c_oracle
    S = p["q"].shape[0]
    rows = sorted(set([0, S // 2, S - 1][:nrows]))
    idx = torch.tensor(rows, dtype=torch.long)
    tables, lens = p["tables_host"][idx], p["lens_host"][idx]
    used = torch.unique(tables)
    remap = torch.zeros(int(used.max()) + 1, dtype=torch.int32)
    remap[used.long()] = torch.arange(len(used), dtype=torch.int32)
    dev = p["q"].device
    kc = p["kc"][used.long().to(dev)].cpu()
    vc = p["vc"][used.long().to(dev)].cpu()
    q = p["q"][idx.to(dev)].cpu().contiguous()
    want = torch.empty(q.shape, dtype=q.dtype)
    c_oracle.decode_attention(want, q, kc, vc, p["scale"], remap[tables.long()].contiguous(), lens.contiguous())
    got = out[idx.to(dev)].float().cpu()
    w = want.float()
    tol = {torch.float16: 1e-3, torch.bfloat16: 1e-2, torch.float32: 1e-4}[p["dtype"]]
    err = (got - w).abs()
    bound = tol * float(w.abs().max()) + tol * w.abs()
    return {"rows": rows, "heads": int(q.shape[1]), "max_abs_err": float(err.max()), "max_abs_want": float(w.abs().max()),
            "tolerance": tol, "ok": bool((err <= bound).all() and torch.isfinite(got).all()),
            "checker": "oracle/paged_oracle.c (double precision) on the pages of these sequences"}


# ---------------------------------------------------------------------------
# the reference's own CUDA kernel on the same tensors (subprocess: this process never maps oracle/_ref)
# ---------------------------------------------------------------------------
def ref_kernel_leg(workload, heads_local, device_index, iters=10):
    """Child process: loads oracle/_ref/ref_attention_ops.so (the reference's csrc/attention compiled unmodified
    for sm_100 by oracle/build_ref.py) and times it next to ours on the same seeded tensors."""
    import importlib.util
    import torch
    so = os.path.join(ROOT, "oracle", "_ref", "ref_attention_ops.so")
    if not os.path.exists(so):
        print(json.dumps({"unavailable": "oracle/_ref/ref_attention_ops.so is not built (oracle/build_ref.py needs /root/reference)"}))
        return
    torch.cuda.set_device(device_index)
    dev = torch.device("cuda", device_index)
    wl = WORKLOADS[workload]
    p = make_problem(wl, heads_local, dev)
    if p["kc"].numel() >= 2 ** 31:
        print(json.dumps({"unavailable": "the reference kernel indexes with 32-bit int (attention_kernels.cu:157,243): "
                                         f"{p['kc'].numel()} cache elements overflow it"}))
        return
    spec = importlib.util.spec_from_file_location("ref_attention_ops", so)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    from vllm_b200 import attention_ops
    out_a = torch.empty(p["q"].shape, dtype=p["dtype"], device=dev)
    out_b = torch.empty_like(out_a)

    def timed(fn):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    ms_ours = timed(lambda: attention_ops.single_query_cached_kv_attention(out_a, p["q"], p["kc"], p["vc"], p["scale"],
                                                                          p["tables"], p["lens"], p["B"], p["L"]))
    ms_ref = timed(lambda: ref.single_query_cached_kv_attention(out_b, p["q"], p["kc"], p["vc"], p["scale"], p["tables"],
                                                                p["lens"], p["B"], p["L"]))
    nbytes = algorithmic_bytes(wl["seqs"], heads_local, wl["head_size"], wl["block_size"], wl["ctx"], p["kc"].element_size())
    diff = float((out_a.float() - out_b.float()).abs().max() / out_b.float().abs().max())
    print(json.dumps({"workload": workload, "heads_per_gpu": heads_local, "ms": ms_ref, "gbs": nbytes / ms_ref / 1e6,
                      "ours_ms": ms_ours, "ours_gbs": nbytes / ms_ours / 1e6, "speedup": ms_ref / ms_ours,
                      "max_rel_diff": diff,
                      "what": "reference csrc/attention/attention_kernels.cu compiled unmodified for sm_100, same tensors, "
                              f"CUDA events over {iters} back-to-back calls each"}))


def gpu_reference_kernel(workload, heads_local, device_index):
    try:
        env = {k: v for k, v in os.environ.items()
               if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS")}
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--leg", "ref-kernel", "--workload", workload,
                            "--heads-local", str(heads_local), "--device-index", str(device_index)],
                           capture_output=True, text=True, timeout=300, env=env)
        lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
        if r.returncode != 0 or not lines:
            return {"unavailable": f"child exited {r.returncode}: {r.stderr.strip()[-300:]}"}
        return json.loads(lines[-1])
    except Exception as e:
        return {"unavailable": repr(e)[:300]}


# ---------------------------------------------------------------------------
# N = 1 extras: configs[4] decode shapes, configs[3] cache movers, the prompt kernel
# ---------------------------------------------------------------------------
def _time_events(fn, iters, warm=3):
    import torch
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def decode_block(name, dev, steps, warmup, peak):
    """One more decode workload on this GPU: device-timed value, roofline against the same measured peak, e2e."""
    import torch
    wl = WORKLOADS[name]
    p = make_problem(wl, wl["heads"], dev)
    nobar = lambda: None
    ms, _, _, out = run_kernel_steps(p, steps, warmup, nobar)
    chk = oracle_row_check(p, out, nrows=2)
    mse, _, _, _ = run_e2e_steps(p, steps, warmup, nobar)
    esize = p["kc"].element_size()
    b = algorithmic_bytes(wl["seqs"], wl["heads"], wl["head_size"], wl["block_size"], wl["ctx"], esize)
    t = ms / steps / 1e3
    blk = {"config": workload_config(name, wl, 1, wl["heads"]), "dtype": wl["dtype"],
           "value": wl["seqs"] / t, "unit": "seqs/s", "ms_per_step": t * 1e3, "hbm_gbs": b / t / 1e9,
           "roofline": {"bound": "hbm", "achieved": b / t / 1e9, "peak": peak, "unit": "GB/s", "frac": b / t / 1e9 / peak,
                        "traffic": traffic_from_profiles(name, wl["heads"]), "algorithmic_bytes_per_launch": b},
           "frac_of_8tbs": b / t / 1e9 / NOMINAL_HBM_GBS, "e2e_value": wl["seqs"] / (mse / steps / 1e3), "check": chk}
    del p, out
    torch.cuda.empty_cache()
    return blk


def movers_block(dev, peak):
    """BASELINE.json configs[3] (OPT-13B, 40 layers x 40 heads x 128, block 16, fp16): copy_blocks for one
    decode step of 64 groups forking 3 siblings each (192 pairs, every layer), swap_blocks of 512 scattered
    blocks each way, reshape_and_cache of a decode step (256 tokens) and of a prompt batch (2560 tokens).
    Algorithmic bytes per SURVEY.md section 8(d).  Every result is checked bit for bit against torch indexing."""
    import torch
    from vllm_b200 import cache_ops
    layers, H, D, B, NB = 40, 40, 128, 16, 1024
    g = torch.Generator(device=dev).manual_seed(7)
    kcs = [torch.empty((NB, H, D // 8, B, 8), dtype=torch.float16, device=dev).normal_(generator=g) for _ in range(layers)]
    vcs = [torch.empty((NB, H, D, B), dtype=torch.float16, device=dev).normal_(generator=g) for _ in range(layers)]
    blk_bytes = H * D * B * 2
    res = {"config": {"workload": "opt13b_cache_movers", "note": "BASELINE.json configs[3]: OPT-13B parallel sampling n=4 / "
                      "beam 4 cache traffic", "layers": layers, "num_heads": H, "head_size": D, "block_size": B,
                      "num_blocks": NB, "kv_dtype": "fp16", "block_bytes": blk_bytes}}
    # copy_blocks
    mapping = {}
    for i in range(192):
        mapping.setdefault(i // 3, []).append(512 + i)
    ms = _time_events(lambda: cache_ops.copy_blocks(kcs, vcs, mapping), iters=20)
    ok = all(torch.equal(kcs[l][512 + i], kcs[l][i // 3]) and torch.equal(vcs[l][512 + i], vcs[l][i // 3])
             for l in (0, layers - 1) for i in (0, 95, 191))
    # unique DRAM bytes: every destination block written once, every source block read once (its three destinations
    # share it through the L2), K and V, every layer
    nbytes = (192 + 64) * layers * 2 * blk_bytes
    res["copy_blocks"] = {"pairs": 192, "layers": layers, "ms": ms, "bit_exact": ok,
                          "roofline": {"bound": "hbm", "achieved": nbytes / ms / 1e6, "peak": peak, "unit": "GB/s",
                                       "frac": nbytes / ms / 1e6 / peak, "traffic": None, "algorithmic_bytes_per_launch": nbytes,
                                       "note": "unique DRAM bytes: 192 blocks written + 64 source blocks read per cache"}}
    # reshape_and_cache
    for N, label in ((256, "decode_step"), (512, "short_prompt"), (2560, "prompt_batch")):
        qkv = torch.empty((N, 3, H, D), dtype=torch.float16, device=dev).normal_(generator=g)
        if label == "decode_step":   # every token is the newest token of its own sequence -> its own block
            slots = (torch.randperm(NB, device=dev)[:N] * B + torch.randint(0, B, (N,), device=dev)).to(torch.int32)
        else:            # prompts of 512 tokens fill consecutive slots of scattered block runs
            starts = torch.randperm(NB // 32, device=dev)[:N // 512] * 32 * B
            slots = (starts[:, None] + torch.arange(512, device=dev)[None, :]).flatten().to(torch.int32)
        kc, vc = kcs[0], vcs[0]
        hint = "scattered" if label == "decode_step" else "runs"      # what PagedAttention passes (num_prompt_tokens)
        ms = _time_events(lambda: cache_ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, slots, slots=hint), iters=50)
        sl = slots.long()
        got_k = kc[sl // B, :, :, sl % B, :].reshape(N, H, D)        # [N, H, D/x, x] -> [N, H, D]
        got_v = vc[sl // B, :, :, sl % B]
        ok = bool(torch.equal(got_k, qkv[:, 1]) and torch.equal(got_v, qkv[:, 2]))
        nbytes = N * H * D * 2 * 2 * 2                     # K and V rows read + written
        res[f"reshape_and_cache_{label}"] = {
            "tokens": N, "ms": ms, "bit_exact": ok,
            "roofline": {"bound": "hbm", "achieved": nbytes / ms / 1e6, "peak": peak, "unit": "GB/s",
                         "frac": nbytes / ms / 1e6 / peak, "traffic": None, "algorithmic_bytes_per_launch": nbytes,
                         "note": "a launch this small is bound by launch + DRAM latency, not bandwidth"}}
    # swap_blocks
    cpu = torch.empty((NB, H, D, B), dtype=torch.float16).pin_memory()
    gp = torch.Generator().manual_seed(3)
    src_ids, dst_ids = torch.randperm(NB, generator=gp)[:512].tolist(), torch.randperm(NB, generator=gp)[:512].tolist()
    mp = {int(a): int(b) for a, b in zip(src_ids, dst_ids)}
    pcie = 63.0          # GB/s, PCIe Gen5 x16 per direction (nominal; no measured figure in MEASURED_PEAKS.json)
    for direction, (src, dst) in (("gpu_to_cpu", (vcs[0], cpu)), ("cpu_to_gpu", (cpu, vcs[1]))):
        ms = _time_events(lambda: cache_ops.swap_blocks(src, dst, mp), iters=5, warm=2)
        torch.cuda.synchronize()
        probe = [(src_ids[i], dst_ids[i]) for i in (0, 255, 511)]
        ok = all(torch.equal(dst[d].cpu(), src[s_].cpu()) for s_, d in probe)
        nbytes = 512 * blk_bytes
        res[f"swap_blocks_{direction}"] = {
            "blocks": 512, "ms": ms, "bit_exact": ok,
            "roofline": {"bound": "pcie", "achieved": nbytes / ms / 1e6, "peak": pcie, "unit": "GB/s",
                         "frac": nbytes / ms / 1e6 / pcie, "traffic": None, "algorithmic_bytes_per_launch": nbytes,
                         "peak_source": "nominal PCIe Gen5 x16, one direction"}}
    del kcs, vcs, cpu
    torch.cuda.empty_cache()
    return res


def prefill_block(dev):
    """The prompt kernel (pa_prefill_attention, tcgen05) next to the library kernel on this GPU (torch SDPA ->
    cuDNN) on the same tensors; roofline against the measured bf16 tensor peak."""
    import torch
    import torch.nn.functional as F
    from vllm_b200 import attention_ops
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            pk = json.load(f)
        peak, peak_src = float(pk["bf16_tflops"]), "measured (MEASURED_PEAKS.json bf16_tflops, cuBLAS burst)"
    except Exception:
        peak, peak_src = 1590.0, "fallback (B200_PROFILING.md 1.59 PFLOP/s)"
    rows = []
    for name, h, d, lens in (("llama13b_5x512", 40, 128, [512] * 5), ("llama13b_8x2048", 40, 128, [2048] * 8)):
        n = sum(lens)
        g = torch.Generator(device=dev).manual_seed(0)
        qkv = torch.empty((n, 3, h, d), dtype=torch.bfloat16, device=dev).normal_(generator=g)
        q, k, v = qkv[:, 0], qkv[:, 1], qkv[:, 2]
        out = torch.empty((n, h, d), dtype=torch.bfloat16, device=dev)
        scale = d ** -0.5
        ms = _time_events(lambda: attention_ops.multi_query_kv_attention(out, q, k, v, lens, scale), iters=20, warm=5)
        fl = sum(4 * d * h * (m * (m + 1) // 2) for m in lens)
        row = {"shape": name, "heads": h, "head_size": d, "prompt_lens": lens, "dtype": "bf16", "ms": ms,
               "roofline": {"bound": "tensor", "achieved": fl / ms / 1e9, "peak": peak, "unit": "TFLOP/s",
                            "frac": fl / ms / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                            "algorithmic_flops_per_launch": fl}}
        try:
            b, L = len(lens), lens[0]
            q4, k4, v4 = (t.reshape(b, L, h, d).transpose(1, 2) for t in (q, k, v))
            sd = lambda: F.scaled_dot_product_attention(q4, k4, v4, is_causal=True, scale=scale)
            ref = sd().transpose(1, 2).reshape(n, h, d)
            ms_sd = _time_events(sd, iters=20, warm=5)
            row["library_sdpa_ms"] = ms_sd
            row["library_sdpa_tflops"] = fl / ms_sd / 1e9
            row["speedup_vs_library"] = ms_sd / ms
            row["max_abs_diff_vs_library"] = float((ref.float() - out.float()).abs().max())
        except Exception as e:
            row["library_sdpa_error"] = repr(e)[:200]
        rows.append(row)
        del qkv, out
        torch.cuda.empty_cache()
    return rows


def cached_prefill_block(dev):
    """Prompt chunks over the paged cache (pa_multi_query_cached_kv_attention: tcgen05 kernel fed by TMA straight from
    the cache pages, long items split over their context) beside what v0.1.0 has to do without the operator --
    recompute the whole prompt.  The last query row of every sequence sees its whole context, so it must agree with
    the decode kernel on the same cache (checked on every shape)."""
    import torch
    from vllm_b200 import attention_ops
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f)["bf16_tflops"])
    except Exception:
        peak = 1590.0
    shapes = (("llama13b_chunk512_of_2048_x8", 40, 128, 16, [(512, 2048)] * 8, torch.bfloat16),
              ("llama13b_chunk2048_of_8192_x4", 40, 128, 16, [(2048, 8192)] * 4, torch.bfloat16),
              ("llama7b_fp16_ragged", 32, 128, 16, [(37, 1500), (911, 911), (128, 4096), (460, 600)], torch.float16))
    rows = []
    for name, h, d, b, pairs, dtype in shapes:
        ctx, qlens = [p[1] for p in pairs], [p[0] for p in pairs]
        cu = [0]
        for n in qlens:
            cu.append(cu[-1] + n)
        nblocks = sum((c + b - 1) // b for c in ctx)
        g = torch.Generator(device=dev).manual_seed(0)
        kc = torch.empty((nblocks, h, d // 8, b, 8), dtype=dtype, device=dev).normal_(generator=g)
        vc = torch.empty((nblocks, h, d, b), dtype=dtype, device=dev).normal_(generator=g)
        q = torch.empty((cu[-1], h, d), dtype=dtype, device=dev).normal_(generator=g)
        out = torch.empty_like(q)
        mb = max((c + b - 1) // b for c in ctx)
        bt = torch.zeros((len(pairs), mb), dtype=torch.int32)
        perm = torch.randperm(nblocks, generator=torch.Generator().manual_seed(1)).int()
        at = 0
        for i, c in enumerate(ctx):
            n = (c + b - 1) // b
            bt[i, :n] = perm[at:at + n]
            at += n
        btd = bt.to(dev)
        scale = d ** -0.5
        call = lambda: attention_ops.multi_query_cached_kv_attention(cu, out, q, kc, vc, scale, btd, ctx, b)
        ms = _time_events(call, iters=20, warm=5)
        fl = sum(4 * d * h * sum(min(c, r + 1 + c - n) for r in range(n)) for n, c in pairs)
        # the last row of each chunk against the decode kernel
        last = torch.tensor([cu[i + 1] - 1 for i in range(len(pairs))], device=dev)
        dec = torch.empty((len(pairs), h, d), dtype=dtype, device=dev)
        attention_ops.single_query_cached_kv_attention(dec, q[last].contiguous(), kc, vc, scale, btd,
                                                       torch.tensor(ctx, dtype=torch.int32, device=dev), b, max(ctx))
        torch.cuda.synchronize()
        err = float((out[last].float() - dec.float()).abs().max())
        tol = (1e-2 if dtype == torch.bfloat16 else 1e-3) * max(1.0, float(dec.float().abs().max()))
        qf = torch.empty((sum(ctx), 3, h, d), dtype=dtype, device=dev).normal_(generator=g)
        of = torch.empty((sum(ctx), h, d), dtype=dtype, device=dev)
        ms_full = _time_events(lambda: attention_ops.multi_query_kv_attention(of, qf[:, 0], qf[:, 1], qf[:, 2], ctx, scale),
                               iters=10, warm=3)
        rows.append({"shape": name, "heads": h, "head_size": d, "block_size": b,
                     "query_and_context_lens": pairs if len(set(pairs)) > 1 else [list(pairs[0]), len(pairs)],
                     "dtype": str(dtype).split(".")[-1], "ms": ms,
                     "roofline": {"bound": "tensor", "achieved": fl / ms / 1e9, "peak": peak, "unit": "TFLOP/s",
                                  "frac": fl / ms / 1e9 / peak, "traffic": None, "algorithmic_flops_per_launch": fl},
                     "recompute_whole_prompt_ms": ms_full, "speedup_vs_recompute": ms_full / ms,
                     "last_rows_vs_decode_kernel": {"max_abs_diff": err, "tolerance": tol, "ok": err <= tol}})
        del kc, vc, q, out, qf, of
        torch.cuda.empty_cache()
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workload", default="llama13b_tp", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU leg (tuning runs)")
    ap.add_argument("--no-also", action="store_true", help="skip the extra workloads, movers, prefill and all-reduce legs")
    ap.add_argument("--leg", default="", help=argparse.SUPPRESS)           # child-process legs
    ap.add_argument("--heads-local", type=int, default=0, help=argparse.SUPPRESS)
    ap.add_argument("--device-index", type=int, default=0, help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.leg == "ref-kernel":
        ref_kernel_leg(args.workload, args.heads_local or WORKLOADS[args.workload]["heads"], args.device_index)
        return
    if args.warmup < 3:
        args.warmup = 3
    wl = WORKLOADS[args.workload]

    if args.impl == "reference":
        run_reference(args, wl)
        return

    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world
    if args.gpus != world and rank == 0:
        print(f"[bench] --gpus {args.gpus} but WORLD_SIZE={world}; using {world}", file=sys.stderr)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if wl["shard"]:
        from vllm_b200 import sharding
        heads_local = sharding.heads_per_gpu(wl["heads"], n_gpus)      # raises unless it divides (vllm/config.py:54-60)
    else:
        heads_local = wl["heads"]

    from vllm_b200 import _cabi
    _cabi.lib()                                       # fails loudly if libpaged_b200.so is missing
    sampler = ClockSampler(local)
    sampler.start()

    p = make_problem(wl, heads_local, dev)
    esize = p["kc"].element_size()
    ms_k, win_k, launches, out = run_kernel_steps(p, args.steps, args.warmup, barrier)
    ms_k = max_over_ranks(ms_k)
    check = None
    if rank == 0:      # checker, outside the timed regions: rank 0's shard of the output against the CPU oracle
        try:
            check = oracle_row_check(p, out)
        except Exception as e:
            check = {"ok": None, "error": repr(e)[:300]}
    del out
    ms_e, win_e, h2d, d2h = run_e2e_steps(p, args.steps, args.warmup, barrier)
    ms_e = max_over_ranks(ms_e)
    ms_p, win_p = run_e2e_pipelined(p, args.steps, args.warmup, barrier)
    ms_p = max_over_ranks(ms_p)
    graph_ok = None
    try:
        ms_g, win_g, graph_ok = run_e2e_graph(p, args.steps, args.warmup, barrier)
        ms_g = max_over_ranks(ms_g)
    except Exception as e:
        print(f"[bench] graph e2e failed: {e!r}", file=sys.stderr)
        ms_g, win_g = max_over_ranks(float("inf")), None
    # the same kernel-only loop held for >= 2 s, so that the headline's short window is not the only figure
    n_sus = max(args.steps, min(20000, int(2.0 / (ms_k / args.steps / 1e3)) + 1))
    ms_s, win_s, _, _o = run_kernel_steps(p, n_sus, 3, barrier)
    del _o
    ms_s = max_over_ranks(ms_s)

    # read ceiling of the kernel's own access pattern (TMA bulk copies, no arithmetic) over the same K pages
    probe_gbs = None
    try:
        lib = _cabi.lib()
        sink = torch.zeros(4, dtype=torch.int32, device=dev)
        nbytes = p["kc"].numel() * esize
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(3):
            _cabi.check(lib.pa_stream_read_probe(p["kc"].data_ptr(), nbytes, 3, sink.data_ptr(), st))
        pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        pe0.record()
        for _ in range(10):
            _cabi.check(lib.pa_stream_read_probe(p["kc"].data_ptr(), nbytes, 3, sink.data_ptr(), st))
        pe1.record()
        torch.cuda.synchronize()
        probe_gbs = nbytes * 10 / (pe0.elapsed_time(pe1) / 1e3) / 1e9
    except Exception as e:   # the probe is an aid, never a reason to lose the bench line
        print(f"[bench] read probe failed: {e}", file=sys.stderr)

    total_heads = wl["heads"] if wl["shard"] else wl["heads"] * n_gpus
    seqs_job = wl["seqs"] if wl["shard"] else wl["seqs"] * n_gpus      # sequences the whole job finishes per step
    t_step = ms_k / args.steps / 1e3
    t_e2e = ms_e / args.steps / 1e3
    t_pipe = ms_p / args.steps / 1e3
    t_graph = ms_g / args.steps / 1e3
    t_sus = ms_s / n_sus / 1e3
    bytes_gpu = algorithmic_bytes(wl["seqs"], heads_local, wl["head_size"], wl["block_size"], wl["ctx"], esize)
    bytes_job = bytes_gpu * n_gpus
    peak, peak_src = measured_peak()
    achieved = bytes_gpu / t_step / 1e9
    kv_gb = 2 * p["kc"].numel() * esize / 1e9

    sampler.stop_flag.set()
    sampler.join(timeout=2)
    clocks = sampler.summary([w for w in (win_k, win_e, win_p, win_g, win_s) if w])
    del p
    torch.cuda.empty_cache()

    def guarded(fn, *a):
        try:
            return fn(*a)
        except Exception as e:        # an extra leg never costs the headline line
            print(f"[bench] {fn.__name__} failed: {e!r}", file=sys.stderr)
            return {"error": repr(e)[:300]}

    also = movers = prefill = cached_prefill = ref_kernel = allreduce = None
    extras = args.workload == "llama13b_tp" and not args.no_also
    if extras and n_gpus == 1:
        k2 = min(args.steps, 100)
        also = [guarded(decode_block, name, dev, k2, args.warmup, peak) for name in ("llama7b", "neox20b_tp8", "opt66b_tp8")]
        movers = guarded(movers_block, dev, peak)
        prefill = guarded(prefill_block, dev)
        cached_prefill = guarded(cached_prefill_block, dev)
        # config 3 at one GPU is out of the reference kernel's reach (32-bit offsets): configs[1] stands in
        ref_kernel = gpu_reference_kernel("llama7b", WORKLOADS["llama7b"]["heads"], local)
    elif extras:
        if rank == 0:
            ref_kernel = gpu_reference_kernel(args.workload, heads_local, local)
        barrier()
        try:
            from tools.allreduce_check import run_leg
            rows = run_leg(rank, world)
            allreduce = {"what": "row-parallel all-reduce after attention (parallel_state.py:506-556): GraphAllReduce"
                                 "(2560, 5120, bf16) -> pa_allreduce, vs torch.distributed.all_reduce (NCCL) on the same buffer",
                         "rows": rows} if rank == 0 else None
        except Exception as e:
            print(f"[bench] all-reduce leg failed on rank {rank}: {e!r}", file=sys.stderr)
            allreduce = {"error": repr(e)[:300]}

    cpu_baseline = None
    if rank == 0 and n_gpus == 1 and not args.no_cpu_baseline:
        nseq, one = cpu_sample_size(wl, budget_s=15.0, cap=wl["seqs"])
        sec = time_cpu_reference(wl, nseq)
        cpu_baseline = {"value": nseq / sec, "unit": "seqs/s", "cores": torch.get_num_threads(), "kind": "port",
                        "sample": f"one pass over {nseq} of {wl['seqs']} sequences (all {wl['heads']} heads, "
                                  f"{wl['ctx']}-token contexts), PyTorch loop reference restated in oracle/paged_oracle.py",
                        "seconds": sec, "host_cpus": os.cpu_count()}
        try:
            cpu_baseline["config1_gpu_seconds"] = time_gpu_config1(dev)
        except Exception as e:
            cpu_baseline["config1_gpu_error"] = repr(e)[:200]
        try:
            c1 = time_cpu_config1()
            cpu_baseline["config1_seconds"] = c1["seconds"]
            cpu_baseline["config1"] = c1
        except Exception as e:
            cpu_baseline["config1_error"] = repr(e)[:200]

    if rank == 0:
        line = {
            "metric": METRIC, "value": seqs_job / t_step, "unit": "seqs/s", "n_gpus": n_gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_step * 1e3, "higher_is_better": True,
            "scaling": "strong" if wl["shard"] else "weak", "vs_baseline": None, "dtype": wl["dtype"],
            "data": "synthetic",
            "config": dict(workload_config(args.workload, wl, n_gpus, heads_local),
                           l2=f"inputs larger than L2: {kv_gb:.2f} GB of KV pages per GPU re-read every step vs 126 MB L2"),
            "hbm_gbs": bytes_job / t_step / 1e9, "hbm_frac_of_8tbs": achieved / NOMINAL_HBM_GBS,
            # frac > 1 is possible: the peak is a read+write copy, the kernel a read-only stream; frac_of_read_probe
            # (pa_stream_read_probe: the same TMA bulk-copy pattern without arithmetic, timed in this run) and
            # hbm_frac_of_8tbs are the bounded readings.  DESIGN.md section 5 explains every key.
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic_from_profiles(args.workload, heads_local), "peak_source": peak_src,
                         "frac_of_read_probe": (achieved / probe_gbs) if probe_gbs else None,
                         "algorithmic_bytes_per_launch": bytes_gpu, "kernel": "pa::paged_decode_kernel",
                         "read_probe_gbs": probe_gbs,
                         "sustained_gbs": bytes_gpu / t_sus / 1e9, "sustained_steps": n_sus,
                         "sustained_seconds": ms_s / 1e3},
            # value = the better of the two serial loops (same contract: this step's inputs host -> device, operator,
            # output -> host, host synchronises every step); pipelined_value relaxes the contract (result of step i on
            # the host before step i+2 is issued) and is reported beside it, not as the value
            "e2e": {"value": seqs_job / min(t_e2e, t_graph), "unit": "seqs/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": min(t_e2e, t_graph) * 1e3,
                    "mode": "graph" if t_graph < t_e2e else "serial",
                    "serial_value": seqs_job / t_e2e, "graph_value": (seqs_job / t_graph) if t_graph < 1e30 else None,
                    "graph_output_ok": graph_ok, "pipelined_value": seqs_job / t_pipe,
                    "frac_of_device_rate": t_step / min(t_e2e, t_graph)},
            "gpu_launches": launches, "clocks": clocks, "check": check, "cpu_baseline": cpu_baseline,
            "gpu_reference_kernel": ref_kernel,
        }
        rf = line["roofline"]
        if check:
            rf["check_ok"], rf["check_max_abs_err"] = check.get("ok"), check.get("max_abs_err")
        if ref_kernel and "gbs" in ref_kernel:      # the reference's own CUDA kernel, recompiled for sm_100, same tensors
            rf["reference_cuda_kernel_gbs"] = ref_kernel["gbs"]
            rf["ours_same_tensors_gbs"] = ref_kernel["ours_gbs"]
            rf["speedup_vs_reference_cuda_kernel"] = ref_kernel["speedup"]
            rf["reference_cuda_kernel_workload"] = f"{ref_kernel['workload']} x {ref_kernel['heads_per_gpu']} heads"
        for blk in also or []:
            if isinstance(blk, dict) and "hbm_gbs" in blk:
                name = blk["config"]["workload"]
                rf[f"{name}_gbs"], rf[f"{name}_frac_of_8tbs"] = blk["hbm_gbs"], blk["frac_of_8tbs"]
                rf[f"{name}_check_ok"] = blk["check"].get("ok")
        rf.update(dispatch_matrix_summary())
        if isinstance(movers, dict):
            for k, v in movers.items():
                if isinstance(v, dict) and "roofline" in v:
                    rf[f"{k}_frac"] = v["roofline"]["frac"]
        for r in prefill if isinstance(prefill, list) else []:
            if isinstance(r, dict) and "speedup_vs_library" in r:
                rf[f"prefill_{r['shape']}_vs_library"] = r["speedup_vs_library"]
                rf[f"prefill_{r['shape']}_tflops"] = r["roofline"]["achieved"]
        if also is not None:
            line["also"] = also
        if movers is not None:
            line["cache_movers"] = movers
        if prefill is not None:
            line["prefill"] = prefill
        if cached_prefill is not None:
            line["cached_prefill"] = cached_prefill
        if allreduce is not None:
            line["allreduce"] = allreduce
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
