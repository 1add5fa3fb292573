"""Context splits of the prompt kernels (plan_splits in csrc/prefill_attention.cu): device time per call with the
host out of the picture (20 calls captured in one CUDA graph, replayed) and host time per call (enqueue only), with
the planner on and off (PA_PREFILL_SPLIT is read once per process, so each setting runs in its own process).

    python tools/prefill_split_probe.py            # both settings, prints one JSON line per shape and setting
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CACHED = [
    ("llama13b_chunk512_of_2048_x8", 40, 128, 16, [(512, 2048)] * 8, "bfloat16"),
    ("llama13b_chunk2048_of_8192_x4", 40, 128, 16, [(2048, 8192)] * 4, "bfloat16"),
    ("llama13b_prefix1024_new256_x10", 40, 128, 16, [(256, 1280)] * 10, "bfloat16"),
    ("llama7b_fp16_ragged", 32, 128, 16, [(37, 1500), (911, 911), (128, 4096), (460, 600)], "float16"),
    ("neox20b_d96_chunk512_of_2048_x8", 64, 96, 16, [(512, 2048)] * 8, "float16"),
    ("llama13b_chunk128_of_8192_x2", 40, 128, 16, [(128, 8192)] * 2, "bfloat16"),
    # unsplit look-alikes of the parts the planner makes of the shape above
    ("probe_chunk128_of_1408_x12", 40, 128, 16, [(128, 1408)] * 12, "bfloat16"),
    ("probe_chunk128_of_8192_x4", 40, 128, 16, [(128, 8192)] * 4, "bfloat16"),
    ("probe_chunk256_of_1408_x12", 40, 128, 16, [(256, 1408)] * 12, "bfloat16"),
]
PLAIN = [
    ("llama13b_5x512", 40, 128, [512] * 5, "bfloat16"),
    ("llama13b_ragged_2560", 40, 128, [37, 911, 128, 1024, 460], "bfloat16"),
    ("llama13b_1x2560", 40, 128, [2560], "bfloat16"),
    ("llama13b_8x2048", 40, 128, [2048] * 8, "bfloat16"),
]


def graph_time(fn, calls=20, replays=10):
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(calls):
        fn()
    host_us = (time.perf_counter() - t0) / calls * 1e6
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for _ in range(calls):
                fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(replays):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (calls * replays) * 1e3, host_us


def worker():
    import torch
    from vllm_b200 import attention_ops
    only = os.environ.get("PROBE_ONLY", "")
    for name, h, d, b, pairs, dt in CACHED:
        if only and only not in name:
            continue
        dtype = getattr(torch, dt)
        ctx = [p[1] for p in pairs]
        cu = [0]
        for p in pairs:
            cu.append(cu[-1] + p[0])
        nblocks = sum((c + b - 1) // b for c in ctx)
        kc = torch.empty((nblocks, h, d // 8, b, 8), dtype=dtype, device="cuda").normal_()
        vc = torch.empty((nblocks, h, d, b), dtype=dtype, device="cuda").normal_()
        q = torch.empty((cu[-1], h, d), dtype=dtype, device="cuda").normal_()
        out = torch.empty_like(q)
        mb = max((c + b - 1) // b for c in ctx)
        bt = torch.zeros((len(pairs), mb), dtype=torch.int32)
        perm = torch.randperm(nblocks).int()
        at = 0
        for i, c in enumerate(ctx):
            n = (c + b - 1) // b
            bt[i, :n] = perm[at:at + n]
            at += n
        btd = bt.cuda()
        fl = sum(4 * d * h * sum(min(c, r + 1 + c - n) for r in range(n)) for n, c in pairs)
        gpu_us, host_us = graph_time(lambda: attention_ops.multi_query_cached_kv_attention(cu, out, q, kc, vc, d ** -0.5, btd, ctx, b))
        print(json.dumps({"shape": name, "kind": "cached", "split": os.environ.get("PA_PREFILL_SPLIT", "1"),
                          "gpu_us": round(gpu_us, 1), "host_us": round(host_us, 1), "tflops": round(fl / gpu_us / 1e6, 1)}), flush=True)
    for name, h, d, lens, dt in PLAIN:
        if only and only not in name:
            continue
        dtype = getattr(torch, dt)
        n = sum(lens)
        qkv = torch.empty((n, 3, h, d), dtype=dtype, device="cuda").normal_()
        out = torch.empty((n, h, d), dtype=dtype, device="cuda")
        fl = sum(4 * d * h * L * (L + 1) // 2 for L in lens)
        gpu_us, host_us = graph_time(lambda: attention_ops.multi_query_kv_attention(out, qkv[:, 0], qkv[:, 1], qkv[:, 2], lens, d ** -0.5))
        print(json.dumps({"shape": name, "kind": "plain", "split": os.environ.get("PA_PREFILL_SPLIT", "1"),
                          "gpu_us": round(gpu_us, 1), "host_us": round(host_us, 1), "tflops": round(fl / gpu_us / 1e6, 1)}), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "worker":
        worker()
    else:
        for setting in ("0", "1"):
            env = dict(os.environ, PA_PREFILL_SPLIT=setting)
            subprocess.run([sys.executable, os.path.abspath(__file__), "worker"], env=env, check=False)
