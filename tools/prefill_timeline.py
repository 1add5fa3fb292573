#!/usr/bin/env python
"""SM-clock timeline of one CTA of the tcgen05 prompt kernel (the longest tile of head 0): where an iteration goes.
    python tools/prefill_timeline.py [prompt_len] [num_prompts]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_b200 import attention_ops  # noqa: E402

L = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nprompts = int(sys.argv[2]) if len(sys.argv) > 2 else 8
h, d, lens = 40, 128, [L] * nprompts
dd = d
n = sum(lens)
qkv = torch.empty((n, 3, h, d), dtype=torch.bfloat16, device="cuda").normal_()
out = torch.empty((n, h, d), dtype=torch.bfloat16, device="cuda")
dbg = torch.zeros(65536, dtype=torch.float32, device="cuda")
for _ in range(3):
    attention_ops.multi_query_kv_attention(out, qkv[:, 0], qkv[:, 1], qkv[:, 2], lens, d ** -0.5, debug=dbg)
torch.cuda.synchronize()
tl = dbg.cpu().numpy().view(np.uint32)[40000:40000 + 3 * 40 * 8].reshape(3, 40, 8).astype(np.int64)
t0 = tl[2, 0, 6]
rel = lambda x: int((x - t0) & 0xffffffff)
d = lambda a, b: int((a - b) & 0xffffffff)
print("softmax group stamps per tile j: 0 loop top, 1 S ready, 2 row max known, 3 turn granted, 7 first half of P arrived, 4 exp done, 5 P arrived")
print("issuer stamps per j: 4 V ready, 0/2 first half of P_A/P_B seen, 1/3 PV+QK of A/B issued")
for j in range(4, 10):
    for g in (0, 1):
        a = tl[g, j]
        print(f"j={j:2d} tile {'AB'[g]}: top {rel(a[0]):7d} S {rel(a[1]):7d} max {rel(a[2]):7d} turn {rel(a[3]):7d} P/2 {rel(a[7]):7d} exp {rel(a[4]):7d} P {rel(a[5]):7d}")
    i = tl[2, j]
    print(f"      issuer: V {rel(i[4]):7d} P_A/2 seen {rel(i[0]):7d} A issued {rel(i[1]):7d} P_B/2 seen {rel(i[2]):7d} B issued {rel(i[3]):7d}")
js = range(4, 18)
avj = lambda f: float(np.mean([f(j) for j in js]))
for g in (0, 1):
    t = tl[g]
    print(f"group {'AB'[g]}: period {avj(lambda j: d(t[j + 1, 1], t[j, 1])):.0f} clk; S wait {avj(lambda j: d(t[j, 1], t[j, 0])):.0f}, ld+max(+rescale) {avj(lambda j: d(t[j, 2], t[j, 1])):.0f},"
          f" turn wait {avj(lambda j: d(t[j, 3], t[j, 2])):.0f}, exp first half {avj(lambda j: d(t[j, 7], t[j, 3])):.0f}, second half {avj(lambda j: d(t[j, 4], t[j, 7])):.0f},"
          f" wait_st+arrive {avj(lambda j: d(t[j, 5], t[j, 4])):.0f}; P arrived -> next S ready {avj(lambda j: d(t[j + 1, 1], t[j, 5])):.0f}")
t = tl[2]
print(f"issuer: P_A/2 seen -> A issued {avj(lambda j: d(t[j, 1], t[j, 0])):.0f}, P_B/2 seen -> B issued {avj(lambda j: d(t[j, 3], t[j, 2])):.0f},"
      f" A issued -> P_B/2 seen {avj(lambda j: d(t[j, 2], t[j, 1])):.0f}, B issued -> next P_A/2 seen {avj(lambda j: d(t[j + 1, 0], t[j, 3])):.0f};"
      f" period {avj(lambda j: d(t[j + 1, 0], t[j, 0])):.0f}")
nj = L // 128
entry, done = tl[2, 38, 7], tl[2, 39, 7]
first_issue, last_issue = tl[2, 0, 6], tl[2, min(nj, 40) - 1, 3]
per = [d(tl[2, j + 1, 0], tl[2, j, 0]) for j in range(0, min(nj, 40) - 1)]
print("issuer period per j:", per)
print(f"group A/B epilogue: loop end -> O ready {d(tl[0, 39, 1], tl[0, 39, 0])} / {d(tl[1, 39, 1], tl[1, 39, 0])}; O ready -> all done {d(done, tl[0, 39, 1])} / {d(done, tl[1, 39, 1])}; first S ready after first MMA {d(tl[0, 0, 1], first_issue)}")
print(f"CTA (head 0, tile 0, {nj} key tiles): entry -> Q, K_0 landed (first MMA) {d(first_issue, entry)} clk; first MMA -> last issue of the loop {d(last_issue, first_issue)};"
      f" last issue -> all warps done (last P.V, O / l, stores) {d(done, last_issue)}; whole CTA {d(done, entry)}")
import subprocess, threading, time
samples = []
def poll():
    while not stop[0]:
        o = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader", "-i", "0"],
                           capture_output=True, text=True).stdout.strip()
        samples.append(o)
        time.sleep(0.05)
stop = [False]
th = threading.Thread(target=poll); th.start()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(600):
    attention_ops.multi_query_kv_attention(out, qkv[:, 0], qkv[:, 1], qkv[:, 2], lens, dd ** -0.5)
e1.record(); torch.cuda.synchronize(); stop[0] = True; th.join()
print(f"600 launches: {e0.elapsed_time(e1) / 600:.4f} ms each; nvidia-smi under load (sm MHz, W, reasons): {samples[len(samples) // 2]} | {samples[-2] if len(samples) > 1 else ''}")
