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
h, d, lens = int(os.environ.get("PF_HEADS", "40")), int(os.environ.get("PF_HEAD_SIZE", "128")), [L] * nprompts
dd = d
n = sum(lens)
qkv = torch.empty((n, 3, h, d), dtype=torch.bfloat16, device="cuda").normal_()
out = torch.empty((n, h, d), dtype=torch.bfloat16, device="cuda")
dbg = torch.zeros(65536, dtype=torch.float32, device="cuda")
for _ in range(3):
    attention_ops.multi_query_kv_attention(out, qkv[:, 0], qkv[:, 1], qkv[:, 2], lens, d ** -0.5, debug=dbg)
torch.cuda.synchronize()
tl = dbg.cpu().numpy().view(np.uint32)[40000:40000 + 4 * 40 * 8].reshape(4, 40, 8).astype(np.int64)
t0 = tl[2, 0, 6]
rel = lambda x: int((x - t0) & 0xffffffff) if x else -1
d = lambda a, b: int((a - b) & 0xffffffff)
nj = L // 128                     # tiles of CTA 0's first item (the longest tile of head 0); its later items follow
print("stamps follow CTA 0 through its work items; index = running tile count (A / B) or running K/V tile (issuer)")
print("group: 0 loop top, 1 S ready, 2 row max known, 3 turn granted, 7 first half of P arrived, 4 exp done, 5 P arrived, 6 O ready (last tile of an item), out = output written")
print("issuer: 6 item's first QK issued, 4 V ready, 0/2 first half of P_A/P_B seen, 1/3 PV+QK of A/B issued, 5 next item's first QK issued early")
lo, hi = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (max(nj - 4, 0), min(nj + 4, 40))
for k in range(lo, hi):
    for g in (0, 1):
        a = tl[g, k]
        print(f"k={k:2d} tile {'AB'[g]}: top {rel(a[0]):7d} S {rel(a[1]):7d} max {rel(a[2]):7d} turn {rel(a[3]):7d} P/2 {rel(a[7]):7d} exp {rel(a[4]):7d} P {rel(a[5]):7d} Oready {rel(a[6]):7d} out {rel(tl[3, k, g]):7d}")
    i = tl[2, k]
    print(f"      issuer g={k:2d}: item top {rel(i[6]):7d} V {rel(i[4]):7d} P_A/2 seen {rel(i[0]):7d} A issued {rel(i[1]):7d} P_B/2 seen {rel(i[2]):7d} B issued {rel(i[3]):7d} early next {rel(i[5]):7d}")
js = range(4, min(nj - 3, 18))
if len(js) > 2:
    avj = lambda f: float(np.mean([f(j) for j in js]))
    for g in (0, 1):
        t = tl[g]
        print(f"group {'AB'[g]} (inside item 0): period {avj(lambda j: d(t[j + 1, 1], t[j, 1])):.0f} clk; S wait {avj(lambda j: d(t[j, 1], t[j, 0])):.0f}, ld+max(+rescale) {avj(lambda j: d(t[j, 2], t[j, 1])):.0f},"
              f" turn wait {avj(lambda j: d(t[j, 3], t[j, 2])):.0f}, exp first half {avj(lambda j: d(t[j, 7], t[j, 3])):.0f}, second half {avj(lambda j: d(t[j, 4], t[j, 7])):.0f},"
              f" wait_st+arrive {avj(lambda j: d(t[j, 5], t[j, 4])):.0f}; P arrived -> next S ready {avj(lambda j: d(t[j + 1, 1], t[j, 5])):.0f}")
per = [d(tl[0, k + 1, 1], tl[0, k, 1]) for k in range(0, 38) if tl[0, k + 1, 1] and tl[0, k, 1]]
print("group A: S ready -> next S ready, per running tile:", per)
