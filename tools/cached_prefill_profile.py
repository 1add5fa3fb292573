"""A few launches of the paged prompt kernel for `ncu` (profiles/): LLaMA-13B bf16, 4 sequences,
2048 new tokens behind 6144 cached ones.   python tools/cached_prefill_profile.py [launches]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_b200 import attention_ops  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
h, d, b = 40, 128, 16
pairs = [(2048, 8192)] * 4
ctx = [p[1] for p in pairs]
cu = [0]
for p in pairs:
    cu.append(cu[-1] + p[0])
nblocks = sum(c // b for c in ctx)
kc = torch.empty((nblocks, h, d // 8, b, 8), dtype=torch.bfloat16, device="cuda").normal_()
vc = torch.empty((nblocks, h, d, b), dtype=torch.bfloat16, device="cuda").normal_()
q = torch.empty((cu[-1], h, d), dtype=torch.bfloat16, device="cuda").normal_()
out = torch.empty_like(q)
bt = torch.randperm(nblocks).int().view(len(pairs), -1).cuda()
for _ in range(reps):
    attention_ops.multi_query_cached_kv_attention(cu, out, q, kc, vc, d ** -0.5, bt, ctx, b)
torch.cuda.synchronize()
print("launched", reps, float(out.float().abs().mean()))
