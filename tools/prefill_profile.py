"""A few launches of the tcgen05 prompt-attention kernel for `ncu` (profiles/): LLaMA-13B bf16, 8 prompts.
    python tools/prefill_profile.py [prompt_len] [launches]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_b200 import attention_ops  # noqa: E402

L = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
h, d, lens = 40, 128, [L] * 8
n = sum(lens)
qkv = torch.empty((n, 3, h, d), dtype=torch.bfloat16, device="cuda").normal_()
out = torch.empty((n, h, d), dtype=torch.bfloat16, device="cuda")
for _ in range(reps):
    attention_ops.multi_query_kv_attention(out, qkv[:, 0], qkv[:, 1], qkv[:, 2], lens, d ** -0.5)
torch.cuda.synchronize()
print("launched", reps, float(out.float().abs().mean()))
