#!/usr/bin/env python
"""HBM read ceiling of the decode kernel's access pattern (pa_stream_read_probe) vs a torch copy."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_b200 import _cabi
from tools.sweep import time_gpu_only
lib = _cabi.lib()
x = torch.empty(10 * 1024 ** 3, dtype=torch.uint8, device="cuda").random_(0, 255)
sink = torch.zeros(4, dtype=torch.int32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
rows = []
for cps in (1, 2, 3, 4, 6):
    ms = time_gpu_only(lambda: _cabi.check(lib.pa_stream_read_probe(x.data_ptr(), x.numel(), cps, sink.data_ptr(), s)), iters=20)
    rows.append(dict(probe="tma_bulk_read", ctas_per_sm=cps, ms=round(ms, 4), gbs=round(x.numel() / ms / 1e6, 1)))
    print(json.dumps(rows[-1]), flush=True)
y = torch.empty(5 * 1024 ** 3, dtype=torch.uint8, device="cuda")
ms = time_gpu_only(lambda: y.copy_(x[:y.numel()]), iters=20)
rows.append(dict(probe="torch_copy_read_plus_write", ms=round(ms, 4), gbs=round(2 * y.numel() / ms / 1e6, 1)))
print(json.dumps(rows[-1]))
json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "read_ceiling.json"), "w"), indent=1)
