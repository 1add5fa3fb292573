#!/usr/bin/env python
"""One launch of the decode kernel per bench workload / heads-per-GPU, for `ncu --set full`
(profiles/roofline_traffic.json is parsed from the capture by tools/parse_traffic.py).
Order: llama13b_tp at 40, 20, 10, 5 heads; llama7b 32; neox20b_tp8 8; opt66b_tp8 9 -- each preceded by a warm launch."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from vllm_b200 import attention_ops  # noqa: E402

DEV = torch.device("cuda", 0)
ORDER = [("llama13b_tp", 40), ("llama13b_tp", 20), ("llama13b_tp", 10), ("llama13b_tp", 5), ("llama7b", 32),
         ("neox20b_tp8", 8), ("opt66b_tp8", 9)]


def main():
    for name, heads in ORDER:
        wl = bench.WORKLOADS[name]
        p = bench.make_problem(wl, heads, DEV)
        out = torch.empty(p["q"].shape, dtype=p["dtype"], device=DEV)
        for _ in range(2):      # launch 2k: warm, launch 2k+1: profiled
            attention_ops.single_query_cached_kv_attention(out, p["q"], p["kc"], p["vc"], p["scale"], p["tables"],
                                                           p["lens"], p["B"], p["L"])
        torch.cuda.synchronize()
        del p, out
        torch.cuda.empty_cache()
    print("launched", 2 * len(ORDER))


if __name__ == "__main__":
    main()
