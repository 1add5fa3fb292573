#!/usr/bin/env python
"""Short and ragged contexts (chat-like decode batches): ours vs the reference kernel, GPU-only time."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from tools.sweep import time_gpu_only
from tools.compare_reference_kernel import load
from vllm_b200 import _cabi, attention_ops
DEV = torch.device("cuda", 0)

def main():
    try:
        ref = load("ref_attention_ops")
    except Exception:
        ref = None
    rows = []
    for name, S, H, L, ragged in (("L64", 256, 32, 64, False), ("L128", 256, 32, 128, False), ("L256", 256, 32, 256, False),
                                  ("L512", 256, 32, 512, False), ("ragged_1_1024", 256, 32, 1024, True),
                                  ("ragged_1_4096", 256, 32, 4096, True), ("S64_ragged_1_2048", 64, 32, 2048, True),
                                  ("S16_L512", 16, 32, 512, False)):
        wl = dict(heads=H, head_size=128, block_size=16, seqs=S, ctx=L, dtype="fp16")
        ncopies = 3                       # rotate over independent copies: >= 2x L2 in total, no dirty-line flush needed
        ps = [bench.make_problem(wl, H, DEV, seed=i) for i in range(ncopies)]
        p = ps[0]
        if ragged:
            g = torch.Generator().manual_seed(1)
            lens = torch.randint(1, L + 1, (S,), generator=g, dtype=torch.int32)
            for q in ps:
                q["lens"] = lens.to(DEV)
        lens = p["lens"].cpu().tolist()
        nbytes = sum(H * (2 * c * 128 * 2 + 2 * 128 * 2) + ((c + 15) // 16) * 4 + 4 for c in lens)
        out = torch.empty(p["q"].shape, dtype=p["dtype"], device=DEV)
        state = {"i": 0}
        def timed(fn):
            def call():
                state["i"] += 1
                fn(ps[state["i"] % ncopies])
            return time_gpu_only(call, iters=30)
        row = dict(shape=name, seqs=S, heads=H, mbytes=round(nbytes / 1e6, 1))
        for label, tun in (("auto", None), ("v2_w4_s2_x1", _cabi.DecodeTuning(2, 4, 2, 1)), ("v2_w2_s2_x1", _cabi.DecodeTuning(2, 2, 2, 1)),
                           ("v2_w1_s4_x1", _cabi.DecodeTuning(2, 1, 4, 1)), ("v2_w1_s2_x1", _cabi.DecodeTuning(2, 1, 2, 1)),
                           ("v2_w2_s4_x1", _cabi.DecodeTuning(2, 2, 4, 1)), ("v1_w4", _cabi.DecodeTuning(1, 4, 0, 1))):
            ms = timed(lambda p: attention_ops.single_query_cached_kv_attention(out, p["q"], p["kc"], p["vc"], p["scale"],
                                                                               p["tables"], p["lens"], 16, L, tuning=tun))
            row[label + "_us"] = round(ms * 1e3, 1)
            row[label + "_gbs"] = round(nbytes / ms / 1e6)
        if ref is not None and p["kc"].numel() < 2 ** 31:      # the reference's int offsets overflow beyond (H1)
            ms = timed(lambda p: ref.single_query_cached_kv_attention(out, p["q"], p["kc"], p["vc"], p["scale"], p["tables"],
                                                                      p["lens"], 16, L))
            row["reference_us"] = round(ms * 1e3, 1)
            row["reference_gbs"] = round(nbytes / ms / 1e6)
        rows.append(row)
        print(json.dumps(row), flush=True)
        del p, ps
        torch.cuda.empty_cache()
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "short_ctx.json"), "w"), indent=1)

if __name__ == "__main__":
    main()
