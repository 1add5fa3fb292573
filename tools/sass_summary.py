#!/usr/bin/env python
"""Counts the Blackwell-specific SASS mnemonics per kernel of lib/libpaged_b200.so (cuobjdump -sass; no GPU needed).

    python tools/sass_summary.py > profiles/r3_sass_summary.txt

UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / .st (tensor memory), UTMALDG = TMA tensor load, UBLKCP = cp.async.bulk
(TMA engine, no tensor map), SYNCS = mbarrier, LDGMC = multimem.ld_reduce / REDG-style multicast, FHFMA = fp16/bf16 x
fp16/bf16 + fp32 fused multiply-add, CREDUX = redux.sync, HMMA = mma.sync (sibling-group tiles), ACQBULK/PREEXIT = PDL."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "vllm-v0.1.0_b200", "lib", "libpaged_b200.so")
WATCH = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "LDGMC", "MULTIMEM", "FHFMA", "CREDUX",
         "HMMA", "LDSM", "MOVM", "ACQBULK", "PREEXIT", "LDGSTS", "MUFU.EX2", "FFMA2", "FADD2"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            counts[cur]["_total"] += 1
            for w in WATCH:
                if op.startswith(w):
                    counts[cur][w] += 1
    dem = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
    fam = collections.OrderedDict()
    for name, d in zip(counts, dem):
        short = re.sub(r"\(.*", "", d)
        short = re.sub(r"^void ", "", short)
        fam.setdefault(re.sub(r"<.*", "", short), []).append((short, counts[name]))
    print(f"# {os.path.relpath(LIB, ROOT)}: SASS mnemonic counts per kernel (cuobjdump -sass, sm_100a)")
    print(f"# kernels: {len(counts)}; watched mnemonics: {' '.join(WATCH)}")
    tot = collections.Counter()
    for f, items in fam.items():
        agg = collections.Counter()
        for _, c in items:
            agg.update(c)
        tot.update(agg)
        hits = " ".join(f"{w}={agg[w]}" for w in WATCH if agg[w])
        print(f"\n{f}  [{len(items)} instantiation(s), {agg['_total']} instructions]  {hits}")
        for short, c in items[:3] if len(items) > 3 else items:
            print("    " + short[:110] + "  " + " ".join(f"{w}={c[w]}" for w in WATCH if c[w]))
        if len(items) > 3:
            print(f"    ... {len(items) - 3} more instantiations")
    print("\n# whole library: " + " ".join(f"{w}={tot[w]}" for w in WATCH if tot[w]))


if __name__ == "__main__":
    main()
