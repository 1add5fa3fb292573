# prompt benches, twice (box noise), bounded
for i in 1 2; do
timeout 120 python tools/prefill_bench.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: continue
    print(r['shape'], r['ours_ms'], r['ours_tflops'], 'sdpa', r.get('torch_sdpa_tflops'))
"
done
timeout 200 python tools/cached_prefill_bench.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: continue
    print(r['shape'], r['ms'], r['tflops'], r['speedup_vs_recompute'])
"
