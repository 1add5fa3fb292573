# prompt kernels with / without programmatic dependent launch (experiment helper)
timeout 600 python -m pytest tests/test_cached_prefill.py tests/test_prefill.py tests/test_gpu_layer.py -m gpu -x -q 2>&1 | tail -3
for PDL in 0 1; do
  echo PDL=$PDL
  PA_PREFILL_PDL=$PDL timeout 200 python tools/prefill_bench.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: continue
    print(r['shape'], r['ours_ms'], r['ours_tflops'], 'sdpa', r.get('torch_sdpa_tflops'))
"
  PA_PREFILL_PDL=$PDL timeout 300 python tools/cached_prefill_bench.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: continue
    print(r['shape'], r['ms'], r['tflops'], r['speedup_vs_recompute'])
"
done
