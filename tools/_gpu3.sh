# cached-prefix prompt attention: parity + regression of the plain prompt kernel + its timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cached_prefill.py tests/test_prefill.py tests/test_gpu_layer.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_cached_prefill.log
timeout 200 python tools/prefill_bench.py 2>&1 | tee gpurun_out/prefill_bench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r['shape'], r['ours_ms'], r['ours_tflops'], 'fa2', r.get('flash_attn2_tflops'))
"
