mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_prefill.py tests/test_cached_prefill.py tests/test_gpu_layer.py -m gpu -x -q 2>&1 | tail -5
python tools/prefill_timeline.py 4096 8 2>&1 | tail -4 | cut -c1-400
timeout 300 python tools/prefill_bench.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r['shape'], r['ours_ms'], r['ours_tflops'], 'sdpa', r.get('torch_sdpa_tflops'), 'diff', r.get('max_abs_diff_vs_flash_attn2'))
"
