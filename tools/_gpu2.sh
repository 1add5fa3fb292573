mkdir -p gpurun_out
for L in 3000 2048 4096 257 129; do timeout 40 python tools/prefill_profile.py $L 30 2>&1 | tail -1; echo "rc=${PIPESTATUS[0]} L=$L"; done
timeout 200 python -m pytest tests/test_prefill.py tests/test_gpu_layer.py -m gpu -x -q 2>&1 | tail -2
timeout 150 python tools/prefill_bench.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r['shape'], r['ours_ms'], r['ours_tflops'], 'fa2', r.get('flash_attn2_tflops'), 'sdpa', r.get('torch_sdpa_tflops'), 'diff', r.get('max_abs_diff_vs_flash_attn2'))
"
