# prompt attention (plain and over the paged cache): parity, then timing of both
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cached_prefill.py tests/test_prefill.py tests/test_gpu_layer.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/pytest_prefill.log
timeout 200 python tools/prefill_bench.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r['shape'], r['ours_ms'], r['ours_tflops'], 'fa2', r.get('flash_attn2_tflops'))
"
timeout 300 python tools/cached_prefill_bench.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r['shape'], 'ms', r['ms'], 'tflops', r['tflops'], 'recompute_ms', r['recompute_whole_prompt_ms'], 'speedup', r['speedup_vs_recompute'])
"
