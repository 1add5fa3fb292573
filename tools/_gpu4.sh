mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_cached_prefill.py -m gpu -x -q -k "golden and bf16-b16-d128" 2>&1 | tail -5
timeout 600 python -m pytest tests/test_cached_prefill.py tests/test_prefill.py tests/test_gpu_layer.py -m gpu -q 2>&1 | grep -E "elements off|passed|failed|FAILED|Error" | head -20 | cut -c1-250
timeout 300 python tools/cached_prefill_bench.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r['shape'], 'ms', r['ms'], 'tflops', r['tflops'], 'recompute_ms', r['recompute_whole_prompt_ms'], 'speedup', r['speedup_vs_recompute'])
"
