timeout 40 python -m pytest tests/test_prefill.py -m gpu -x -q -k tiles_raw 2>&1 | tail -15 || exit 1
timeout 120 python -m pytest tests/test_prefill.py tests/test_gpu_layer.py -m gpu -q 2>&1 | tail -25
timeout 120 python tools/prefill_bench.py 2>&1 | tail -20 | cut -c1-330
