timeout 60 python -m pytest tests/test_prefill.py -m gpu -x -q -k tiles_raw 2>&1 | tail -15
timeout 120 python -m pytest tests/test_prefill.py -m gpu -x -q 2>&1 | tail -15
timeout 120 python tools/prefill_bench.py 2>&1 | tail -20
