timeout 300 python -m pytest tests/test_prefill.py -m gpu -x -q -k "tiles_raw" 2>&1 | tail -30
timeout 600 python -m pytest tests/test_prefill.py -m gpu -q 2>&1 | tail -30
