timeout 90 python -m pytest tests/test_allreduce.py -m gpu -x -q 2>&1 | tail -5
