mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_step_inputs.py tests/test_cabi.py -m gpu -x -q 2>&1 | tail -15
timeout 300 compute-sanitizer --tool memcheck python -m pytest tests/test_step_inputs.py -m gpu -x -q -k "edge or trace" 2>&1 | tail -6
timeout 300 python tools/step_inputs_bench.py --json gpurun_out/step_inputs_bench.json 2>&1 | tail -5
