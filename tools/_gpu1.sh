set -x
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/s4_pytest.log
cat gpurun_out/s4_pytest.log
timeout 600 python bench.py > gpurun_out/s4_bench.json 2> gpurun_out/s4_bench.err; tail -c 3000 gpurun_out/s4_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/s4_bench_ref.json 2> gpurun_out/s4_bench_ref.err; tail -c 1500 gpurun_out/s4_bench_ref.json
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
