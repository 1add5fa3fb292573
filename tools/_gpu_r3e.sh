# round 3, final check of the tree: GPU tests, smoke, both bench arms, ncu launch list of the bench command
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r3_final_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2 | tee gpurun_out/r3_final_smoke.log
timeout 900 python bench.py 2> gpurun_out/r3_final_bench.err | tail -1 > gpurun_out/r3_final_bench_n1.json; cut -c1-300 gpurun_out/r3_final_bench_n1.json; echo
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 | tee gpurun_out/r3_final_bench_ref_n1.json | cut -c1-400
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r3_bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r3_ncu_bench.log 2>&1; wc -l gpurun_out/r3_bench_launches.csv
