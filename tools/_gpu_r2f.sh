mkdir -p gpurun_out
python tools/capture_traffic.py > gpurun_out/plain_traffic.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:paged_decode -f -o gpurun_out/r2_decode_traffic python tools/capture_traffic.py > gpurun_out/ncu_traffic.log 2>&1
tail -2 gpurun_out/ncu_traffic.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_bench_launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
tail -2 gpurun_out/ncu_bench.log | cut -c1-300
