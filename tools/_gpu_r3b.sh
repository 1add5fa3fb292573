# round 3, call B: forced launch shapes on the weak dispatch cells, ncu on the B=8 anomaly, prompt-kernel tests after the
# barrier fix, the bench line with the new keys
mkdir -p gpurun_out
timeout 600 python tools/dispatch_matrix.py --no-ref --cells fp16:8:64,bf16:8:64,fp16:8:96,fp16:8:80,fp16:32:128,fp32:32:128,fp32:32:64 \
  --tunings 2.4.2,2.4.3,2.4.4,2.2.4,2.2.8,2.8.2,2.1.8,2.8.1,2.3.2,2.6.2,2.2.2,2.3.3,1.4.1,1.8.1,3.4.2 --out gpurun_out/dispatch_tune.json 2>&1 | tail -12 | tee gpurun_out/r3_dispatch_tune.log
timeout 900 python -m pytest tests/test_prefill.py tests/test_cached_prefill.py tests/test_prefill_soak.py tests/test_gpu_graphs.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r3_prefill_tests.log
timeout 900 python bench.py > gpurun_out/r3_bench_n1.json 2> gpurun_out/r3_bench_n1.err; tail -c 600 gpurun_out/r3_bench_n1.err; python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r3_bench_n1.json").read().strip().splitlines()[-1])
    print(json.dumps({k: d[k] for k in ("value", "ms_per_step", "roofline", "e2e", "cpu_baseline", "gpu_launches")})[:6000])
    print(len(json.dumps(d)), "bytes in the line")
except Exception as e:
    print("bench line unreadable:", e)
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_decode -c 8 -f -o gpurun_out/r3_b8_cells python tools/dispatch_ncu.py fp16:8:64 fp16:8:80 fp16:8:96 fp32:32:128 > gpurun_out/r3_ncu.log 2>&1; tail -3 gpurun_out/r3_ncu.log
