# round 3, call C1: whole GPU suite on the new plan rules, the dispatch matrix again, the changed cells at a longer context
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r3_pytest_gpu_c.log
timeout 900 python tools/dispatch_matrix.py --out gpurun_out/dispatch_matrix_after.json 2>&1 | tail -3 | cut -c1-1500 | tee gpurun_out/r3_dispatch_matrix_after.log
timeout 600 python tools/dispatch_matrix.py --no-ref --seqs 64 --ctx 4096 --cells fp16:8:64,fp16:8:80,fp16:8:96,fp16:32:128,fp32:32:64,fp32:32:128,bf16:8:64 \
  --tunings 2.4.2,2.4.3,2.4.4,2.2.2,2.3.2 --out gpurun_out/dispatch_tune_L4096.json 2>&1 | tail -9 | cut -c1-700 | tee gpurun_out/r3_dispatch_tune_L4096.log
