N=$1
export NCCL_DEBUG=WARN
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/allreduce_check.py --json gpurun_out/allreduce_check_n$N.json 2>&1 | grep -v "^W1\|^\*\*\*\|OMP_NUM" | tail -45 | cut -c1-330
for M in 8 4; do
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $M --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $M --steps 200 --warmup 20 2>&1 | grep -v "^W1\|^\*\*\*\|OMP_NUM" | tail -3 > gpurun_out/bench_n$M.json; cut -c1-400 gpurun_out/bench_n$M.json
done
