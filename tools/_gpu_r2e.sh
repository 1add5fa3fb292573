mkdir -p gpurun_out
nvidia-smi -L | head -3
timeout 900 python -m pytest tests/test_allreduce_multigpu.py -m gpu -x -q 2>&1 | tail -5
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "rc=$?"; tail -5 gpurun_out/bench_n2.err
python - <<'PY'
import json
l=json.loads([x for x in open('gpurun_out/bench_n2.json').read().strip().splitlines() if x.startswith('{')][-1])
print({k:(v if not isinstance(v,(dict,list)) else '...') for k,v in l.items()})
print('check', l['check']); print('e2e', l['e2e']); print('ref', l['gpu_reference_kernel']); print('roofline', l['roofline'])
for r in (l.get('allreduce') or {}).get('rows', []): print('ar', r)
print((l.get('allreduce') or {}).get('error'))
PY
