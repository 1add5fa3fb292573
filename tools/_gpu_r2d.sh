mkdir -p gpurun_out
timeout 900 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"; tail -3 gpurun_out/bench_n1.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/bench_n1.json').read().strip().splitlines()[-1])
print({k:(v if not isinstance(v,(dict,list)) else '...') for k,v in l.items()})
print('check', l['check']); print('e2e', l['e2e']); print('ref', l['gpu_reference_kernel'])
for a in l.get('also',[]): print('also', a.get('config',{}).get('workload'), a.get('ms_per_step'), a.get('hbm_gbs'), a.get('check',{}).get('ok'), a.get('error'))
print('movers', json.dumps(l.get('cache_movers'))[:1500])
print('prefill', json.dumps(l.get('prefill'))[:1200])
print('cpu', l['cpu_baseline'])
PY
