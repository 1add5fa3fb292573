mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_decode.py tests/test_gpu_sibling_groups.py tests/test_gpu_full_size.py -m gpu -x -q 2>&1 | tail -5
echo "--- PDL on (default), L2 hints"; python tools/wave_probe.py 5 0 2>&1 | grep -E '"seqs": (88|89|177|256|266|267|888|1065),'
echo "--- PDL off"; PA_DECODE_PDL=0 python tools/wave_probe.py 5 0 2>&1 | grep -E '"seqs": (88|89|177|256|266|267|888|1065),'
echo "--- short ctx"; python tools/short_ctx.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(r['shape'], 'auto', r['auto_us'], r['auto_gbs'], 'w2', r['v2_w2_s2_x1_us'], 'ref', r.get('reference_us'))
"
echo "--- bench"; timeout 600 python bench.py --no-cpu-baseline 2>&1 | tail -1 | cut -c1-900
