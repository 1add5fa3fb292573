# round 3, call C: compute-sanitizer over the GPU tests of the hot path; summaries -> gpurun_out/r3_sanitizer_*.log
# NOTE: compute-sanitizer is closed on this pool (exit 86 before the target starts); kept as the record of what was asked for.
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
run() {  # tool, time limit, label, pytest args...
  tool=$1; lim=$2; label=$3; shift 3
  log=gpurun_out/r3_sanitizer_${tool}_${label}.log
  timeout $lim $CS --tool $tool --target-processes all --print-limit 5 python -m pytest "$@" -m gpu -q -x -p no:cacheprovider > $log 2>&1
  rc=$?
  echo "== $tool $label rc=$rc: $(grep -c 'ERROR SUMMARY' $log) summaries; $(grep 'ERROR SUMMARY' $log | sort | uniq -c | tr '\n' ';') pytest: $(grep -E 'passed|failed|error' $log | tail -1)" | tee -a gpurun_out/r3_sanitizer_summary.txt
}
: > gpurun_out/r3_sanitizer_summary.txt
run memcheck 500 decode tests/test_gpu_decode.py -k "golden or launch_shapes or reference_matrix"
run memcheck 400 movers tests/test_gpu_cache.py tests/test_rope.py tests/test_zz_no_alloc.py tests/test_gpu_sibling_groups.py
run memcheck 400 prefill tests/test_prefill.py tests/test_cached_prefill.py -k golden
run synccheck 400 decode tests/test_gpu_decode.py -k "golden or launch_shapes" tests/test_gpu_sibling_groups.py tests/test_zz_no_alloc.py
run synccheck 400 prefill tests/test_prefill.py tests/test_cached_prefill.py -k golden
run initcheck 400 decode tests/test_gpu_decode.py -k "golden" tests/test_gpu_cache.py tests/test_zz_no_alloc.py
run racecheck 400 decode_movers tests/test_gpu_decode.py -k "golden" tests/test_gpu_cache.py tests/test_gpu_sibling_groups.py
cat gpurun_out/r3_sanitizer_summary.txt
