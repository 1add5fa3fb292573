# round 3, call A: the no-allocation entry point alone (fresh process), then the whole GPU suite, then the dispatch matrix
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r3_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r3_pytest_gpu.log
timeout 900 python tools/dispatch_matrix.py 2>&1 | tail -40 | tee gpurun_out/r3_dispatch_matrix.log
timeout 300 python tools/reshape_hint_bench.py 2>&1 | tail -8 | tee gpurun_out/r3_reshape_hint.log
