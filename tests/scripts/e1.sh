python -m pytest tests/test_gpu_cpp_examples.py -m gpu -q -x -k "verbatim or rotating or unknown" 2>&1 | tail -25
tests/profiler/bin/ckProfiler gemm_w4a16 0 2 0 1 1 11008 4096 0 3 30 | grep "Perf"
CKP_ROTATING_MB=512 tests/profiler/bin/ckProfiler gemm_w4a16 0 2 0 1 1 11008 4096 0 3 30 | grep "Perf"
