python -m pytest tests -m gpu -x -q 2>&1 | tail -4
tests/profiler/bin/ckProfiler gemm_multi_abd 0 1 1 0 2 0 1 16 11008 4096 -1 -1 -1 | grep "^Perf\|^Best\|found"
tests/profiler/bin/ckProfiler fpAintB_gemm 0 2 0 1 16 11008 4096 | grep "^Perf"
tests/profiler/bin/ckProfiler gemm_universal_reduce 0 0 0 2 0 1 16 4096 4096 -1 -1 -1 4 | grep "^Perf"
