# Device time (CUDA-graph replay, programmatic dependent launch) of the W4A16 decode kernels on the bench shapes.
S="w4 1 4096 4096 w4 1 11008 4096 w4 1 4096 11008 w4 1 28672 8192 w4 4 4096 4096 w4 4 11008 4096 w4 4 4096 11008 w4 16 4096 4096 w4 16 11008 4096 w4 16 4096 11008 w4 16 28672 8192"
export PERF_GRAPH=1
echo "== default dispatch"; python tests/perf_probe.py $S
echo "== rotating-producer kernel forced (CKB_GEMV=2)"; CKB_GEMV=2 python tests/perf_probe.py $S
echo "== single-producer kernel forced (CKB_GEMV=1)"; CKB_GEMV=1 python tests/perf_probe.py $S
