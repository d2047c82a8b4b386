V=composable_kernel_b200/lib/variants/libckb_$1.so
S="w4 1 4096 4096 w4 1 11008 4096 w4 1 4096 11008 w4 4 4096 4096 w4 4 11008 4096 w4 16 4096 4096"
export PERF_GRAPH=1
for r in 1 2 3; do
echo "== release, run $r"; python tests/perf_probe.py $S
echo "== $1, run $r"; CKB_LIB=$V python tests/perf_probe.py $S
done
echo "== parity suites on $1"
CKB_LIB=$V python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py tests/test_gpu_guards.py -m gpu -q -p no:cacheprovider 2>&1 | tail -5
