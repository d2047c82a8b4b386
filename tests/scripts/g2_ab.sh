# A/B of a wq_gemv2 variant library against the release one: graph-replay device time on the bench shapes, then the parity suites on the variant.
#   tests/scripts/g2_ab.sh red4
V=composable_kernel_b200/lib/variants/libckb_$1.so
S="w4 1 4096 4096 w4 1 11008 4096 w4 1 4096 11008 w4 4 4096 4096 w4 4 11008 4096 w4 4 4096 11008 w4 16 4096 4096 w4 16 11008 4096 w4 16 4096 11008 w4 1 28672 8192 w8 1 4096 4096 w8 16 11008 4096"
export PERF_GRAPH=1
for r in 1 2; do
echo "== release, run $r"; python tests/perf_probe.py $S
echo "== $1, run $r"; CKB_LIB=$V python tests/perf_probe.py $S
done
echo "== parity suites on $1"
CKB_LIB=$V python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py tests/test_gpu_guards.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -15
