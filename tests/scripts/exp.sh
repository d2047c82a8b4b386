S="w8 4096 14336 4096 w4 4096 11008 4096 w4 4096 4096 11008 w4 4096 28672 8192 w4 4096 4096 4096"
V=$PWD/composable_kernel_b200/lib/variants
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k prefill --timeout 100 2>&1 | tail -1
echo "== s4 tg4"; timeout 200 python tests/perf_probe.py $S
echo "== s5 tg4"; CKB_LIB=$V/libckb_tg4s5.so timeout 200 python tests/perf_probe.py $S
echo "== s4 tg2"; CKB_LIB=$V/libckb_tg2.so timeout 200 python tests/perf_probe.py $S
