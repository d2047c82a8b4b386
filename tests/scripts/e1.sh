python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py tests/test_gpu_guards.py -m gpu -q -x -k "w4 or golden or cfg2 or contract or boundary or fixed_k or window" 2>&1 | tail -3
export PERF_GRAPH=1
S="w4 1 4096 4096 w4 1 11008 4096 w4 1 4096 11008 w4 4 4096 4096 w4 16 4096 4096"
echo "== current"; python tests/perf_probe.py $S
echo "== c3d3a8e"; CKB_LIB=composable_kernel_b200/lib/variants/libckb_c3d3a8e.so python tests/perf_probe.py $S
echo "== current, 2 of 3 independent (EXT kernel)"; PERF_NODEP=3 python tests/perf_probe.py $S
