timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "w4a16 or host_buffer or golden" 2>&1 | tail -3
for d in 0 1; do echo "== DBG $d"; CKB_GEMV_DBG=$d timeout 120 python tests/perf_probe.py w4 1 4096 4096 w4 16 4096 4096 w4 1 11008 4096 w4 16 11008 4096 w4 4 4096 11008 w4 1 28672 8192 w4 16 28672 8192 2>&1 | tail -7; done
CKB_LIB=composable_kernel_b200/lib/libckb_trace.so timeout 120 python tests/gtrace_probe.py 1 4096 4096 1 11008 4096 2>&1 | tail -24
