# M = 16 decode: where does the time go?  trace build, graph-timed; dbg 1 = copies only, 8 = math only, 9 = neither
L=$PWD/composable_kernel_b200/lib/libckb_trace.so
S="w4 16 4096 4096 w4 16 11008 4096 w4 16 4096 11008 w4 8 4096 4096 w4 4 4096 4096"
for d in 0 1 8 9; do
  echo "== gemv2 forced, dbg $d"
  CKB_GEMV=2 CKB_LIB=$L CKB_GEMV_DBG=$d PERF_GRAPH=1 timeout 120 python tests/perf_probe.py $S 2>&1 | tail -5
done
