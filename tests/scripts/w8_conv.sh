# cheaper int8 x scale conversion (one FFMA per code): parity of every path that uses it, then timings
# (compare with profiles/r2_bench_n1.json cfg3 and profiles/r2_w8_decode_hmma_vs_streamk.txt)
timeout 900 python -m pytest tests -x -q -m gpu -k "w8 or W8 or multiply or grouped or client or cfg3 or i8 or abd or multi or passthrough" --timeout 600 2>&1 | tail -3
PERF_GRAPH=1 timeout 200 python tests/perf_probe.py w8 1 4096 4096 w8 16 4096 4096 w8 1 11008 4096 w8 16 11008 4096 w8 16 4096 11008 2>&1 | tail -5
timeout 200 python tests/perf_probe.py w8 4096 14336 4096 w8 4096 4096 4096 2>&1 | tail -2
P=tests/profiler/bin/ckProfiler
M16=48,48,48,48,48,48,48,48,48,48,48,48,48,48,48,48
for lay in 0 1; do $P grouped_gemm_fixed_nk 0 $lay 1 2 0 1 $M16 1024 512 -1 -1 -1 1 5 200 2>&1 | grep "Best Perf\|rror" | cut -c1-62; done
