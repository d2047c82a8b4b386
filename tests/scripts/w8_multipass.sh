# multi-pass W8 decode schedule (N = 14336, 28672): parity, then HMMA N-split form (default now) against the stream-K form
timeout 900 python -m pytest tests -x -q -m gpu -k "w8 or W8 or multiply or cfg3 or abd or passthrough or fpaintb or no_alloc or static" --timeout 600 2>&1 | tail -3
S="w8 1 14336 4096 w8 4 14336 4096 w8 16 14336 4096 w8 1 28672 8192 w8 16 28672 8192 w8 32 14336 4096"
echo "== default"; PERF_GRAPH=1 timeout 200 python tests/perf_probe.py $S 2>&1 | tail -6
echo "== stream-K"; CKB_W8_DECODE=streamk PERF_GRAPH=1 timeout 200 python tests/perf_probe.py $S 2>&1 | tail -6
