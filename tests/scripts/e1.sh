python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "chain or w4 or golden" 2>&1 | tail -5
export PERF_GRAPH=1
S="w4 1 4096 4096 w4 1 11008 4096 w4 1 4096 11008 w4 4 4096 4096 w4 16 4096 4096"
echo "== hardware PDL chain"; python tests/perf_probe.py $S
echo "== software dependency chain"; PERF_CHAIN=1 python tests/perf_probe.py $S
