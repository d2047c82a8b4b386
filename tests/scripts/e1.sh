timeout 300 python tests/slab_probe.py 2>&1 | grep -v "mbarrier wait" | tail -100
python -m pytest tests/test_gpu_baseline_sizes.py -m gpu -q -x -k "fixed_k or contract or boundary" 2>&1 | tail -5
PERF_GRAPH=1 PERF_RTN=0 python tests/perf_probe.py w4 256 4096 4096
