timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "w4a16 or host_buffer or golden" 2>&1 | tail -3
PERF_GRAPH=1 timeout 120 python tests/perf_probe.py w4 1 4096 4096 w4 4 4096 4096 w4 16 4096 4096 w4 1 11008 4096 w4 16 11008 4096 w4 1 4096 11008 w4 16 4096 11008 w4 1 28672 8192 w4 16 28672 8192 2>&1 | tail -9
python bench.py --no-cpu-baseline > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; tail -3 gpurun_out/bench_graph.err
