timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
python bench.py --no-cpu-baseline > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; tail -3 gpurun_out/bench_graph.err
