python -m pytest tests -m gpu -x -q > gpurun_out/s4_pytest.log 2>&1; tail -3 gpurun_out/s4_pytest.log
python -c "import __graft_entry__ as g; g.smoke()"
python bench.py > gpurun_out/s4_bench.json 2> gpurun_out/s4_bench.err; echo bench rc=$?
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s4_bench_ref.json 2> gpurun_out/s4_bench_ref.err; echo ref rc=$?; tail -c 600 gpurun_out/s4_bench_ref.json
