# round-1 evidence: bench line, ncu launch list, ncu full capture of the decode kernel (run under gpurun)
set -x
python bench.py > gpurun_out/r1_bench_n1.json 2> gpurun_out/r1_bench_n1.err
python bench.py --steps 2 --warmup 3 --layers 4 --no-graph --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 252 -c 168 --csv --log-file gpurun_out/r1_launches.csv \
    python bench.py --steps 2 --warmup 3 --layers 4 --no-graph --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:wq_gemv -s 300 -c 3 -o gpurun_out/r1_gemv_full \
    python bench.py --steps 2 --warmup 3 --layers 4 --no-graph --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
