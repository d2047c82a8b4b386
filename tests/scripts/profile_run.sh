# round-2 evidence (run under gpurun): ncu launch list of the bench command (kernel shares of the step) and one full capture of
# the W4A16 decode kernels on the bench shapes.  Each ncu pass only after the same command has exited 0 without ncu.
set -x
B="python bench.py --steps 2 --warmup 3 --layers 4 --no-graph --no-cpu-baseline --no-extras"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 252 -c 168 --csv --log-file gpurun_out/r2_launches_bench_layers4_nograph.csv $B > gpurun_out/ncu_list.log 2>&1
tail -2 gpurun_out/ncu_list.log
bash tests/scripts/ncu_gemv.sh
