# ONE ncu --set full capture covering the W4 decode kernels on the bench shapes (run under gpurun, after the plain run exits 0):
# three launches per shape, the third of each is the warm one.
S="w4 1 4096 4096 w4 1 11008 4096 w4 1 4096 11008 w4 4 4096 4096 w4 16 4096 4096 w4 16 11008 4096"
PERF_ONCE=1 python tests/perf_probe.py $S > gpurun_out/plain_once.log 2>&1 && \
PERF_ONCE=1 ncu --set full --clock-control none --import-source on -k regex:wq_gemv -c 18 -f -o gpurun_out/${TAG:-r2}_gemv_all \
    python tests/perf_probe.py $S > gpurun_out/ncu_all.log 2>&1
tail -3 gpurun_out/ncu_all.log
