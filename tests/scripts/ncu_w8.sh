# one ncu --set full capture of the int8-weight decode kernels (HMMA N-split form and the stream-K tcgen05 form), third launch of each
S="w8 16 11008 4096 w8 1 4096 4096"
PERF_ONCE=1 python tests/perf_probe.py $S > gpurun_out/plain_once.log 2>&1 && \
PERF_ONCE=1 ncu --set full --clock-control none --import-source on -k regex:wq_ -c 6 -f -o gpurun_out/r2_w8_hmma python tests/perf_probe.py $S > gpurun_out/ncu_w8.log 2>&1
PERF_ONCE=1 CKB_W8_DECODE=streamk ncu --set full --clock-control none --import-source on -k regex:wq_ -c 6 -f -o gpurun_out/r2_w8_streamk python tests/perf_probe.py $S > gpurun_out/ncu_w8b.log 2>&1
tail -2 gpurun_out/ncu_w8b.log
