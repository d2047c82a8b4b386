# K4 grouped kernel: fixed cost vs K (client-31 groups), then one ncu --set full capture of the third launch
P=tests/profiler/bin/ckProfiler
M16=48,48,48,48,48,48,48,48,48,48,48,48,48,48,48,48
for f in 0 1; do
  for K in 128 256 512 1024 2048; do
    echo -n "kernel=$f layout=1 K=$K: "; CKB_GROUPED_KERNEL=$f $P grouped_gemm_fixed_nk 0 1 0 2 0 1 $M16 1024 $K -1 -1 -1 1 5 200 2>&1 | grep "Best Perf" | cut -c1-60
  done
done
$P grouped_gemm_fixed_nk 0 1 0 2 0 1 $M16 1024 512 -1 -1 -1 1 1 2 > gpurun_out/gg_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:grouped -s 2 -c 1 -f -o gpurun_out/r2_grouped_v3 $P grouped_gemm_fixed_nk 0 1 0 2 0 1 $M16 1024 512 -1 -1 -1 1 1 2 > gpurun_out/ncu_gg.log 2>&1
tail -2 gpurun_out/ncu_gg.log
