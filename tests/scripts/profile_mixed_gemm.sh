#!/bin/bash
# Shape sweep of the mixed-precision GEMM profiler ops, B200 build of the reference's script/profile_mixed_gemm.sh: the same
# 3 x 3 x 3 (M, N, K) grid, one ckProfiler invocation per shape, "Perf:" and "PerfJSON:" lines on stdout.
#   usage: profile_mixed_gemm.sh <op> <datatype> <layout> <verify> <init> <log> <time> <KBatch>
#   e.g.   profile_mixed_gemm.sh gemm_universal_reduce 0 0 1 2 0 1 4        (bf16 x int8, mk_kn_mn, KBatch 4)
#          CKP_ROTATING_MB=512 profile_mixed_gemm.sh gemm_w4a16 ...          (cold-cache timing)
DRIVER="$(dirname "$0")/../profiler/bin/ckProfiler"
OP=$1; DATATYPE=$2; LAYOUT=$3; VERIFY=$4; INIT=$5; LOG=$6; TIME=$7; KBatch=$8
for M in 16 2048 8192; do
  for N in 16 2048 8192; do
    for K in 1024 8192 65536; do
      if [ "$OP" = "gemm_w4a16" ]; then
        [ $N -lt 128 ] && continue
        $DRIVER $OP $VERIFY $INIT $LOG $TIME $M $N $K 0
      else
        $DRIVER $OP $DATATYPE $LAYOUT $VERIFY $INIT $LOG $TIME $M $N $K -1 -1 -1 $KBatch
      fi
    done
  done
done
