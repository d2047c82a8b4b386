# K4 grouped fixed-NK kernel: parity tests, then A/B timing of the two forms
# (CKB_GROUPED_KERNEL=1 = register-prefetch form, 3 = four-stages-in-flight form, unset = dispatch on K)
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_cpp_examples.py tests/test_gpu_guards.py -x -q -m gpu -k "grouped or client" --timeout 300 2>&1 | tail -3
for f in 1 3; do CKB_GROUPED_KERNEL=$f timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "grouped" --timeout 300 2>&1 | tail -1; done
P=tests/profiler/bin/ckProfiler
M16=48,48,48,48,48,48,48,48,48,48,48,48,48,48,48,48
MR=33,61,40,52,35,63,47,58,32,44,39,55,50,36,60,42
M8=512,512,512,512,512,512,512,512
run() { echo -n "kernel=${F:-auto} layout=$1 Ms=$2 N=$3 K=$4: "; CKB_GROUPED_KERNEL=$F $P grouped_gemm_fixed_nk 0 $1 $5 2 0 1 $2 $3 $4 -1 -1 -1 1 5 $6 2>&1 | grep "Best Perf\|rror" | cut -c1-62; }
for F in "" 1 3; do
  for lay in 0 1; do
    run $lay $M16 1024 512 1 200
    run $lay $MR 1024 512 1 200
    run $lay $M16 4096 4096 1 50
    run $lay $M8 4096 4096 0 20
  done
done
