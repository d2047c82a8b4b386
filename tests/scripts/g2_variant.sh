# Builds lib/variants/libckb_<NAME>.so: the release objects with the six wq_gemv2 instantiation units recompiled with extra defines.
#   tests/scripts/g2_variant.sh red4 -DCKB_G2_RED4=1      (select with CKB_LIB=composable_kernel_b200/lib/variants/libckb_red4.so)
set -e
NAME=$1; shift
cd "$(dirname "$0")/../../composable_kernel_b200"
make -j8 -s
mkdir -p build/variants lib/variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
OBJS=""
for i in f0_nt1 f0_nt2 f0_nt4 f1_nt1 f1_nt2 f1_nt4; do
  f16=${i:1:1}; nt=${i##*nt}
  nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC "$@" -DGV_INST_F16=$f16 -DGV_INST_NT=$nt -c csrc/wq_gemv2_inst.cu -o build/variants/wq_gemv2_inst_${i}_$NAME.o &
  OBJS="$OBJS build/variants/wq_gemv2_inst_${i}_$NAME.o"
done
wait
REST=$(ls build/*.o | grep -v wq_gemv2_inst_)
nvcc $ARCH -shared -o lib/variants/libckb_$NAME.so $REST $OBJS -cudart static
echo "built lib/variants/libckb_$NAME.so"
