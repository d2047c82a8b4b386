// wq_gemv_inst.cu -- explicit instantiations of the W4A16 decode kernel (wq_gemv_kernel.cuh), one translation unit per
// (activation type, token-tile count): built six times with -DGV_INST_F16={0,1} -DGV_INST_NT={1,2,4}; NT = 4 also carries NT = 8.
#include "wq_gemv_kernel.cuh"

#ifndef GV_INST_F16
#error "compile with -DGV_INST_F16=0|1 -DGV_INST_NT=1|2|4"
#endif

namespace ckb {

#define GV_JOIN2(a, b) a##b
#define GV_JOIN(a, b) GV_JOIN2(a, b)
#define GV_FN GV_JOIN(GV_JOIN(launch_gemv_f, GV_INST_F16), GV_JOIN(_nt, GV_INST_NT))

int GV_FN(int nt, int gw, const GVPlan* gp, const GVArgs& a, int grid, cudaStream_t st)
{
    constexpr bool F16 = GV_INST_F16 != 0;
#define GV_CASE(NT_, GW_)                                                                                      \
    if(nt == NT_ && gw == GW_)                                                                                 \
        return a.has_zero ? launch_gemv<NT_, GW_, F16, true>(gp, a, grid, st) : launch_gemv<NT_, GW_, F16, false>(gp, a, grid, st);
#if GV_INST_NT == 1
    GV_CASE(1, 1) GV_CASE(1, 2) GV_CASE(1, 3) GV_CASE(1, 4) GV_CASE(1, 5) GV_CASE(1, 6) GV_CASE(1, 7) GV_CASE(1, 8)
    GV_CASE(1, 9) GV_CASE(1, 10) GV_CASE(1, 11) GV_CASE(1, 12) GV_CASE(1, 13) GV_CASE(1, 14) GV_CASE(1, 15) GV_CASE(1, 16)
#elif GV_INST_NT == 2
    GV_CASE(2, 1) GV_CASE(2, 2) GV_CASE(2, 3) GV_CASE(2, 4) GV_CASE(2, 5) GV_CASE(2, 6) GV_CASE(2, 7) GV_CASE(2, 8)
#else
    GV_CASE(4, 1) GV_CASE(4, 2) GV_CASE(4, 3) GV_CASE(4, 4) GV_CASE(8, 1) GV_CASE(8, 2)
#endif
#undef GV_CASE
    return set_error(CKB_ERR_INVALID, "no W4 decode kernel variant for nt %d gw %d", nt, gw);
}

} // namespace ckb
