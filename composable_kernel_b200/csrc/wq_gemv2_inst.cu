// wq_gemv2_inst.cu -- explicit instantiations of the W4A16 decode kernel (wq_gemv2_kernel.cuh), one translation unit per
// (activation type, token-tile count): built six times with -DGV_INST_F16={0,1} -DGV_INST_NT={1,2,4}.
#include "wq_gemv2_kernel.cuh"

#ifndef GV_INST_F16
#error "compile with -DGV_INST_F16=0|1 -DGV_INST_NT=1|2|4"
#endif

namespace ckb {

#define GV_JOIN2(a, b) a##b
#define GV_JOIN(a, b) GV_JOIN2(a, b)
#define GV_FN GV_JOIN(GV_JOIN(launch_gemv2_f, GV_INST_F16), GV_JOIN(_nt, GV_INST_NT))

int GV_FN(int nt, int gw, int kbu, const G2Plan* gp, const G2Args& a, int grid, cudaStream_t st)
{
    constexpr bool F16 = GV_INST_F16 != 0;
    const bool zp      = a.has_zero != 0;
    if(a.wf != 0)
    {
        // int8-weight formats (K-major int8 streamed by TMA): Multiply / PassThrough for bf16 A, fpAintB / PassThrough for fp16 A
#define GV_CASE8(NT_, GW_)                                                                                                   \
    if(nt == NT_ && gw == GW_)                                                                                               \
    {                                                                                                                        \
        if(a.wf == 3) return launch_gemv2<NT_, GW_, 2, F16, false, false, false, 3>(gp, a, grid, st);                       \
        return launch_gemv2<NT_, GW_, 2, F16, false, false, false, (F16 ? 2 : 1)>(gp, a, grid, st);                          \
    }
#if GV_INST_NT == 1
        GV_CASE8(1, 1) GV_CASE8(1, 2) GV_CASE8(1, 3) GV_CASE8(1, 4)
#elif GV_INST_NT == 2
        GV_CASE8(2, 1) GV_CASE8(2, 2) GV_CASE8(2, 3) GV_CASE8(2, 4)
#else
        GV_CASE8(4, 1) GV_CASE8(4, 2)
#endif
#undef GV_CASE8
        return set_error(CKB_ERR_INVALID, "no W8 decode kernel variant for nt %d gw %d", nt, gw);
    }
#define GV_CASE(NT_, GW_)                                                                                                    \
    if(nt == NT_ && gw == GW_)                                                                                               \
    {                                                                                                                        \
        if(a.epi) return zp ? launch_gemv2<NT_, GW_, 2, F16, true, false, false, 4>(gp, a, grid, st) : launch_gemv2<NT_, GW_, 2, F16, false, false, false, 4>(gp, a, grid, st); \
        if(a.ext) return zp ? launch_gemv2<NT_, GW_, 2, F16, true, false, true>(gp, a, grid, st) : launch_gemv2<NT_, GW_, 2, F16, false, false, true>(gp, a, grid, st); \
        if(a.rtn) return zp ? launch_gemv2<NT_, GW_, 2, F16, true, true, false>(gp, a, grid, st) : launch_gemv2<NT_, GW_, 2, F16, false, true, false>(gp, a, grid, st); \
        if(kbu == 2) return zp ? launch_gemv2<NT_, GW_, 2, F16, true, false, false>(gp, a, grid, st) : launch_gemv2<NT_, GW_, 2, F16, false, false, false>(gp, a, grid, st); \
        return zp ? launch_gemv2<NT_, GW_, 1, F16, true, false, false>(gp, a, grid, st) : launch_gemv2<NT_, GW_, 1, F16, false, false, false>(gp, a, grid, st); \
    }
#if GV_INST_NT == 1
    GV_CASE(1, 1) GV_CASE(1, 2) GV_CASE(1, 3) GV_CASE(1, 4)
#elif GV_INST_NT == 2
    GV_CASE(2, 1) GV_CASE(2, 2) GV_CASE(2, 3) GV_CASE(2, 4)
#else
    GV_CASE(4, 1) GV_CASE(4, 2)
#endif
#undef GV_CASE
    return set_error(CKB_ERR_INVALID, "no W4 decode kernel variant for nt %d gw %d", nt, gw);
}

} // namespace ckb
