// kernels.h -- internal interface between the C-ABI front end (api.cu) and the kernel families.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>

#include "../../include/ckb.h"

namespace ckb {

struct EpiParams;
void fill_epi_params(const ckb_gemm_problem& p, EpiParams& ep);

enum KernelId : int32_t {
    KID_NONE       = 0,
    KID_GEMM_SS    = 1, // K3: TMA + tcgen05 SS (int8 x int8 -> s32, bf16/f16 dense)
    KID_WQ_DECODE  = 2, // K2: weight-dequant -> TMEM A operand, stream-K, M <= 64
    KID_WQ_PREFILL = 3, // K1: weight dequant pass + dense tcgen05 SS GEMM
    KID_WQ_GEMV    = 4, // round-1 K2 (W4A16): N-split HMMA kernel, CTA-wide stages (kept behind CKB_GEMV=1 for comparison)
    KID_WQ_GEMV2   = 5, // K2 (W4A16): N-split HMMA kernel for M <= 32 per launch, rotating producer, no cross-CTA reduction
    KID_WQ_FUSED   = 6  // K1: fused TMA -> dequant -> tcgen05 cta_group::2 prefill kernel (no workspace)
};

// ---- K3 ---------------------------------------------------------------------------------------
struct SSPlan
{
    CUtensorMap tmA, tmB;
    CUtensorMap tmBh; // B with a 128-row box: the half of a 256-column tile one CTA of a pair loads
    int32_t pair;     // 1 = 2-CTA (cta_group::2) kernel, 256x256 tiles per CTA pair
};
int plan_gemm_ss(const ckb_gemm_problem& p, ckb_gemm_plan& plan);
int run_gemm_ss(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream);

// ---- K1 / K2 ----------------------------------------------------------------------------------
struct WQPlan
{
    CUtensorMap tmAct; // activations A[M][K] (MMA B operand)
    CUtensorMap tmW;   // plain int8/u8 weights [N][K] (unused for pre-packed int4)
    int32_t n_mma;     // MMA N (padded M tile)
    int32_t units_total, units_per_tile;
};
int plan_wq_fused(const ckb_gemm_problem& p, ckb_gemm_plan& plan);
int run_wq_fused(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream);
int plan_wq_decode(const ckb_gemm_problem& p, ckb_gemm_plan& plan);
int run_wq_decode(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, void* workspace, cudaStream_t stream);
// the same launch with the scratch in the library's own static device array (ckb_gemm_run_static_scratch, ckb.h): no allocation
int run_wq_decode_static(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream);
int plan_wq_gemv(const ckb_gemm_problem& p, ckb_gemm_plan& plan);
int run_wq_gemv(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream);
// schedule of the W4 decode kernel (host arithmetic only: exercised on CPU through ckb_gemv_schedule)
struct G2Sched
{
    int nt, rb, n_groups, kb_total, grid, passes, n_slots, slot_q, slot_r, g_hi, wn_log2, gw, kbu, upw, n_chunks, ring;
    int stage_bytes, off_w, off_s, off_z, red_off, bar_off, smem_bytes;
};
// mode bits: 1 = fixed K order (CKB_FLAG_FIXED_K_ORDER: 2 x 8 / 1 x 8 warps, one 128-k unit per warp and stage, whatever N is),
//            2 = RTN numerics (CKB_BOP_GROUP_RTN: scale applied per weight, 128-k units only)
int schedule_wq_gemv2(int M, int N, int K, bool zp, int sms, G2Sched& s, int mode = 0);
int plan_wq_gemv2(const ckb_gemm_problem& p, ckb_gemm_plan& plan);
int run_wq_gemv2(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream);
int launch_peer_wait(void* flag, void* epoch_state, unsigned per_epoch, cudaStream_t stream);
int plan_wq_prefill(const ckb_gemm_problem& p, ckb_gemm_plan& plan);
int run_wq_prefill(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, void* workspace, cudaStream_t stream);

// ---- K4: grouped GEMM, fixed N,K, device-resident argument array ---------------------------------
int grouped_arg_bytes(int num_b, int num_d);
int check_grouped(const ckb_gemm_problem& proto, int group_count);
int run_grouped(const ckb_gemm_problem& proto, const void* dev_args, int group_count, int sum_m, cudaStream_t stream);

// ---- pre-pack ---------------------------------------------------------------------------------
int launch_prepack_u8(const void* src, void* dst, long long n, cudaStream_t stream);
int launch_prepack_int4(const void* src, long long sk, long long sn, int N, int K, void* dst, cudaStream_t stream);
int launch_transpose_16(const void* src, long long ld_src, int rows, int cols, void* dst, long long ld_dst, cudaStream_t stream);
int launch_transpose_i8(const void* src, long long ld_src, int rows, int cols, void* dst, long long ld_dst, cudaStream_t stream);

} // namespace ckb
