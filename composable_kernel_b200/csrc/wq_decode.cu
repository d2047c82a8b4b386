// wq_decode.cu -- K2: weight-quantized GEMM for decode shapes (M <= 64), HBM-bound.
//
//   E[M,N] = cde( A[M,K] * dequant(B)[K,N], D0, D1 )
//
// Roles are swapped with respect to CK (whose kernels put the weight on the B side of the MFMA):
// the WEIGHT is the tcgen05 A operand (MMA-M = 128 output channels n) and the ACTIVATIONS are the
// tcgen05 B operand (MMA-N = padded M), so one CTA streams 128 channels x K of weights.
//   * weights: HBM --TMA/bulk copy--> smem ring --ld.shared--> registers --unpack/dequant-->
//     tcgen05.st --> TMEM (A operand; no second trip through shared memory)
//   * activations: TMA, 128B swizzle, K-major tcgen05 B operand
//   * accumulators: fp32 in TMEM, drained once per 128-k unit and combined in registers with the
//     (k-group, n) scale -- exactly the K-blocked scale placement of CK's ab_scale GEMM
//     (grid/gridwise_gemm_xdl_cshuffle_v3_multi_d_ab_scale.hpp), here with ScaleBlockK = 128.
//   * stream-K: the (n-tile, k-unit) space is cut into gridDim.x equal contiguous ranges
//     (BlockToCTileMap_GemmStreamK, grid/block_to_ctile_map.hpp:1016-1400); partial tiles are reduced
//     with fp32 red.add into a workspace and finished by the last arriving CTA
//     (StreamKReductionStrategy::Atomic + the fix-up of DeviceGemm_Xdl_CShuffleV3R1's reduce kernel).
// Replaces kernel_fpAintB_gemm_wmma / kernel_gemm_xdl_cshuffle_v3 (multi-ABD) for M <= 64.
#include "common.cuh"
#include "dequant.cuh"
#include "epilogue.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace ckb {
namespace {

constexpr int DQ_UNIT_N  = 128;
constexpr int DQ_UNIT_K  = 128;
constexpr int DQ_AST     = 3;  // TMEM A-operand stages (64 columns each)
constexpr int DQ_DBUF    = 4;  // TMEM accumulator buffers
constexpr int DQ_THREADS = 512;
constexpr int DQ_WARP_EPI0 = 4, DQ_WARP_DQ0 = 8;
constexpr int DQ_SMEM_BUDGET = 100 * 1024;

struct DQArgs
{
    int M, N, K;
    int n_tiles, k_units;
    long long total_units;
    const uint8_t* w_packed;  // W4: pre-packed tiles [n_tile][k_unit][8192]
    const void* scale;        // GROUP: [k_units][N]; per-channel: [N]
    const uint8_t* zero;      // GROUP: [k_units][N] or null
    int zero_const;
    float* ws;                // [n_tiles][N_MMA][128] fp32, zero on entry and on exit
    int* counters;            // [n_tiles], zero on entry and on exit
    uint32_t idesc;
    int stages;
    EpiParams ep;
};

template <int WF>
struct WTraits
{
    static constexpr int kUnitBytes = (WF == WF_W4_GROUP) ? 8192 : 16384;
};

__device__ __forceinline__ void
store_one(const EpiParams& ep, int m, int n, float acc)
{
    float d0 = 0.f, d1 = 0.f;
    if(ep.num_d > 0) d0 = load_d_f32(ep.d_dtype[0], ep.d[0], (long long)m * ep.d_sm[0] + (long long)n * ep.d_sn[0]);
    if(ep.num_d > 1) d1 = load_d_f32(ep.d_dtype[1], ep.d[1], (long long)m * ep.d_sm[1] + (long long)n * ep.d_sn[1]);
    const float v       = cde_apply_f32(ep, acc, d0, d1);
    const long long off = (long long)m * ep.lde + n;
    if(ep.e_dtype == CKB_BF16) static_cast<uint16_t*>(ep.e)[off] = f32_to_bf16_bits(v, ep.flags);
    else if(ep.e_dtype == CKB_F16) static_cast<__half*>(ep.e)[off] = __float2half_rn(v);
    else static_cast<float*>(ep.e)[off] = v;
}

template <int N_MMA, int WF, bool F16>
__global__ void __launch_bounds__(DQ_THREADS, 1)
wq_decode_kernel(const __grid_constant__ CUtensorMap tmAct, const __grid_constant__ CUtensorMap tmW,
                 const DQArgs args)
{
    constexpr int ACT_BYTES   = N_MMA * 256;               // two boxes of [N_MMA][64 k] 16-bit
    constexpr int W_BYTES     = WTraits<WF>::kUnitBytes;
    constexpr int AUX_BYTES   = 1024;                      // scales (256) + zero points (128)
    constexpr int STAGE_BYTES = ACT_BYTES + W_BYTES + AUX_BYTES;
    constexpr int TMEM_COLS   = (DQ_AST * 64 + DQ_DBUF * N_MMA) <= 256 ? 256 : 512;
    constexpr int D_COL0      = DQ_AST * 64;
    constexpr bool GROUP      = (WF == WF_W4_GROUP);

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem            = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t smem_base = ptx::smem_u32(smem);
    const int STAGES         = args.stages;
    const uint32_t bar_base  = smem_base + STAGES * STAGE_BYTES;
    auto full_bar   = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar  = [&](int s) { return bar_base + 8u * (8 + s); };
    auto afull_bar  = [&](int a) { return bar_base + 8u * (16 + a); };
    auto aempty_bar = [&](int a) { return bar_base + 8u * (20 + a); };
    auto dfull_bar  = [&](int b) { return bar_base + 8u * (24 + b); };
    auto dempty_bar = [&](int b) { return bar_base + 8u * (28 + b); };
    const uint32_t tmem_slot = bar_base + 8u * 32;
    volatile int* flag_smem  = reinterpret_cast<volatile int*>(smem + STAGES * STAGE_BYTES + 8 * 33);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if(warp == 0 && lane == 0) ptx::prefetch_tmap(&tmAct);
    if(warp == 1 && lane == 0)
    {
        for(int s = 0; s < STAGES; ++s)
        {
            ptx::mbar_init(full_bar(s), 1);
            ptx::mbar_init(empty_bar(s), 8 + 1 + 4); // dequant warps + MMA commit + epilogue warps
        }
        for(int a = 0; a < DQ_AST; ++a)
        {
            ptx::mbar_init(afull_bar(a), 8);
            ptx::mbar_init(aempty_bar(a), 1);
        }
        for(int b = 0; b < DQ_DBUF; ++b)
        {
            ptx::mbar_init(dfull_bar(b), 1);
            ptx::mbar_init(dempty_bar(b), 4);
        }
        ptx::fence_mbar_init();
    }
    if(warp == 2)
    {
        ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
        ptx::tmem_relinquish();
    }
    ptx::tcgen05_fence_before_sync();
    __syncthreads();
    ptx::tcgen05_fence_after_sync();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    ptx::pdl_launch_dependents();

    const long long u_begin = args.total_units * blockIdx.x / gridDim.x;
    const long long u_end   = args.total_units * (blockIdx.x + 1) / gridDim.x;
    const int n_my          = int(u_end - u_begin);
    const int k_units       = args.k_units;

    if(warp == 0)
    {
        // ===================== producer =====================
        if(lane == 0)
        {
            const uint64_t pol_stream = ptx::policy_evict_first();
            const int elem = 2;
            auto issue_weights = [&](int i) {
                const long long u = u_begin + i;
                const int nt = int(u / k_units), kg = int(u % k_units);
                const int s  = i % STAGES;
                const uint32_t st = smem_base + s * STAGE_BYTES;
                const int n0 = nt * DQ_UNIT_N;
                const int rows = min(DQ_UNIT_N, args.N - n0);
                uint32_t bytes = ACT_BYTES + W_BYTES;
                if(GROUP) bytes += rows * elem + (args.zero ? rows : 0);
                ptx::mbar_arrive_expect_tx(full_bar(s), bytes);
                if(GROUP)
                {
                    ptx::bulk_load_1d(st + ACT_BYTES, args.w_packed + (size_t(nt) * k_units + kg) * W_BYTES,
                                      W_BYTES, full_bar(s), pol_stream);
                    ptx::bulk_load_1d(st + ACT_BYTES + W_BYTES,
                                      static_cast<const uint8_t*>(args.scale) + (size_t(kg) * args.N + n0) * elem,
                                      rows * elem, full_bar(s), pol_stream);
                    if(args.zero)
                        ptx::bulk_load_1d(st + ACT_BYTES + W_BYTES + 256, args.zero + size_t(kg) * args.N + n0,
                                          rows, full_bar(s), pol_stream);
                }
                else
                {
                    ptx::tma_load_2d_hint(st + ACT_BYTES, &tmW, full_bar(s), kg * DQ_UNIT_K, n0, pol_stream);
                }
            };
            auto issue_act = [&](int i) {
                const long long u = u_begin + i;
                const int kg = int(u % k_units);
                const int s  = i % STAGES;
                const uint32_t st = smem_base + s * STAGE_BYTES;
                ptx::tma_load_2d(st, &tmAct, full_bar(s), kg * DQ_UNIT_K, 0);
                ptx::tma_load_2d(st + ACT_BYTES / 2, &tmAct, full_bar(s), kg * DQ_UNIT_K + 64, 0);
            };
            // weights do not depend on the previous kernel in the stream: start them before the
            // programmatic-dependent-launch wait, activations after it.
            const int pre = min(STAGES, n_my);
            for(int i = 0; i < pre; ++i) issue_weights(i);
            ptx::pdl_wait_prior_grids();
            for(int i = 0; i < pre; ++i) issue_act(i);
            for(int i = pre; i < n_my; ++i)
            {
                ptx::mbar_wait(empty_bar(i % STAGES), ((i / STAGES) & 1) ^ 1);
                issue_weights(i);
                issue_act(i);
            }
        }
    }
    else if(warp == 1)
    {
        // ===================== MMA issuer =====================
        for(int i = 0; i < n_my; ++i)
        {
            const int s = i % STAGES, a = i % DQ_AST, b = i % DQ_DBUF;
            ptx::mbar_wait(full_bar(s), (i / STAGES) & 1);
            ptx::mbar_wait(afull_bar(a), (i / DQ_AST) & 1);
            ptx::mbar_wait(dempty_bar(b), ((i / DQ_DBUF) & 1) ^ 1);
            ptx::tcgen05_fence_after_sync();
            if(lane == 0)
            {
                const uint32_t act    = smem_base + s * STAGE_BYTES;
                const uint32_t d_tmem = tmem_base + D_COL0 + b * N_MMA;
                const uint32_t a_tmem = tmem_base + a * 64;
#pragma unroll
                for(int j = 0; j < DQ_UNIT_K / 16; ++j)
                {
                    const uint64_t db = ptx::make_smem_desc_sw128(act + (j / 4) * (ACT_BYTES / 2) + (j % 4) * 32, 1024);
                    ptx::mma_f16_ts(d_tmem, a_tmem + j * 8, db, args.idesc, j != 0);
                }
                ptx::mma_commit(empty_bar(s));
                ptx::mma_commit(aempty_bar(a));
                ptx::mma_commit(dfull_bar(b));
            }
            __syncwarp();
        }
    }
    else if(warp >= DQ_WARP_DQ0)
    {
        // ===================== dequant: smem -> registers -> TMEM (A operand) =====================
        const int q   = warp & 3;            // TMEM lane quarter
        const int h   = (warp - DQ_WARP_DQ0) >> 2; // which 64-k half of the unit
        const int row = q * 32 + lane;       // channel inside the tile
        float ch_scale_f = 1.f;
        uint32_t ch_scale2 = 0;
        int cur_nt = -1;
        for(int i = 0; i < n_my; ++i)
        {
            const long long u = u_begin + i;
            const int nt = int(u / k_units);
            const int s = i % STAGES, a = i % DQ_AST;
            if((WF == WF_W8_MUL || WF == WF_W8_FPAINTB) && nt != cur_nt)
            {
                cur_nt       = nt;
                const int n  = min(nt * DQ_UNIT_N + row, args.N - 1);
                const uint16_t sb = static_cast<const uint16_t*>(args.scale)[n];
                ch_scale_f   = bf16_bits_to_f32(sb);
                ch_scale2    = uint32_t(sb) | (uint32_t(sb) << 16);
            }
            ptx::mbar_wait(full_bar(s), (i / STAGES) & 1);
            const uint8_t* st = smem + s * STAGE_BYTES + ACT_BYTES;
            uint32_t out[32];
            if(GROUP)
            {
                // tile layout [chunk j = 0..3][row][16 B]; this thread: chunks 2h, 2h+1 (32 k each)
                const uint4 c0 = *reinterpret_cast<const uint4*>(st + (2 * h) * 2048 + row * 16);
                const uint4 c1 = *reinterpret_cast<const uint4*>(st + (2 * h + 1) * 2048 + row * 16);
                const int z    = args.zero ? int(st[W_BYTES + 256 + row]) : args.zero_const;
                __syncwarp();
                if(lane == 0) ptx::mbar_arrive(empty_bar(s));
                const uint32_t zm = F16 ? (0x6400u + z) * 0x10001u : (0x4300u + z) * 0x10001u;
                dequant_w4_word<F16>(c0.x, zm, out + 0);
                dequant_w4_word<F16>(c0.y, zm, out + 4);
                dequant_w4_word<F16>(c0.z, zm, out + 8);
                dequant_w4_word<F16>(c0.w, zm, out + 12);
                dequant_w4_word<F16>(c1.x, zm, out + 16);
                dequant_w4_word<F16>(c1.y, zm, out + 20);
                dequant_w4_word<F16>(c1.z, zm, out + 24);
                dequant_w4_word<F16>(c1.w, zm, out + 28);
            }
            else
            {
                // TMA SWIZZLE_128B tile [row][128 B]: 16-byte chunk c of row r lives at c ^ (r & 7)
                uint4 c[4];
#pragma unroll
                for(int jj = 0; jj < 4; ++jj)
                    c[jj] = *reinterpret_cast<const uint4*>(st + row * 128 + (((4 * h + jj) ^ (row & 7)) * 16));
                __syncwarp();
                if(lane == 0) ptx::mbar_arrive(empty_bar(s));
#pragma unroll
                for(int jj = 0; jj < 4; ++jj)
                {
                    const uint32_t w4[4] = {c[jj].x, c[jj].y, c[jj].z, c[jj].w};
#pragma unroll
                    for(int t = 0; t < 4; ++t)
                    {
                        uint32_t* o = out + jj * 8 + t * 2;
                        if(WF == WF_W8_MUL) dequant_s8x4_mul_bf16(w4[t], ch_scale_f, o);
                        else if(WF == WF_W8_FPAINTB) dequant_u8x4_fpaintb(w4[t], ch_scale2, o);
                        else dequant_s8x4_pass<F16>(w4[t], o);
                    }
                }
            }
            ptx::mbar_wait(aempty_bar(a), ((i / DQ_AST) & 1) ^ 1);
            ptx::tcgen05_fence_after_sync();
            const uint32_t taddr = tmem_base + a * 64 + h * 32 + (uint32_t(q * 32) << 16);
            ptx::tmem_st_x16(taddr, out);
            ptx::tmem_st_x16(taddr + 16, out + 16);
            ptx::tmem_st_wait();
            ptx::tcgen05_fence_before_sync();
            __syncwarp();
            if(lane == 0) ptx::mbar_arrive(afull_bar(a));
        }
    }
    else if(warp >= DQ_WARP_EPI0)
    {
        // ===================== epilogue: TMEM -> registers, group scale, stream-K reduction =====================
        ptx::pdl_wait_prior_grids();
        const int q   = warp & 3;
        const int row = q * 32 + lane;
        const EpiParams& ep = args.ep;
        float acc[N_MMA];
#pragma unroll
        for(int j = 0; j < N_MMA; ++j) acc[j] = 0.f;
        int seg_units = 0;
        for(int i = 0; i < n_my; ++i)
        {
            const long long u = u_begin + i;
            const int nt = int(u / k_units), kg = int(u % k_units);
            const int s = i % STAGES, b = i % DQ_DBUF;
            float sc = 1.f;
            ptx::mbar_wait(full_bar(s), (i / STAGES) & 1);
            if(GROUP)
            {
                const uint16_t sb = *reinterpret_cast<const uint16_t*>(smem + s * STAGE_BYTES + ACT_BYTES + W_BYTES + row * 2);
                sc = F16 ? __half2float(__ushort_as_half(sb)) : bf16_bits_to_f32(sb);
            }
            __syncwarp();
            if(lane == 0) ptx::mbar_arrive(empty_bar(s));
            ptx::mbar_wait(dfull_bar(b), (i / DQ_DBUF) & 1);
            ptx::tcgen05_fence_after_sync();
            const uint32_t taddr = tmem_base + D_COL0 + b * N_MMA + (uint32_t(q * 32) << 16);
#pragma unroll
            for(int c = 0; c < N_MMA; c += 16)
            {
                uint32_t r[16];
                ptx::tmem_ld_x16(taddr + c, r);
                ptx::tmem_ld_wait();
#pragma unroll
                for(int j = 0; j < 16; ++j) acc[c + j] = fmaf(sc, __uint_as_float(r[j]), acc[c + j]);
            }
            ptx::tcgen05_fence_before_sync();
            __syncwarp();
            if(lane == 0) ptx::mbar_arrive(dempty_bar(b));
            ++seg_units;

            const bool seg_end = (i + 1 == n_my) || (kg + 1 == k_units);
            if(!seg_end) continue;
            const int n = nt * DQ_UNIT_N + row;
            if(seg_units == k_units)
            {
                // this CTA owned the whole K range of the tile: finish directly
                if(n < args.N)
                {
#pragma unroll
                    for(int j = 0; j < N_MMA; ++j)
                        if(j < args.M) store_one(ep, j, n, acc[j]);
                }
            }
            else
            {
                float* ws = args.ws + (size_t(nt) * N_MMA) * DQ_UNIT_N + row;
#pragma unroll
                for(int j = 0; j < N_MMA; ++j)
                    if(j < args.M) ptx::red_add_f32(ws + j * DQ_UNIT_N, acc[j]);
                __threadfence();
                ptx::named_bar_sync(1, 128);
                if(warp == DQ_WARP_EPI0 && lane == 0)
                {
                    const int old = atomicAdd(args.counters + nt, seg_units);
                    *flag_smem    = (old + seg_units == k_units) ? 1 : 0;
                }
                ptx::named_bar_sync(1, 128);
                const int last = *flag_smem;
                ptx::named_bar_sync(1, 128);
                if(last)
                {
                    __threadfence();
#pragma unroll
                    for(int j = 0; j < N_MMA; ++j)
                    {
                        if(j < args.M)
                        {
                            const float v = __ldcg(ws + j * DQ_UNIT_N);
                            __stcg(ws + j * DQ_UNIT_N, 0.f);
                            if(n < args.N) store_one(ep, j, n, v);
                        }
                    }
                    if(warp == DQ_WARP_EPI0 && lane == 0) args.counters[nt] = 0;
                }
            }
#pragma unroll
            for(int j = 0; j < N_MMA; ++j) acc[j] = 0.f;
            seg_units = 0;
        }
    }

    ptx::tcgen05_fence_before_sync();
    __syncthreads();
    if(warp == 2)
    {
        ptx::tcgen05_fence_after_sync();
        ptx::tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

template <int N_MMA>
constexpr int dq_stage_bytes(int wf) { return N_MMA * 256 + (wf == WF_W4_GROUP ? 8192 : 16384) + 1024; }

} // namespace

// ---------------------------------------------------------------------------------------------
namespace {

int wq_weight_format(const ckb_gemm_problem& p)
{
    switch(p.b_op)
    {
    case CKB_BOP_GROUP: return WF_W4_GROUP;
    case CKB_BOP_MULTIPLY: return WF_W8_MUL;
    case CKB_BOP_FPAINTB: return WF_W8_FPAINTB;
    default: return WF_W8_PASS;
    }
}

int decode_n_mma(int M) { return M <= 16 ? 16 : (M <= 32 ? 32 : 64); }

int decode_stages(int n_mma, int wf)
{
    const int stage  = n_mma * 256 + (wf == WF_W4_GROUP ? 8192 : 16384) + 1024;
    const int budget = (n_mma == 16 && wf == WF_W4_GROUP) ? DQ_SMEM_BUDGET : 200 * 1024;
    return std::max(2, std::min(8, budget / stage));
}

template <int N_MMA, int WF, bool F16>
int launch_decode(const WQPlan* wp, const DQArgs& a, const ckb_gemm_plan& plan, cudaStream_t stream)
{
    auto kfn = wq_decode_kernel<N_MMA, WF, F16>;
    static int smem_set = 0;
    if(smem_set < plan.smem_bytes)
    {
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, plan.smem_bytes));
        smem_set = plan.smem_bytes;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim          = dim3(plan.grid);
    cfg.blockDim         = dim3(DQ_THREADS);
    cfg.dynamicSmemBytes = plan.smem_bytes;
    cfg.stream           = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    cfg.numAttrs = 1;
    CKB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kfn, wp->tmAct, wp->tmW, a));
    return CKB_OK;
}

template <int N_MMA>
int dispatch_decode(int wf, bool f16, const WQPlan* wp, const DQArgs& a, const ckb_gemm_plan& plan, cudaStream_t st)
{
    switch(wf)
    {
    case WF_W4_GROUP:
        return f16 ? launch_decode<N_MMA, WF_W4_GROUP, true>(wp, a, plan, st)
                   : launch_decode<N_MMA, WF_W4_GROUP, false>(wp, a, plan, st);
    case WF_W8_MUL: return launch_decode<N_MMA, WF_W8_MUL, false>(wp, a, plan, st);
    case WF_W8_FPAINTB: return launch_decode<N_MMA, WF_W8_FPAINTB, true>(wp, a, plan, st);
    default:
        return f16 ? launch_decode<N_MMA, WF_W8_PASS, true>(wp, a, plan, st)
                   : launch_decode<N_MMA, WF_W8_PASS, false>(wp, a, plan, st);
    }
}

} // namespace

int plan_wq_decode(const ckb_gemm_problem& p, ckb_gemm_plan& plan)
{
    static_assert(sizeof(WQPlan) <= sizeof(plan.opaque), "opaque too small");
    WQPlan* wp      = reinterpret_cast<WQPlan*>(plan.opaque);
    const int wf    = wq_weight_format(p);
    const int n_mma = decode_n_mma(p.M);
    const bool f16  = p.a_dtype == CKB_F16;
    const CUtensorMapDataType adt = f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    if(int rc = encode_tmap_2d(&wp->tmAct, adt, 2, p.a, p.M, p.K, p.lda, n_mma, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    if(wf != WF_W4_GROUP)
    {
        if(int rc = encode_tmap_2d(&wp->tmW, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p.b, p.N, p.K, p.ldb, DQ_UNIT_N, 128,
                                   CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    }
    else
        wp->tmW = wp->tmAct;
    const int n_tiles = (p.N + DQ_UNIT_N - 1) / DQ_UNIT_N, k_units = (p.K + DQ_UNIT_K - 1) / DQ_UNIT_K;
    wp->n_mma          = n_mma;
    wp->units_total    = n_tiles * k_units;
    wp->units_per_tile = k_units;
    const int stages   = decode_stages(n_mma, wf);
    const int sms      = device_info().sm_count > 0 ? device_info().sm_count : 148;
    int per_sm         = 1;
    if(const char* e = getenv("CKB_DECODE_CTAS_PER_SM")) per_sm = std::max(1, atoi(e));
    plan.grid            = std::max(1, std::min(sms * per_sm, wp->units_total));
    plan.block           = DQ_THREADS;
    plan.smem_bytes      = stages * (n_mma * 256 + (wf == WF_W4_GROUP ? 8192 : 16384) + 1024) + 1024 + 512;
    plan.split_units     = stages;
    plan.workspace_bytes = uint64_t(n_tiles) * n_mma * DQ_UNIT_N * 4 + uint64_t(n_tiles) * 4;
    static const char* wfn[] = {"w4g128", "w8mul", "w8fpaintb", "w8pass"};
    snprintf(plan.kernel_name, sizeof plan.kernel_name, "wq_decode_%s_%s_tmemA_streamk_128x%dx128_s%d",
             wfn[wf], f16 ? "f16" : "bf16", n_mma, stages);
    return CKB_OK;
}

int run_wq_decode(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, void* workspace, cudaStream_t stream)
{
    const WQPlan* wp = reinterpret_cast<const WQPlan*>(plan.opaque);
    if(!workspace) return set_error(CKB_ERR_WORKSPACE, "decode kernel needs %llu bytes of zeroed workspace",
                                    (unsigned long long)plan.workspace_bytes);
    const int wf   = wq_weight_format(p);
    const bool f16 = p.a_dtype == CKB_F16;
    DQArgs a{};
    a.M = p.M; a.N = p.N; a.K = p.K;
    a.n_tiles     = (p.N + DQ_UNIT_N - 1) / DQ_UNIT_N;
    a.k_units     = (p.K + DQ_UNIT_K - 1) / DQ_UNIT_K;
    a.total_units = (long long)a.n_tiles * a.k_units;
    a.w_packed    = static_cast<const uint8_t*>(p.b);
    a.scale       = p.scale;
    a.zero        = static_cast<const uint8_t*>(p.zero);
    a.zero_const  = p.zero_const;
    a.ws          = static_cast<float*>(workspace);
    a.counters    = reinterpret_cast<int*>(static_cast<uint8_t*>(workspace) + size_t(a.n_tiles) * wp->n_mma * DQ_UNIT_N * 4);
    a.idesc       = ptx::make_idesc(1, f16 ? 0 : 1, f16 ? 0 : 1, 128, wp->n_mma);
    a.stages      = plan.split_units;
    fill_epi_params(p, a.ep);
    switch(wp->n_mma)
    {
    case 16: return dispatch_decode<16>(wf, f16, wp, a, plan, stream);
    case 32: return dispatch_decode<32>(wf, f16, wp, a, plan, stream);
    default: return dispatch_decode<64>(wf, f16, wp, a, plan, stream);
    }
}

} // namespace ckb
