// wq_decode.cu -- K2: weight-quantized GEMM for decode shapes (M <= 64), HBM-bound.
//
//   E[M,N] = cde( A[M,K] * dequant(B)[K,N], D0, D1 )
//
// Roles are swapped with respect to CK (whose kernels put the weight on the B side of the MFMA):
// the WEIGHT is the tcgen05 A operand (MMA-M = 128 output channels n) and the ACTIVATIONS are the
// tcgen05 B operand (MMA-N = padded M), so one CTA streams 128 channels x K of weights.
//   * weights: HBM --bulk copy / TMA--> smem ring --ld.shared--> registers --unpack, (q-zp)*scale-->
//     tcgen05.st --> TMEM (the MMA A operand; the dequantised tile never returns to shared memory)
//   * activations: TMA, 128B swizzle, K-major tcgen05 B operand, own small smem ring
//   * accumulator: fp32 in TMEM for a whole (tile, K-range) segment, double buffered
//   * stream-K: the (n-tile, k-unit) space is cut into gridDim.x equal contiguous ranges
//     (BlockToCTileMap_GemmStreamK, grid/block_to_ctile_map.hpp:1016-1400); partial tiles are reduced
//     with fp32 red.add into a workspace and finished by the last arriving CTA
//     (StreamKReductionStrategy::Atomic + the fix-up of DeviceGemm_Xdl_CShuffleV3R1's reduce kernel).
// Replaces kernel_fpAintB_gemm_wmma / kernel_gemm_xdl_cshuffle_v3 (multi-ABD) for M <= 64.
//
// The pipeline is latency- rather than issue-bound (each barrier hand-off costs 100-400 cycles), so
// the design keeps the per-CTA pipeline short and runs TWO CTAs per SM:
//   warps 0 .. 4T-1  dequant teams (T teams of 4 warps; team t owns units t, t+T, ...; the team that
//                    dequantised a segment's last unit also drains the accumulator and runs the epilogue)
//   warp  4T         MMA issuer (+ TMEM allocation)
//   warp  4T+1       producer (bulk / TMA copies)
// Barriers: wfull[s] -> team -> wread[s] (weight slot free) ; team -> afull[a] ; actfull[t] ;
//           MMA commit -> mdone[m] (TMEM A stage + activation slot free) and, at a segment end, dfull[d];
//           epilogue -> dempty[d].
#include "common.cuh"
#include "dequant.cuh"
#include "epilogue.cuh"
#include "kernels.h"
#include "ptx.cuh"
#include "tile_map.cuh"

#include <vector>

namespace ckb {
namespace {

constexpr int DQ_UNIT_N  = 128;
constexpr int DQ_UNIT_K  = 128;
constexpr int DQ_TEAMS   = 2;
constexpr int DQ_THREADS = 128 * DQ_TEAMS + 64;
constexpr int DQ_WARP_MMA  = 4 * DQ_TEAMS;
constexpr int DQ_WARP_PROD = 4 * DQ_TEAMS + 1;
constexpr int DQ_MD        = 4;   // ring of per-unit "MMAs complete" barriers
constexpr int DQ_TMEM_COLS = 256;
constexpr int DQ_SMEM_EXCLUSIVE = 116 * 1024; // dynamic shared memory of a launch that must own its SM (see plan_wq_decode)

struct DQArgs
{
    int M, N, K;
    int n_tiles, k_units;
    long long total_units;
    const uint8_t* w_packed; // W4: pre-packed tiles [n_tile][k_unit][8192]
    const void* scale;       // GROUP: [k_units][N]; per-channel: [N]
    const uint8_t* zero;     // GROUP: [k_units][N] or null
    int zero_const;
    float* ws;               // [n_tiles][N_MMA][128] fp32, zero on entry and on exit
    int* counters;           // [n_tiles], zero on entry and on exit
    uint32_t idesc;
    long long* trace;        // debug timeline (CKB_TRACE), null in production
    long long* gtrace;       // debug: per-CTA globaltimer stamps (CKB_GTRACE), null in production
    EpiParams ep;
};

__host__ __device__ constexpr int dq_w_bytes(int wf) { return wf == WF_W4_GROUP ? 8192 : 16384; }
__host__ __device__ constexpr int dq_ast(int n_mma) { return n_mma <= 32 ? 3 : 2; }   // TMEM A stages
__host__ __device__ constexpr int dq_wst(int wf) { return wf == WF_W4_GROUP ? 10 : 5; } // 80 KB of weights in flight
__host__ __device__ constexpr int dq_actst(int n_mma) { return n_mma <= 32 ? 4 : 2; }
__host__ __device__ constexpr int dq_smem_bytes(int n_mma, int wf)
{
    return dq_actst(n_mma) * n_mma * 256 + dq_wst(wf) * dq_w_bytes(wf) + 1024 /*barriers*/ + 1024 /*align*/;
}

static __device__ __noinline__ void store_one(const EpiParams& ep, int m, int n, float acc)
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

// debug timeline: role r, unit i, slot k  (only block 0 records; 64 units max)
#ifdef CKB_ENABLE_TRACE
#define CKB_TR(r, i, k)                                                                                   \
    do {                                                                                                  \
        if(args.trace && blockIdx.x == 0 && (i) < 64) args.trace[((r)*64 + (i)) * 4 + (k)] = clock64();   \
    } while(0)
#define CKB_GT(k)                                                                                         \
    do {                                                                                                  \
        if(args.gtrace) args.gtrace[blockIdx.x * 8 + (k)] = (long long)ptx::global_timer_ns();            \
    } while(0)
#else
#define CKB_TR(r, i, k) do { } while(0)
#define CKB_GT(k) do { } while(0)
#endif

// ring position helper: advance by `step` without integer division
struct Ring
{
    int s;
    uint32_t ph;
    __device__ __forceinline__ void advance(int step, int size)
    {
        s += step;
        while(s >= size) { s -= size; ph ^= 1; }
    }
};

// Segment epilogue, deliberately NOT inlined and not unrolled: it runs once per (CTA, tile) and its
// code (all CDE functors x all dtypes) must stay out of the hot loops' instruction-cache footprint.
//   wait for the segment's MMAs, TMEM -> registers, then either the CDE epilogue (whole K range owned)
//   or the stream-K reduction: fp32 red.add into the workspace, last arriving CTA finishes the tile.
template <int N_MMA>
static __device__ __noinline__ void decode_drain(const DQArgs& args, int nt, int kg, int nt0, int kg0, uint32_t tmem_d0,
                                          uint32_t dfull0, uint32_t dempty0, volatile int* flag, int bar_id, int q,
                                          int lane)
{
    const EpiParams& ep = args.ep;
    const int row       = q * 32 + lane;
    const int sidx      = nt - nt0;
    const int d         = sidx & 1;
    const int kg_first  = (nt == nt0) ? kg0 : 0;
    const int seg_units = kg - kg_first + 1;
    const int k_units   = args.k_units;
    if(q == 0 && lane == 0) CKB_TR(4, sidx, 0);
    ptx::mbar_wait(dfull0 + 8u * d, (sidx >> 1) & 1);
    if(q == 0 && lane == 0) CKB_TR(4, sidx, 1);
    if(q == 0 && lane == 0) CKB_GT(6);
    ptx::tcgen05_fence_after_sync();
    float acc[N_MMA];
    const uint32_t taddr = tmem_d0 + d * N_MMA + (uint32_t(q * 32) << 16);
#pragma unroll
    for(int c = 0; c < N_MMA; c += 16)
    {
        uint32_t v[16];
        ptx::tmem_ld_x16(taddr + c, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for(int j = 0; j < 16; ++j) acc[c + j] = __uint_as_float(v[j]);
    }
    ptx::tcgen05_fence_before_sync();
    __syncwarp();
    if(lane == 0) ptx::mbar_arrive(dempty0 + 8u * d);
    const int n = nt * DQ_UNIT_N + row;
    // common case (PassThrough, 16-bit E, no D tensors): branch-free stores; otherwise the generic functor path
    const bool fast = ep.cde_op == CKB_CDE_PASSTHROUGH && (ep.e_dtype == CKB_BF16 || ep.e_dtype == CKB_F16);
    auto emit = [&](const float* v) {
        if(n >= args.N) return;
        if(fast)
        {
            uint16_t* e = static_cast<uint16_t*>(ep.e) + n;
#pragma unroll
            for(int j = 0; j < N_MMA; ++j)
                if(j < args.M)
                    e[(long long)j * ep.lde] = ep.e_dtype == CKB_BF16 ? f32_to_bf16_bits(v[j], ep.flags)
                                                                       : __half_as_ushort(__float2half_rn(v[j]));
        }
        else
        {
#pragma unroll
            for(int j = 0; j < N_MMA; ++j)
                if(j < args.M) store_one(ep, j, n, v[j]);
        }
    };
    if(seg_units == k_units)
    {
        emit(acc); // this CTA owned the whole K range of the tile
        return;
    }
    // partial tile: workspace layout [tile][row][N_MMA] so each thread reduces with 128-bit red.add
    // (rows >= M hold exact zeros: the activation box is zero-filled by TMA)
    float* ws = args.ws + ((size_t)nt * DQ_UNIT_N + row) * N_MMA;
#pragma unroll
    for(int j = 0; j < N_MMA; j += 4) ptx::red_add_v4_f32(ws + j, acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
    if(q == 0 && lane == 0) CKB_TR(4, sidx, 2);
    __threadfence();
    if(q == 0 && lane == 0) CKB_TR(4, sidx, 3);
    if(bar_id == 1) ptx::named_bar_sync<1>(128); else ptx::named_bar_sync<2>(128);
    if(q == 0 && lane == 0)
    {
        // release by the SIGNALLING thread, after the barrier (cumulative over the reds of the threads it has synchronised with;
        // the per-thread fences above sit in front of the barrier and do not order thread 0's arrival behind the others' reds)
        __threadfence();
        const int old = atomicAdd(args.counters + nt, seg_units);
        *flag         = (old + seg_units == k_units) ? 1 : 0;
    }
    if(bar_id == 1) ptx::named_bar_sync<1>(128); else ptx::named_bar_sync<2>(128);
    const int last = *flag;
    if(bar_id == 1) ptx::named_bar_sync<1>(128); else ptx::named_bar_sync<2>(128);
    if(q == 0 && lane == 0) CKB_TR(4, 32 + sidx, 0);
    if(last)
    {
        __threadfence();
        float4 v4[N_MMA / 4];
#pragma unroll
        for(int j = 0; j < N_MMA / 4; ++j) v4[j] = __ldcg(reinterpret_cast<const float4*>(ws) + j);
#pragma unroll
        for(int j = 0; j < N_MMA / 4; ++j) __stcg(reinterpret_cast<float4*>(ws) + j, make_float4(0.f, 0.f, 0.f, 0.f));
        if(q == 0 && lane == 0) args.counters[nt] = 0;
#pragma unroll
        for(int j = 0; j < N_MMA / 4; ++j)
        {
            acc[4 * j] = v4[j].x; acc[4 * j + 1] = v4[j].y; acc[4 * j + 2] = v4[j].z; acc[4 * j + 3] = v4[j].w;
        }
        emit(acc);
    }
}

template <int N_MMA, int WF, bool F16>
__global__ void __launch_bounds__(DQ_THREADS, 2)
wq_decode_kernel(const __grid_constant__ CUtensorMap tmAct, const __grid_constant__ CUtensorMap tmW,
                 const DQArgs args)
{
    constexpr int ACT_BYTES = N_MMA * 256; // two boxes of [N_MMA][64 k] 16-bit
    constexpr int W_BYTES   = dq_w_bytes(WF);
    constexpr int WST       = dq_wst(WF);      // weight ring stages
    constexpr int ACT_ST    = dq_actst(N_MMA); // activation ring stages
    constexpr int AST       = dq_ast(N_MMA);   // TMEM A-operand stages (64 columns each)
    constexpr int D_COL0    = AST * 64;        // two accumulators of N_MMA columns follow
    constexpr bool GROUP    = (WF == WF_W4_GROUP);
    static_assert(WST >= DQ_TEAMS && DQ_MD >= AST && DQ_MD % ACT_ST == 0 && AST * 64 + 2 * N_MMA <= DQ_TMEM_COLS,
                  "ring/TMEM budget");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem            = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t smem_base = ptx::smem_u32(smem);
    const uint32_t act_base  = smem_base;                      // [ACT_ST][ACT_BYTES]
    const uint32_t w_base    = smem_base + ACT_ST * ACT_BYTES; // [WST][W_BYTES]
    constexpr int BAR_OFF    = ACT_ST * ACT_BYTES + WST * W_BYTES;
    const uint32_t bar_base  = smem_base + BAR_OFF;
    auto wfull_bar   = [&](int s) { return bar_base + 8u * s; };        // weights landed
    auto wread_bar   = [&](int s) { return bar_base + 8u * (16 + s); }; // weights read by the team
    auto actfull_bar = [&](int t) { return bar_base + 8u * (32 + t); }; // activations landed
    auto mdone_bar   = [&](int m) { return bar_base + 8u * (36 + m); }; // unit's MMAs complete
    auto afull_bar   = [&](int a) { return bar_base + 8u * (40 + a); }; // TMEM A stage written
    auto dfull_bar   = [&](int d) { return bar_base + 8u * (44 + d); }; // segment accumulated
    auto dempty_bar  = [&](int d) { return bar_base + 8u * (46 + d); }; // accumulator drained
    const uint32_t tmem_slot = bar_base + 8u * 48;
    volatile int* flag_smem  = reinterpret_cast<volatile int*>(smem + BAR_OFF + 8 * 50); // one int per team

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if(threadIdx.x == 0) CKB_TR(5, 0, 0);
    if(threadIdx.x == 0) CKB_GT(0);

    if(warp == DQ_WARP_PROD && lane == 0)
    {
        ptx::prefetch_tmap(&tmAct);
        for(int s = 0; s < WST; ++s)
        {
            ptx::mbar_init(wfull_bar(s), 1);
            ptx::mbar_init(wread_bar(s), 4);
        }
        for(int t = 0; t < ACT_ST; ++t) ptx::mbar_init(actfull_bar(t), 1);
        for(int m = 0; m < DQ_MD; ++m) ptx::mbar_init(mdone_bar(m), 1);
        for(int a = 0; a < AST; ++a) ptx::mbar_init(afull_bar(a), 4);
        for(int d = 0; d < 2; ++d)
        {
            ptx::mbar_init(dfull_bar(d), 1);
            ptx::mbar_init(dempty_bar(d), 4);
        }
        ptx::fence_mbar_init();
    }
    if(warp == DQ_WARP_MMA)
    {
        ptx::tmem_alloc<DQ_TMEM_COLS>(tmem_slot);
        ptx::tmem_relinquish();
    }
    ptx::tcgen05_fence_before_sync();
    __syncthreads();
    ptx::tcgen05_fence_after_sync();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    ptx::pdl_launch_dependents();
    if(threadIdx.x == 0) CKB_TR(5, 0, 1);
    if(threadIdx.x == 0) CKB_GT(1);

    long long u_begin, u_end;
    streamk_range(args.total_units, int(blockIdx.x), int(gridDim.x), u_begin, u_end);
    const int n_my          = int(u_end - u_begin);
    const int k_units       = args.k_units;
    const int nt0           = int(u_begin / k_units); // first tile / k-unit of this CTA (one division)
    const int kg0           = int(u_begin - (long long)nt0 * k_units);
    auto adv = [&](int& nt_, int& kg_, int step) {
        kg_ += step;
        while(kg_ >= k_units) { kg_ -= k_units; ++nt_; }
    };

    if(warp == DQ_WARP_PROD)
    {
        // ===================== producer =====================
        if(lane == 0)
        {
            const uint64_t pol_stream = ptx::policy_evict_first();
            int ntw = nt0, kgw = kg0; // weight cursor
            int nta = nt0, kga = kg0; // activation cursor
            auto issue_w = [&](int s) {
                ptx::mbar_arrive_expect_tx(wfull_bar(s), W_BYTES);
                if(GROUP)
                    ptx::bulk_load_1d(w_base + s * W_BYTES, args.w_packed + ((size_t)ntw * k_units + kgw) * W_BYTES,
                                      W_BYTES, wfull_bar(s), pol_stream);
                else
                    ptx::tma_load_2d_hint(w_base + s * W_BYTES, &tmW, wfull_bar(s), kgw * DQ_UNIT_K, ntw * DQ_UNIT_N,
                                          pol_stream);
                adv(ntw, kgw, 1);
            };
            auto issue_a = [&](int t) {
                const uint32_t st = act_base + t * ACT_BYTES;
                ptx::mbar_arrive_expect_tx(actfull_bar(t), ACT_BYTES);
                ptx::tma_load_2d(st, &tmAct, actfull_bar(t), kga * DQ_UNIT_K, 0);
                ptx::tma_load_2d(st + ACT_BYTES / 2, &tmAct, actfull_bar(t), kga * DQ_UNIT_K + 64, 0);
                adv(nta, kga, 1);
            };
            // weights do not depend on the previous kernel in the stream: fill the weight ring before the
            // programmatic-dependent-launch wait, activations only after it.
            int iw = 0, ia = 0;
            for(; iw < min(WST, n_my); ++iw) issue_w(iw);
            CKB_TR(0, 0, 0);
            ptx::pdl_wait_prior_grids();
            CKB_TR(0, 0, 1);
            CKB_GT(2);
            for(; ia < min(ACT_ST, n_my); ++ia) issue_a(ia);
            Ring rw{0, 0}; // weight slot of unit iw; parity of its previous use's wread phase
            Ring rt{0, 0}; // activation slot of unit ia
            Ring rm{0, 0}; // mdone slot of unit ia - ACT_ST
            // Non-blocking polling of both rings: blocking on one could starve the other (the MMA that
            // frees an activation slot may itself be waiting for activations not issued yet).
            while(iw < n_my || ia < n_my)
            {
                bool progress = false;
                if(iw < n_my && ptx::mbar_test_wait(wread_bar(rw.s), rw.ph))
                {
                    issue_w(rw.s);
                    rw.advance(1, WST);
                    ++iw;
                    progress = true;
                }
                if(ia < n_my && ptx::mbar_test_wait(mdone_bar(rm.s), rm.ph))
                {
                    issue_a(rt.s);
                    rt.advance(1, ACT_ST);
                    rm.advance(1, DQ_MD);
                    ++ia;
                    progress = true;
                }
                if(!progress) __nanosleep(20);
            }
        }
    }
    else if(warp == DQ_WARP_MMA)
    {
        // ===================== MMA issuer (warp-uniform loop, one elected lane issues) =====================
        // A single warp retires roughly one dependent instruction every 4-6 cycles, so this loop is kept
        // tiny: it is unrolled over lcm(AST, ACT_ST, DQ_MD) units, which turns every ring slot, barrier
        // address, TMEM column and smem descriptor into a compile-time offset.
        constexpr int UNR = (AST == 3) ? 12 : 4;
        static_assert(UNR % AST == 0 && UNR % ACT_ST == 0 && UNR % DQ_MD == 0, "unroll must cover all rings");
        constexpr uint32_t A_WRAPS_ODD = (UNR / AST) & 1, T_WRAPS_ODD = (UNR / ACT_ST) & 1;
        const uint64_t db_base = ptx::make_smem_desc_sw128(act_base, 1024);
        int nt = nt0, kg = kg0;
        uint32_t outer = 0; // parity of the unrolled-block index
        for(int i0 = 0; i0 < n_my; i0 += UNR, outer ^= 1)
        {
#pragma unroll
            for(int u = 0; u < UNR; ++u)
            {
                const int i = i0 + u;
                if(i >= n_my) break;
                constexpr int dummy = 0; (void)dummy;
                const int a_s = u % AST, t_s = u % ACT_ST, m_s = u % DQ_MD;
                const uint32_t a_ph = ((u / AST) & 1) ^ (A_WRAPS_ODD & outer);
                const uint32_t t_ph = ((u / ACT_ST) & 1) ^ (T_WRAPS_ODD & outer);
                const bool seg_first = (i == 0) || (kg == 0);
                const bool seg_last  = (i + 1 == n_my) || (kg + 1 == k_units);
                const int sidx       = nt - nt0; // segment index inside this CTA
                const int d          = sidx & 1;
                if(seg_first) ptx::mbar_wait(dempty_bar(d), ((sidx >> 1) & 1) ^ 1);
                ptx::mbar_wait2(afull_bar(a_s), a_ph, actfull_bar(t_s), t_ph);
                if(lane == 0) CKB_TR(1, i, 0);
                ptx::tcgen05_fence_after_sync();
                const uint32_t d_tmem = tmem_base + D_COL0 + d * N_MMA;
                const uint32_t a_tmem = tmem_base + a_s * 64;
                if(ptx::elect_one())
                {
#pragma unroll
                    for(int j = 0; j < DQ_UNIT_K / 16; ++j)
                    {
                        // descriptor start address is in 16-byte units: +32 B per k-step, second box after ACT_BYTES/2
                        const uint64_t db = db_base + uint64_t((t_s * ACT_BYTES + (j / 4) * (ACT_BYTES / 2) + (j % 4) * 32) >> 4);
                        ptx::mma_f16_ts(d_tmem, a_tmem + j * 8, db, args.idesc, (j != 0 || !seg_first) ? 1u : 0u);
                    }
                    ptx::mma_commit(mdone_bar(m_s));
                    if(seg_last) ptx::mma_commit(dfull_bar(d));
                }
                __syncwarp();
                if(lane == 0) CKB_TR(1, i, 1);
                if(++kg == k_units) { kg = 0; ++nt; }
            }
        }
    }
    else
    {
        // ===================== dequant teams: smem -> registers -> TMEM (A operand); segment epilogue ============
        const int q    = warp & 3;  // TMEM lane quarter
        const int team = warp >> 2;
        const int row  = q * 32 + lane; // channel inside the tile
        const bool tr  = (q == 0 && lane == 0);
        float ch_scale_f   = 1.f;
        uint32_t ch_scale2 = 0;
        int cur_nt = -1;
        bool pdl_waited = false;

        auto load_scale2 = [&](int nt_, int kg_) -> uint32_t {
            const int n_      = min(nt_ * DQ_UNIT_N + row, args.N - 1);
            const uint32_t sb = __ldg(static_cast<const uint16_t*>(args.scale) + (size_t)kg_ * args.N + n_);
            return sb | (sb << 16);
        };
        auto load_zero = [&](int nt_, int kg_) -> int {
            if(!args.zero) return args.zero_const;
            return int(__ldg(args.zero + (size_t)kg_ * args.N + min(nt_ * DQ_UNIT_N + row, args.N - 1)));
        };

        auto drain = [&](int nt, int kg) {
            if(!pdl_waited) { ptx::pdl_wait_prior_grids(); pdl_waited = true; }
            decode_drain<N_MMA>(args, nt, kg, nt0, kg0, tmem_base + D_COL0, dfull_bar(0), dempty_bar(0), flag_smem + team,
                                1 + team, q, lane);
        };

        // ---- one unit: ring positions are derived from the unit index with compile-time ring sizes ----
        auto unit = [&](int i, int nt, int kg, uint32_t sc2, int z) {
            if((WF == WF_W8_MUL || WF == WF_W8_FPAINTB) && nt != cur_nt)
            {
                cur_nt            = nt;
                const int n       = min(nt * DQ_UNIT_N + row, args.N - 1);
                const uint16_t sb = static_cast<const uint16_t*>(args.scale)[n];
                ch_scale_f        = bf16_bits_to_f32(sb);
                ch_scale2         = uint32_t(sb) | (uint32_t(sb) << 16);
            }
            const int ws_slot      = i % WST;
            const uint32_t ws_ph   = uint32_t(i / WST) & 1;
            const int a_s          = i % AST;
            const int ip           = i - AST; // previous user of this TMEM A stage
            if(tr) CKB_TR(2 + team, i, 0);
            ptx::mbar_wait(wfull_bar(ws_slot), ws_ph);
            if(tr) CKB_TR(2 + team, i, 1);
            if(tr && i == 0) CKB_GT(3);
            if(tr && i + DQ_TEAMS >= n_my) CKB_GT(4 + team);
            const uint8_t* st    = smem + ACT_ST * ACT_BYTES + ws_slot * W_BYTES;
            const uint32_t taddr = tmem_base + a_s * 64 + (uint32_t(q * 32) << 16);
            if(GROUP)
            {
                // tile layout [chunk j = 0..3][row][16 B]: chunk j = k 32j .. 32j+31 of this channel
                uint4 c[4];
#pragma unroll
                for(int j = 0; j < 4; ++j) c[j] = *reinterpret_cast<const uint4*>(st + j * 2048 + row * 16);
                __syncwarp();
                if(lane == 0) ptx::mbar_arrive(wread_bar(ws_slot));
                const uint32_t zm = F16 ? (0x6400u + z) * 0x10001u : (0x4300u + z) * 0x10001u;
                if(ip >= 0) ptx::mbar_wait(mdone_bar(ip % DQ_MD), uint32_t(ip / DQ_MD) & 1);
                if(tr) CKB_TR(2 + team, i, 2);
                ptx::tcgen05_fence_after_sync();
#pragma unroll
                for(int j = 0; j < 4; ++j)
                {
                    uint32_t out[16];
                    dequant_w4_word_scaled<F16>(c[j].x, zm, sc2, out + 0);
                    dequant_w4_word_scaled<F16>(c[j].y, zm, sc2, out + 4);
                    dequant_w4_word_scaled<F16>(c[j].z, zm, sc2, out + 8);
                    dequant_w4_word_scaled<F16>(c[j].w, zm, sc2, out + 12);
                    ptx::tmem_st_x16(taddr + j * 16, out);
                }
            }
            else
            {
                // TMA SWIZZLE_128B tile [row][128 B]: 16-byte chunk c of row r lives at c ^ (r & 7)
                uint4 c[8];
#pragma unroll
                for(int j = 0; j < 8; ++j)
                    c[j] = *reinterpret_cast<const uint4*>(st + row * 128 + ((j ^ (row & 7)) * 16));
                __syncwarp();
                if(lane == 0) ptx::mbar_arrive(wread_bar(ws_slot));
                if(ip >= 0) ptx::mbar_wait(mdone_bar(ip % DQ_MD), uint32_t(ip / DQ_MD) & 1);
                ptx::tcgen05_fence_after_sync();
#pragma unroll
                for(int jj = 0; jj < 8; jj += 2)
                {
                    uint32_t out[16];
#pragma unroll
                    for(int h = 0; h < 2; ++h)
                    {
                        const uint32_t w4[4] = {c[jj + h].x, c[jj + h].y, c[jj + h].z, c[jj + h].w};
#pragma unroll
                        for(int t = 0; t < 4; ++t)
                        {
                            uint32_t* o = out + h * 8 + t * 2;
                            if(WF == WF_W8_MUL) dequant_s8x4_mul_bf16(w4[t], ch_scale_f, o);
                            else if(WF == WF_W8_FPAINTB) dequant_u8x4_fpaintb(w4[t], ch_scale2, o);
                            else dequant_s8x4_pass<F16>(w4[t], o);
                        }
                    }
                    ptx::tmem_st_x16(taddr + (jj / 2) * 16, out);
                }
            }
            ptx::tmem_st_wait();
            ptx::tcgen05_fence_before_sync();
            __syncwarp();
            if(lane == 0) ptx::mbar_arrive(afull_bar(a_s));
            if(tr) CKB_TR(2 + team, i, 3);
            if((i + 1 == n_my) || (kg + 1 == k_units))
            {
                drain(nt, kg);
                if(tr) CKB_TR(5, 1 + i, team);
            }
        };

        // Group scales / zero points come straight from global memory, two team-iterations ahead, through
        // registers that are only read one full iteration after the load was issued (loop unrolled by 2).
        int nt = nt0, kg = kg0;
        adv(nt, kg, team);
        int ntp = nt, kgp = kg; // prefetch cursor
        uint32_t scA = 0, scB = 0;
        int zA = args.zero_const, zB = args.zero_const;
        if(GROUP)
        {
            if(team < n_my) { scA = load_scale2(ntp, kgp); zA = load_zero(ntp, kgp); }
            adv(ntp, kgp, DQ_TEAMS);
            if(team + DQ_TEAMS < n_my) { scB = load_scale2(ntp, kgp); zB = load_zero(ntp, kgp); }
            adv(ntp, kgp, DQ_TEAMS);
        }
        for(int i = team; i < n_my; i += 2 * DQ_TEAMS)
        {
            unit(i, nt, kg, scA, zA);
            adv(nt, kg, DQ_TEAMS);
            if(GROUP && i + 2 * DQ_TEAMS < n_my) { scA = load_scale2(ntp, kgp); zA = load_zero(ntp, kgp); }
            adv(ntp, kgp, DQ_TEAMS);
            if(i + DQ_TEAMS < n_my)
            {
                unit(i + DQ_TEAMS, nt, kg, scB, zB);
                adv(nt, kg, DQ_TEAMS);
                if(GROUP && i + 3 * DQ_TEAMS < n_my) { scB = load_scale2(ntp, kgp); zB = load_zero(ntp, kgp); }
                adv(ntp, kgp, DQ_TEAMS);
            }
        }
    }

    ptx::tcgen05_fence_before_sync();
    __syncthreads();
    if(threadIdx.x == 0) CKB_GT(7);
    if(warp == DQ_WARP_MMA)
    {
        ptx::tcgen05_fence_after_sync();
        ptx::tmem_dealloc<DQ_TMEM_COLS>(tmem_base);
    }
}

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

template <int N_MMA, int WF, bool F16>
int launch_decode(const WQPlan* wp, const DQArgs& a, const ckb_gemm_plan& plan, cudaStream_t stream)
{
    auto kfn = wq_decode_kernel<N_MMA, WF, F16>;
    constexpr int smem = dq_smem_bytes(N_MMA, WF);
    static PerDeviceOnce attr_once;
    uint64_t dev_bit = 0;
    if(attr_once.need(dev_bit))
    {
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, std::max(smem, DQ_SMEM_EXCLUSIVE)));
        // without this the driver sizes the L1/shared carve-out for ONE block and the second CTA never fits
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kfn, cudaFuncAttributePreferredSharedMemoryCarveout,
                                            cudaSharedmemCarveoutMaxShared));
        attr_once.mark(dev_bit);
        if(getenv("CK_LOGGING"))
        {
            int nb = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kfn, DQ_THREADS, smem);
            fprintf(stderr, "ckb: wq_decode<%d,%d,%d> smem %d B, %d threads, resident CTAs/SM = %d\n", N_MMA, WF, int(F16),
                    smem, DQ_THREADS, nb);
        }
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim          = dim3(plan.grid);
    cfg.blockDim         = dim3(DQ_THREADS);
    cfg.dynamicSmemBytes = std::max(smem, int(plan.smem_bytes));
    cfg.stream           = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    // Programmatic dependent launch only while the stream is being captured (kernel -> kernel edges of a graph), as in wq_gemv2:
    // an EAGER launch with the attribute costs ~20 us of host time against ~6 us without (tests/host_cost_probe.py).
    // CKB_DECODE_PDL=0/1 overrides.
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if(cudaStreamIsCapturing(stream, &cs) != cudaSuccess) { cudaGetLastError(); cs = cudaStreamCaptureStatusNone; }
    int pdl = cs == cudaStreamCaptureStatusActive;
    if(const char* e = getenv("CKB_DECODE_PDL")) pdl = atoi(e);
    cfg.numAttrs = pdl ? 1 : 0;
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

int dispatch_decode_n(int n_mma, int wf, bool f16, const WQPlan* wp, const DQArgs& a, const ckb_gemm_plan& plan,
                      cudaStream_t st)
{
    switch(n_mma)
    {
    case 16: return dispatch_decode<16>(wf, f16, wp, a, plan, st);
    case 32: return dispatch_decode<32>(wf, f16, wp, a, plan, st);
    default: return dispatch_decode<64>(wf, f16, wp, a, plan, st);
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
    const int sms      = device_info().sm_count > 0 ? device_info().sm_count : 148;
    // ONE CTA per SM.  With two co-resident CTAs (the round-1 choice for n_mma <= 32) 2 - 40 % of the launches on 11008 x 4096 and
    // 14336 x 4096 returned tiles with a few wrong k-chunks (timing-dependent; not the stream-K reduction: returning atomics,
    // strong loads and atomic read-and-clear changed nothing, tests/streamk_stress_probe.py, profiles/r2_streamk_two_cta_defect.txt);
    // with one CTA per SM: 0 of 1800.  The cause inside the TMEM-A pipeline is not found yet; one CTA is also 6 - 10 % FASTER up to
    // 14336 x 4096 and 5 - 9 % slower at 28672 x 8192.  CKB_DECODE_CTAS_PER_SM=2 reproduces the defect.
    int per_sm         = 1;
    if(const char* e = getenv("CKB_DECODE_CTAS_PER_SM")) per_sm = std::max(1, std::min(2, atoi(e)));
    if(dq_smem_bytes(n_mma, wf) > 113 * 1024) per_sm = 1;
    plan.grid            = std::max(1, std::min(sms * per_sm, wp->units_total));
    plan.block           = DQ_THREADS;
    // exclusive SM: the launch asks for more than half of the shared memory, so that a CTA of the NEXT stream-K launch (programmatic
    // dependent launch under graph replay) cannot become co-resident either
    plan.smem_bytes      = per_sm == 1 ? std::max(dq_smem_bytes(n_mma, wf), DQ_SMEM_EXCLUSIVE) : dq_smem_bytes(n_mma, wf);
    plan.split_units     = k_units;
    plan.workspace_bytes = uint64_t(n_tiles) * n_mma * DQ_UNIT_N * 4 + uint64_t(n_tiles) * 4;
    static const char* wfn[] = {"w4g128", "w8mul", "w8fpaintb", "w8pass"};
    snprintf(plan.kernel_name, sizeof plan.kernel_name, "wq_decode_%s_%s_tmemA_streamk_128x%dx128_t%dx%d",
             wfn[wf], f16 ? "f16" : "bf16", n_mma, DQ_TEAMS, per_sm);
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
    a.trace       = nullptr;
    a.gtrace      = nullptr;
    fill_epi_params(p, a.ep);
    if(const char* gf = getenv("CKB_GTRACE"))
    {
        // debug only: per-CTA globaltimer stamps of this launch, dumped as text (one line per CTA)
        static long long* gbuf = nullptr;
        const int NW = 8 * 1024;
        if(!gbuf) cudaMalloc(&gbuf, NW * sizeof(long long));
        cudaDeviceSynchronize();
        cudaMemset(gbuf, 0, NW * sizeof(long long));
        cudaDeviceSynchronize();
        a.gtrace = gbuf;
        const int rc = dispatch_decode_n(wp->n_mma, wf, f16, wp, a, plan, stream);
        cudaDeviceSynchronize();
        std::vector<long long> hb(NW);
        cudaMemcpy(hb.data(), gbuf, NW * sizeof(long long), cudaMemcpyDeviceToHost);
        if(FILE* f = fopen(gf, "a"))
        {
            long long t0 = 0;
            for(int i = 0; i < NW; ++i) if(hb[i] && (!t0 || hb[i] < t0)) t0 = hb[i];
            fprintf(f, "# M%d N%d K%d grid %d\n", p.M, p.N, p.K, plan.grid);
            for(int b = 0; b < plan.grid && b < 1024; ++b)
            {
                fprintf(f, "%d", b);
                for(int k = 0; k < 8; ++k) fprintf(f, " %lld", hb[b * 8 + k] ? hb[b * 8 + k] - t0 : -1);
                fprintf(f, "\n");
            }
            fclose(f);
        }
        return rc;
    }
    if(const char* tf = getenv("CKB_TRACE"))
    {
        // debug only: record block 0's timeline of this launch and dump it as text
        static long long* dbuf = nullptr;
        constexpr int NW = 6 * 64 * 4;
        static long long buf[NW];
        if(!dbuf) cudaMalloc(&dbuf, NW * sizeof(long long));
        cudaDeviceSynchronize();
        cudaMemset(dbuf, 0, NW * sizeof(long long));
        cudaDeviceSynchronize();
        a.trace = dbuf;
        const int rc = dispatch_decode_n(wp->n_mma, wf, f16, wp, a, plan, stream);
        cudaDeviceSynchronize();
        cudaMemcpy(buf, dbuf, NW * sizeof(long long), cudaMemcpyDeviceToHost);
        if(FILE* f = fopen(tf, "w"))
        {
            long long t0 = 0;
            for(int i = 0; i < NW; ++i) if(buf[i] && (!t0 || buf[i] < t0)) t0 = buf[i];
            const char* names[6] = {"prod", "mma", "dq0", "dq1", "drain", "misc"};
            for(int r = 0; r < 6; ++r)
                for(int i = 0; i < 64; ++i)
                {
                    const long long* e = buf + (r * 64 + i) * 4;
                    if(!(e[0] | e[1] | e[2] | e[3])) continue;
                    fprintf(f, "%s %d", names[r], i);
                    for(int k = 0; k < 4; ++k) fprintf(f, " %lld", e[k] ? e[k] - t0 : -1);
                    fprintf(f, "\n");
                }
            fclose(f);
        }
        return rc;
    }
    return dispatch_decode_n(wp->n_mma, wf, f16, wp, a, plan, stream);
}

// ---- ckb_gemm_run_static_scratch (ckb.h): the stream-K scratch in a static device array, for callers that may not allocate ----
// One array per device that loads the library (held until exit), shared by every call on that device.  The NAME below is a
// host-side stand-in in host code, not a device address: the address the kernel gets comes from cudaGetSymbolAddress, asked once
// per device with that device current.
static __device__ __align__(256) unsigned char g_static_scratch[CKB_STATIC_SCRATCH_BYTES];

int run_wq_decode_static(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream)
{
    if(plan.workspace_bytes > uint64_t(CKB_STATIC_SCRATCH_BYTES))
        return set_error(CKB_ERR_WORKSPACE, "static scratch: this problem needs %llu bytes of scratch, the built-in array holds %u (use ckb_gemm_run with a caller workspace)",
                         (unsigned long long)plan.workspace_bytes, unsigned(CKB_STATIC_SCRATCH_BYTES));
    if(getenv("CKB_GTRACE") || getenv("CKB_TRACE"))
        return set_error(CKB_ERR_INVALID, "static scratch: the CKB_GTRACE / CKB_TRACE debug recorders allocate device memory; unset them for this entry point");
    constexpr int MAX_DEV = 64;
    static void* dev_ptr[MAX_DEV]; // per device ordinal; written once (the address never changes), read with acquire
    int dev = 0;
    CKB_CUDA_CHECK(cudaGetDevice(&dev));
    if(dev < 0 || dev >= MAX_DEV) return set_error(CKB_ERR_INVALID, "static scratch: device ordinal %d is out of range", dev);
    void* ptr = __atomic_load_n(&dev_ptr[dev], __ATOMIC_ACQUIRE);
    if(!ptr)
    {
        CKB_CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_static_scratch)); // the current device's copy of the array
        if(!ptr) return set_error(CKB_ERR_INVALID, "static scratch: cudaGetSymbolAddress returned a null address");
        __atomic_store_n(&dev_ptr[dev], ptr, __ATOMIC_RELEASE);
    }
    // partial tiles and arrival counters must start at zero; the array keeps whatever the last call (possibly a failed or
    // differently shaped one) left behind, so the part this launch uses is cleared every time, in stream order
    CKB_CUDA_CHECK(cudaMemsetAsync(ptr, 0, size_t(plan.workspace_bytes), stream));
    return run_wq_decode(p, plan, ptr, stream);
}

} // namespace ckb
