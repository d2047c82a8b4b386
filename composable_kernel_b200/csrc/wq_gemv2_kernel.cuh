// wq_gemv2_kernel.cuh -- K2 (round 2): W4A16 group-128 GEMM for decode shapes (M <= 32 per launch), HBM-bound.
//
//   E[M,N] = cde( A[M,K] * ((q[K,N] - zp) * s), D0, D1 )
//
// Same arithmetic and the same pre-packed weight layout as the round-1 kernel (register-resident HMMA: ld.shared ->
// lop3/sub -> A fragment -> mma.sync, group scale on fp32 partial sums; see DESIGN.md for why tcgen05 is not used at
// M <= 64).  What round 2 measured (profiles/r2_micro_pipe_mix2.txt, r2_micro_g2_loop*.txt, r2_g2trace_*.txt) and what
// this kernel does about it:
//   * the unpack + HMMA loop needs 4 warps per SM sub-partition to hide its dependent chains (14-17 clk per warp-word
//     with 16 warps per SM, 23-26 with 8; HBM budget 22.9);
//   * a decode GEMM is so short (1.3 - 3.6 us of HBM time) that the SM is ISSUE-bound on the TOTAL dynamic instruction
//     count of a launch, prologue and epilogue included: under programmatic dependent launch the next GEMM's CTA is
//     co-resident and its prologue competes with this GEMM's math for issue slots.  A version with one private TMA
//     pipeline per warp (48 copies per CTA, ~400 prologue instructions per warp) lost 40 % of the SM to prologues;
//   * so: 16 math warps of <= 64 registers (two CTAs per SM for every variant), CTA-wide ring stages filled by THREE TMA
//     tensor copies each, issued by one lane of a different warp per stage (rotating producer duty), a prologue of a few
//     dozen instructions per warp, one cross-warp reduction per pass.
// Decomposition: n16 channel groups spread evenly over grid x passes slots (<= WN x GW groups each, whole K).  The warps of
// a CTA form a WN x WK grid (M <= 8: 16 warps, 1 x 16 or 2 x 8; M > 8: 8 warps of 128 registers, 1 x 8): in every stage
// warp (wn, wk) takes K unit wk (KBU = 1 or 2 64-k blocks) of its half of the slot's groups.  One ring stage =
//     [activations (WK KBU) x RB x 128 B] [packed int4 G x (WK KBU) x 512 B] [scales rows x 16 G] [zero points rows x 16 G]
// (weights / scales before the PDL wait, activations after).  full[s]: TMA transaction barrier; empty[s]: one arrival per warp.
// Replaces kernel_fpAintB_gemm_wmma / kernel_gemm_xdl_cshuffle_v3 (multi-ABD) for the [EXT] W4A16 row a20; the role of
// gridwise_fpAintB_gemm_wmma.hpp:735-771 (dequant inside the B copy) + gridwise_gemm_pipeline_v1.hpp:558-654.
#pragma once
#include "common.cuh"
#include "dequant.cuh"
#include "epilogue.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace ckb {

constexpr int G2_MAX_RING = 8;            // ring stages per CTA
__host__ __device__ constexpr int g2_warps(int nt) { return nt == 1 ? 16 : 8; }
__host__ __device__ constexpr int g2_threads(int nt) { return 32 * g2_warps(nt); }
__host__ __device__ constexpr int g2_gw_max(int nt) { return nt <= 2 ? 4 : (nt == 4 ? 2 : 1); } // groups per warp

struct G2Args
{
    int M, N, K;
    int kb_total;   // ceil(K / 64)
    int passes;     // slots per CTA; slot = pass * gridDim.x + blockIdx.x
    int slot_q, slot_r; // n16 groups per slot: slots < slot_r hold slot_q + 1, the others slot_q
    int g_hi;       // groups in the larger slots (box of tmW0 / tmS0 / tmZ0; tm*1 hold g_hi - 1)
    int wn_log2;    // WN = 1 << wn_log2 channel-group halves, WK = warps / WN K units per stage
    int upw;        // K units per warp and stage
    int n_chunks;   // stages per pass = ceil(kb_total / (WK upw kbu))
    int ring;       // ring stages
    int stage_bytes, off_w, off_s, off_z;   // one stage: activations at 0, then weights, scales, zero points
    int red_off;    // cross-warp reduction buffer: the (drained) ring when passes == 1, else its own area
    int bar_off;    // mbarriers: full[ring], empty[ring]
    int rb;         // rows of the activation box (>= M; rows past M are zero-filled by TMA)
    int zero_const;
    int has_zero;
    int pdl;        // launch with programmatic stream serialization
    int wf;         // weight format: 0 = packed int4 group-128 (W4A16); 1 = int8 x per-channel bf16 scale (Multiply), 2 = biased uint8 x
                    // fp16 scale (fpAintB), 3 = int8 PassThrough -- K-major int8 streamed as [kb][row][64 B] boxes (W8 instantiations)
    const uint16_t* ch_scale; // W8 formats 1, 2: per-channel scale [N] in A's type
    int ext;        // launch the EXT instantiation (peers or nodep requested; 128-k units, fp32-scale numerics)
    int nodep;      // CKB_FLAG_INPUTS_READY (EXT instantiation): do not wait for the previous grid before reading A; wait before exit
    int epi;        // W4, functor epilogue (D tensors / CDE functor / fp32 E): launch the WF = 4 instantiation, which has g2_store_one inlined
    int rtn;        // CKB_BOP_GROUP_RTN: (q - zp) * s rounded once into A's type before the MMA (RTN kernel instantiation)
    int dbg;        // debug knobs (CKB_GEMV_DBG): 1 = skip the math, 8 = skip the copies and waits
    int peer_rank;  // this rank's index among the peers
    int num_peers;  // fused all-gather: > 0 = store E to e_peer[0..num_peers) and signal flag_peer[*] (ckb.h)
    void* e_peer[CKB_MAX_PEERS];
    unsigned* flag_peer[CKB_MAX_PEERS];
    long long* trace; // trace build: [cta][warp][10] = globaltimer at entry / prefill issued / PDL wait done / first stage full /
                      // K loop done / exit, then clock cycles waiting on full barriers / issuing refills / in the K loop / total
    EpiParams ep;
};

static __device__ __forceinline__ void g2_store_one_body(const EpiParams& ep, int m, int n, float acc)
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
// Out of line for the plain W4 instantiations: the headline path (PassThrough, 16-bit E) never calls it and is 2-5 % slower with the
// code inlined.  Inlined (INL) for the int8-weight instantiations, whose callers usually carry a bias, and for the WF = 4
// instantiations that serve W4 problems with a functor epilogue: out of line, `ep` is a generic pointer into the kernel's parameter
// window and every field read is a cold global-latency load (0.8 - 1.7 us per launch; W8 + bias decode 4-12 % faster inlined,
// profiles/r2_w8_store_inline.txt, r2_gemv_red4.txt; a per-thread local copy of the parameters is slower still, r2_w4_bias_epilogue.txt).
static __device__ __noinline__ void g2_store_one_call(const EpiParams& ep, int m, int n, float acc) { g2_store_one_body(ep, m, n, acc); }
template <bool INL>
static __device__ __forceinline__ void g2_store_one(const EpiParams& ep, int m, int n, float acc)
{
    if(INL) g2_store_one_body(ep, m, n, acc);
    else g2_store_one_call(ep, m, n, acc);
}

// EXT instantiation only: the same element into this rank's slot of every rank's gathered buffer (fused all-gather, ckb.h)
static __device__ __noinline__ void g2_store_peers(const EpiParams& ep, int m, int n, float acc, int num_peers, void* const* e_peer)
{
    float d0 = 0.f, d1 = 0.f;
    if(ep.num_d > 0) d0 = load_d_f32(ep.d_dtype[0], ep.d[0], (long long)m * ep.d_sm[0] + (long long)n * ep.d_sn[0]);
    if(ep.num_d > 1) d1 = load_d_f32(ep.d_dtype[1], ep.d[1], (long long)m * ep.d_sm[1] + (long long)n * ep.d_sn[1]);
    const float v       = cde_apply_f32(ep, acc, d0, d1);
    const long long off = (long long)m * ep.lde + n;
    for(int g = 0; g < num_peers; ++g)
    {
        void* e = e_peer[g];
        if(ep.e_dtype == CKB_BF16) static_cast<uint16_t*>(e)[off] = f32_to_bf16_bits(v, ep.flags);
        else if(ep.e_dtype == CKB_F16) static_cast<__half*>(e)[off] = __float2half_rn(v);
        else static_cast<float*>(e)[off] = v;
    }
}

#ifdef CKB_ENABLE_TRACE
#define G2_DBG(bit) (args.dbg & (bit))
#define G2_T(k) do { if(args.trace && lane == 0) args.trace[(blockIdx.x * 16 + warp) * 10 + (k)] = (long long)ptx::global_timer_ns(); } while(0)
#define G2_CLK(var, stmt) do { const long long _c0 = clock64(); stmt; var += clock64() - _c0; } while(0)
#define G2_TV(k, v) do { if(args.trace && lane == 0) args.trace[(blockIdx.x * 16 + warp) * 10 + (k)] = (v); } while(0)
#else
#define G2_DBG(bit) false
#define G2_T(k) do { } while(0)
#define G2_CLK(var, stmt) do { stmt; } while(0)
#define G2_TV(k, v) do { } while(0)
#endif

// EXT = the instantiation with the optional paths compiled in (peer-store epilogue of the fused all-gather, CKB_FLAG_INPUTS_READY):
// the kernel is so sensitive to its dynamic instruction count that even never-taken branches cost 2-5 % (A/B against the build
// without them, profiles/r2_gemv_ab_ext.txt), so the default instantiation does not contain them.
template <int NT, int GW, int KBU, bool F16, bool ZP, bool RTN, bool EXT, int WF = 0>
__global__ void __launch_bounds__(g2_threads(NT), 2)
wq_gemv2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW0,
                const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmS0,
                const __grid_constant__ CUtensorMap tmS1, const __grid_constant__ CUtensorMap tmZ0,
                const __grid_constant__ CUtensorMap tmZ1, const __grid_constant__ G2Args args)
{
    constexpr uint32_t MAGIC16 = F16 ? 0x6400u : 0x4300u;
    constexpr bool W8 = WF >= 1 && WF <= 3; // WF == 4: the W4 format again, with the functor epilogue inlined (G2Args::epi)
    constexpr int NW  = g2_warps(NT);
    constexpr int NCH = GW == 1 && NT == 1 && !RTN ? 2 : 1; // independent HMMA accumulation chains per group
    // M <= 8: the zero point is folded out of the unpack (magic + q goes to the tensor core, 7 instead of 11 instructions per
    // 8 weights: the SM is issue-bound) and (magic + zp) * sum_k a_k is removed from the fp32 partial sums afterwards
    constexpr bool FOLD = false; // zero-point folding (magic + q into the MMA, correction by an extra ones-MMA): measured slower, profiles/r2_gemv_variants.txt

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem            = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
    const uint32_t bar_base  = smem_base + args.bar_off;
    auto full_bar  = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (G2_MAX_RING + s); };

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int WK   = NW >> args.wn_log2;
    const int wk   = warp & (WK - 1);     // K unit inside a stage
    const int wn   = warp >> (NT == 1 ? 4 - args.wn_log2 : 3 - args.wn_log2); // channel-group half
    constexpr int kbu = KBU;          // 64-k blocks per K unit (1: the group scale is applied per 64 k, 2: per 128 k)
    const int RB   = args.rb;
    const int upw  = args.upw;            // K units per warp and stage
    const int CHK  = WK * upw * kbu;      // 64-k blocks per stage
    const int total = args.passes * args.n_chunks;

    G2_T(0);
    long long wait_clk = 0, issue_clk = 0, clk_entry = clock64(), clk_loop = 0;
    (void)wait_clk; (void)issue_clk; (void)clk_entry; (void)clk_loop;
    if(threadIdx.x == 0)
    {
        for(int s = 0; s < args.ring; ++s)
        {
            ptx::mbar_init(full_bar(s), 1);
            ptx::mbar_init(empty_bar(s), NW);
        }
        ptx::fence_mbar_init();
    }
    // groups of the CTA's slot in pass ps
    auto slot_groups = [&](int ps, int& sb, int& sg) {
        const int slot = ps * gridDim.x + blockIdx.x;
        sb = slot * args.slot_q + min(slot, args.slot_r);
        sg = args.slot_q + (slot < args.slot_r ? 1 : 0);
    };
    // ===================== producer duty: stage j is issued by lane 0 of warp j % NW =====================
    uint64_t pol_stream = 0;
    auto issue_weights = [&](int s, int ps, int c) {
        int sb, sg;
        slot_groups(ps, sb, sg);
        const uint32_t st  = smem_base + s * args.stage_bytes;
        const uint32_t bar = full_bar(s);
        const bool hi      = (sg == args.g_hi);
        const int rows     = (CHK + 1) >> 1; // scale rows per stage
        if(W8)
        {
            // int8 weights, K-major in HBM: one box of [CHK k-blocks][16 sg rows][64 B] (per-channel scales are read once per pass)
            ptx::mbar_arrive_expect_tx(bar, CHK * RB * 128u + sg * CHK * 1024u);
            ptx::tma_load_3d_hint(st + args.off_w, hi ? &tmW0 : &tmW1, bar, 0, sb * 16, c * CHK, pol_stream);
            return;
        }
        uint32_t bytes     = CHK * RB * 128u + sg * (CHK * 512u + rows * 32u);
        if(ZP) bytes += sg * rows * 16u;
        ptx::mbar_arrive_expect_tx(bar, bytes);
        ptx::tma_load_3d_hint(st + args.off_w, hi ? &tmW0 : &tmW1, bar, 0, c * CHK, sb, pol_stream);
        ptx::tma_load_2d(st + args.off_s, hi ? &tmS0 : &tmS1, bar, sb * 16, (c * CHK) >> 1);
        if(ZP) ptx::tma_load_2d(st + args.off_z, hi ? &tmZ0 : &tmZ1, bar, sb * 16, (c * CHK) >> 1);
    };
    auto issue_acts = [&](int s, int c) {
        ptx::tma_load_3d(smem_base + s * args.stage_bytes, &tmA, full_bar(s), 0, 0, c * CHK);
    };
    __syncthreads();
    // Ring prefill.  Weights / scales / zero points do not depend on the previous kernel in the stream and are issued
    // BEFORE the programmatic-dependent-launch wait; activations only after it.
    const int pre     = min(args.ring, total);
    const bool my_pre = lane == 0 && warp < pre && !G2_DBG(8);
    int my_c = 0;
    if(my_pre)
    {
        ptx::prefetch_tmap(&tmW0);
        ptx::prefetch_tmap(&tmA);
        pol_stream      = ptx::policy_evict_first();
        const int my_ps = warp / args.n_chunks;
        my_c            = warp - my_ps * args.n_chunks;
        issue_weights(warp, my_ps, my_c);
    }
    G2_T(1);
    ptx::pdl_launch_dependents();
    if(!(EXT && args.nodep)) ptx::pdl_wait_prior_grids();
    G2_T(2);
    if(my_pre) issue_acts(warp, my_c);

    // ===================== consumer side =====================
    const int g = lane >> 2, t = lane & 3;
    const int tid = threadIdx.x;
    const EpiParams& ep = args.ep;
    // Activation fragment chunks of this warp's K unit, relative to the stage: box rows are [kb][m][64 k], 128B-swizzled for
    // RB >= 4, linear for RB <= 2.  Lanes whose token row does not exist read the last row (a broadcast; the output columns
    // of those tokens are never stored), so the loads need no predicates.
    uint32_t a_addr[KBU][NT][2];
#pragma unroll
    for(int nt = 0; nt < NT; ++nt)
    {
        const int r = min(8 * nt + g, RB - 1);
#pragma unroll
        for(int b = 0; b < KBU; ++b)
        {
            const uint32_t rix = (wk * upw * kbu + b) * RB + r; // first unit; unit uu adds uu * kbu * RB rows (same swizzle phase)
            const uint32_t sw  = RB >= 4 ? (rix & 7) : 0;
            a_addr[b][nt][0]   = rix * 128 + (((2 * t) ^ sw) << 4);
            a_addr[b][nt][1]   = rix * 128 + (((2 * t + 1) ^ sw) << 4);
        }
    }
    const int kb_w       = wk * upw * kbu;                              // this warp's first 64-k block inside a stage
    const uint32_t w_off = args.off_w + kb_w * 512 + lane * 16;         // + gi * (CHK * 512) + (uu * kbu + b) * 512

    int rs = 0;
    uint32_t ph = 0;
    int it = 0;
    for(int ps = 0; ps < args.passes; ++ps)
    {
        int sb, sg;
        slot_groups(ps, sb, sg);
        // this warp's half of the slot's groups (the first half takes the odd group)
        const int first = args.wn_log2 ? (sg + 1) >> 1 : sg;
        const int g0    = wn == 0 ? 0 : first;             // first group inside the slot
        const int G     = wn == 0 ? first : sg - first;    // groups of this warp: GW or GW - 1 (or 0)
        float acc[GW][NT][4];
#pragma unroll
        for(int gi = 0; gi < GW; ++gi)
#pragma unroll
            for(int nt = 0; nt < NT; ++nt)
#pragma unroll
                for(int e = 0; e < 4; ++e) acc[gi][nt][e] = 0.f;
        const uint32_t w_g0 = w_off + g0 * (CHK * 512);
        // W8 formats: the per-channel scales of this lane's two rows (g, g + 8) of each group, packed lo | hi << 16 in A's type
        uint32_t sc8[GW];
        if(WF == 1 || WF == 2)
        {
#pragma unroll
            for(int gi = 0; gi < GW; ++gi)
            {
                const int n_lo = 16 * (sb + g0 + gi) + g;
                const uint32_t lo = (gi < G && n_lo < args.N) ? args.ch_scale[n_lo] : 0u;
                const uint32_t hi = (gi < G && n_lo + 8 < args.N) ? args.ch_scale[n_lo + 8] : 0u;
                sc8[gi] = lo | (hi << 16);
            }
        }
        const uint32_t s_g0 = args.off_s + (g0 * 16 + g) * 2;
        const uint32_t z_g0 = args.off_z + g0 * 16 + g;

        for(int c = 0; c < args.n_chunks; ++c, ++it)
        {
            const uint32_t st = smem_base + rs * args.stage_bytes;
            if(!G2_DBG(8)) G2_CLK(wait_clk, ptx::mbar_wait(full_bar(rs), ph));
            if(it == 0) { G2_T(3); clk_loop = clock64(); }
            if(G > 0 && !G2_DBG(1))
            for(int uu = 0; uu < upw; ++uu)
            {
                // K units past the end of K hold zero-filled activations and weights: skipping them is only a shortcut
                const int kb_u = kb_w + uu * kbu;
                if(c * CHK + kb_u >= args.kb_total) break;
                const uint32_t a_u = st + uu * (kbu * RB * 128);
                const uint32_t w_u = st + w_g0 + uu * (kbu * 512);
                const uint32_t s_u = st + s_g0 + (kb_u >> 1) * (32 * sg);
                const uint32_t z_u = st + z_g0 + (kb_u >> 1) * (16 * sg);
                uint4 bf[KBU][NT][2];
#pragma unroll
                for(int b = 0; b < KBU; ++b)
#pragma unroll
                    for(int nt = 0; nt < NT; ++nt)
                    {
                        bf[b][nt][0] = ptx::lds128(a_u + a_addr[b][nt][0]);
                        bf[b][nt][1] = ptx::lds128(a_u + a_addr[b][nt][1]);
                    }
                float asum0 = 0.f, asum1 = 0.f; // FOLD: sum_k a[k][token 2t], a[k][token 2t + 1] over this unit
                if(FOLD)
                {
                    constexpr uint32_t ONE2 = F16 ? 0x3C003C00u : 0x3F803F80u;
                    const uint32_t ones[4] = {ONE2, ONE2, ONE2, ONE2};
                    float as[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for(int b = 0; b < KBU; ++b)
#pragma unroll
                        for(int ks = 0; ks < 4; ++ks)
                        {
                            const uint4 bb = bf[b][0][ks >> 1];
                            if(ks & 1) ptx::mma_m16n8k16<F16>(as, ones, bb.z, bb.w);
                            else ptx::mma_m16n8k16<F16>(as, ones, bb.x, bb.y);
                        }
                    asum0 = as[0];
                    asum1 = as[1];
                }
                // A warp holds GW or GW - 1 groups: the first GW - 1 run unconditionally with no branches between them (the
                // scheduler overlaps one group's HMMA chain with the next group's unpack); only the last one is guarded.
#pragma unroll
                for(int gi = 0; gi < GW; ++gi)
                {
                    if(GW > 1 && gi == GW - 1 && G < GW) break;
                    if(W8)
                    {
                        // int8 weights: rows g and g + 8 of the group, 16 bytes = the lane's k slice of the 64-k block (k = 16 t + 4 ks + i,
                        // the same permutation the activation fragments use); dequantised with the reference functor's arithmetic and
                        // multiplied straight into the accumulators
                        const uint32_t row_b = st + args.off_w + (uint32_t(g0 + gi) * 16u + uint32_t(g)) * 64u + uint32_t(t) * 16u;
                        const uint32_t kb_st = uint32_t(sg) * 1024u; // bytes of one k-block of the slot's rows
                        const uint32_t s2_lo = (sc8[gi] & 0xFFFFu) * 0x10001u, s2_hi = (sc8[gi] >> 16) * 0x10001u; // fp16 pairs
#pragma unroll
                        for(int b = 0; b < KBU; ++b)
                        {
                            const uint32_t ad  = row_b + uint32_t(kb_u + b) * kb_st;
                            const uint4 lo4    = ptx::lds128(ad), hi4 = ptx::lds128(ad + 512);
                            const uint32_t lw[4] = {lo4.x, lo4.y, lo4.z, lo4.w}, hw[4] = {hi4.x, hi4.y, hi4.z, hi4.w};
#pragma unroll
                            for(int ks = 0; ks < 4; ++ks)
                            {
                                uint32_t ol[2], oh[2];
                                if(WF == 2) { dequant_u8x4_fpaintb(lw[ks], s2_lo, ol); dequant_u8x4_fpaintb(hw[ks], s2_hi, oh); }
                                else if(WF == 1)
                                {
                                    dequant_s8x4_mul_bf16(lw[ks], __uint_as_float(sc8[gi] << 16), ol);
                                    dequant_s8x4_mul_bf16(hw[ks], __uint_as_float(sc8[gi] & 0xFFFF0000u), oh);
                                }
                                else { dequant_s8x4_pass<F16>(lw[ks], ol); dequant_s8x4_pass<F16>(hw[ks], oh); }
                                const uint32_t a[4] = {ol[0], oh[0], ol[1], oh[1]};
#pragma unroll
                                for(int nt = 0; nt < NT; ++nt)
                                {
                                    const uint4 bb = bf[b][nt][ks >> 1];
                                    if(ks & 1) ptx::mma_m16n8k16<F16>(acc[gi][nt], a, bb.z, bb.w);
                                    else ptx::mma_m16n8k16<F16>(acc[gi][nt], a, bb.x, bb.y);
                                }
                            }
                        }
                        (void)s2_lo; (void)s2_hi;
                        continue;
                    }
                    const uint32_t sl = ptx::lds_u16(s_u + gi * 32), sh = ptx::lds_u16(s_u + gi * 32 + 16);
                    uint32_t zm_lo, zm_hi; // magic + zero point of the two channel rows (one 16-bit value; doubled for the exact path)
                    if(ZP)
                    {
                        zm_lo = MAGIC16 + ptx::lds_u8(z_u + gi * 16);
                        zm_hi = MAGIC16 + ptx::lds_u8(z_u + gi * 16 + 8);
                    }
                    else
                        zm_lo = zm_hi = MAGIC16 + uint32_t(args.zero_const);
                    const float s_lo = F16 ? __half2float(__ushort_as_half((unsigned short)sl)) : __uint_as_float(sl << 16);
                    const float s_hi = F16 ? __half2float(__ushort_as_half((unsigned short)sh)) : __uint_as_float(sh << 16);
                    if(!FOLD) { zm_lo *= 0x10001u; zm_hi *= 0x10001u; }
                    // RTN: the weight the MMA sees is (q - zp) * s rounded ONCE (nearest-even) into A's type, the operand a 16-bit
                    // dequantizing GEMM streams (same definition as the fused prefill kernel); products go straight into acc
                    const uint32_t sl2 = sl * 0x10001u, sh2 = sh * 0x10001u;
                    // FOLD: magic + q, else exact (q - zp), in A's type -> fp32 partial sums of this unit (inside one scale group)
                    float part[NCH][NT][4];
#pragma unroll
                    for(int ch = 0; ch < NCH; ++ch)
#pragma unroll
                        for(int nt = 0; nt < NT; ++nt) part[ch][nt][0] = part[ch][nt][1] = part[ch][nt][2] = part[ch][nt][3] = 0.f;
#pragma unroll
                    for(int b = 0; b < KBU; ++b)
                    {
                        const uint4 wv       = ptx::lds128(w_u + gi * (CHK * 512) + b * 512);
                        const uint32_t w4[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                        for(int ks = 0; ks < 4; ++ks)
                        {
                            uint32_t a[4];
                            if(FOLD) dequant_w4_word_magic<F16>(w4[ks], a);
                            else if(RTN) dequant_w4_word_scaled2<F16>(w4[ks], zm_lo, zm_hi, sl2, sh2, a);
                            else dequant_w4_word_exact2<F16>(w4[ks], zm_lo, zm_hi, a);
#pragma unroll
                            for(int nt = 0; nt < NT; ++nt)
                            {
                                const uint4 bb = bf[b][nt][ks >> 1];
                                float* pp      = RTN ? acc[gi][nt] : part[NCH == 2 ? (ks & 1) : 0][nt];
                                if(ks & 1) ptx::mma_m16n8k16<F16>(pp, a, bb.z, bb.w);
                                else ptx::mma_m16n8k16<F16>(pp, a, bb.x, bb.y);
                            }
                        }
                    }
                    if(!RTN)
#pragma unroll
                    for(int nt = 0; nt < NT; ++nt)
                    {
                        float p0 = part[0][nt][0], p1 = part[0][nt][1], p2 = part[0][nt][2], p3 = part[0][nt][3];
                        if(NCH == 2) { p0 += part[NCH - 1][nt][0]; p1 += part[NCH - 1][nt][1]; p2 += part[NCH - 1][nt][2]; p3 += part[NCH - 1][nt][3]; }
                        if(FOLD)
                        {
                            // (magic + zp) is exact in A's type; as fp32: bf16 bits << 16, fp16 by conversion
                            const float zf_lo = F16 ? __half2float(__ushort_as_half((unsigned short)zm_lo)) : __uint_as_float(zm_lo << 16);
                            const float zf_hi = F16 ? __half2float(__ushort_as_half((unsigned short)zm_hi)) : __uint_as_float(zm_hi << 16);
                            p0 = fmaf(-zf_lo, asum0, p0); p1 = fmaf(-zf_lo, asum1, p1);
                            p2 = fmaf(-zf_hi, asum0, p2); p3 = fmaf(-zf_hi, asum1, p3);
                        }
                        acc[gi][nt][0] = fmaf(s_lo, p0, acc[gi][nt][0]);
                        acc[gi][nt][1] = fmaf(s_lo, p1, acc[gi][nt][1]);
                        acc[gi][nt][2] = fmaf(s_hi, p2, acc[gi][nt][2]);
                        acc[gi][nt][3] = fmaf(s_hi, p3, acc[gi][nt][3]);
                    }
                }
            }
            // release the stage; the warp on producer duty for stage it + ring refills it once all warps have released it
            const int nxt = it + args.ring;
            if(nxt < total && !G2_DBG(8))
            {
                __syncwarp();
                if(lane == 0)
                {
                    ptx::mbar_arrive(empty_bar(rs));
                    if(warp == (nxt & (NW - 1)))
                    {
                        G2_CLK(issue_clk,
                               ptx::mbar_wait(empty_bar(rs), ph);
                               if(pol_stream == 0) pol_stream = ptx::policy_evict_first();
                               const int nps = nxt / args.n_chunks; const int nc = nxt - nps * args.n_chunks;
                               issue_weights(rs, nps, nc); issue_acts(rs, nc));
                    }
                }
            }
            if(++rs == args.ring) { rs = 0; ph ^= 1; }
        }

        if(ps == args.passes - 1) { G2_T(4); G2_TV(6, wait_clk); G2_TV(7, issue_clk); G2_TV(8, clock64() - clk_loop); }
        // ---- pass end: the WK warps' partial sums of each half meet in shared memory, then the CDE epilogue ----
        // (single pass: the ring is drained once every warp is here, and the sums are parked in it; otherwise a dedicated buffer)
        float4* red     = reinterpret_cast<float4*>(smem + args.red_off);
        const int n_vec = GW * NT * 32; // vectors per warp (dead groups included: fixed stride)
        __syncthreads();
#pragma unroll
        for(int gi = 0; gi < GW; ++gi)
            if(gi < G)
#pragma unroll
                for(int nt = 0; nt < NT; ++nt)
                    red[warp * n_vec + (gi * NT + nt) * 32 + lane] =
                        make_float4(acc[gi][nt][0], acc[gi][nt][1], acc[gi][nt][2], acc[gi][nt][3]);
        __syncthreads();
        const bool fast = ep.cde_op == CKB_CDE_PASSTHROUGH && (ep.e_dtype == CKB_BF16 || ep.e_dtype == CKB_F16) && !(EXT && args.num_peers > 0);
        // thread -> (half h, vector v): the WK partial sums of a vector are added and stored by one thread
        const int n_out = sg * NT * 32;
        for(int o = tid; o < n_out; o += NW * 32)
        {
            const int l2 = o & 31, q = o >> 5, nt = q % NT, gs = q / NT; // gs = group inside the slot
            const int h  = gs >= first ? 1 : 0;
            const int gi = gs - (h ? first : 0);
            const float4* src = red + (h * WK) * n_vec + (gi * NT + nt) * 32 + l2;
            float4 v = src[0];
            for(int w = 1; w < WK; ++w)
            {
                const float4 p = src[w * n_vec];
                v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
            }
            const int n_lo = 16 * (sb + gs) + (l2 >> 2), n_hi = n_lo + 8;
            const int m0 = 8 * nt + 2 * (l2 & 3), m1 = m0 + 1;
            const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for(int r = 0; r < 4; ++r)
            {
                const int m = (r & 1) ? m1 : m0, n = (r & 2) ? n_hi : n_lo;
                if(m < args.M && n < args.N)
                {
                    if(fast)
                        static_cast<uint16_t*>(ep.e)[(long long)m * ep.lde + n] =
                            ep.e_dtype == CKB_BF16 ? f32_to_bf16_bits(vv[r], ep.flags) : __half_as_ushort(__float2half_rn(vv[r]));
                    else
                        if(EXT && args.num_peers > 0)
                            g2_store_peers(ep, m, n, vv[r], (args.dbg & 32) ? 1 : args.num_peers, (args.dbg & 32) ? &args.e_peer[args.peer_rank] : args.e_peer);
                        else
                            g2_store_one<WF != 0>(ep, m, n, vv[r]);
                }
            }
        }
    }
    // inputs were declared independent of the previous grid: keep completion order along the stream (a later kernel that waits for
    // THIS grid must also be ordered after the one before it)
    if(EXT && args.nodep) ptx::pdl_wait_prior_grids();
    if(EXT && args.num_peers > 0 && !(args.dbg & 128))
    {
        // fused all-gather: every store of this CTA is ordered before its arrival on every rank's counter (CTA barrier, then
        // one thread: system-scope fence + release adds over NVLink); the consumer side is ckb_peer_wait
        // (a system-scope release costs ~2.5 us on this part, tests/peer_probe.py, so only the LAST CTA of the grid pays it: CTAs
        // count themselves on a device-scope counter next to this rank's own flag word; the one that completes the grid
        // publishes one arrival per rank)
        volatile int* s_last_p = reinterpret_cast<volatile int*>(smem); // the ring is drained: its first word carries the verdict
        __syncthreads();
        if(threadIdx.x == 0)
        {
            const unsigned old = ptx::atom_acq_rel_gpu_add_u32(args.flag_peer[args.peer_rank] + 1, 1u);
            *s_last_p = ((old + 1u) % gridDim.x) == 0u;
        }
        __syncthreads();
        if(*s_last_p && threadIdx.x < args.num_peers)
            ptx::red_release_sys_add_u32(args.flag_peer[(args.dbg & 64) ? args.peer_rank : threadIdx.x], 1u);
    }
    G2_T(5);
    G2_TV(9, clock64() - clk_entry);
}


struct G2Plan
{
    CUtensorMap tmA, tmW[2], tmS[2], tmZ[2];
    int32_t nt, gw, passes, slot_q, slot_r, g_hi, wn_log2, kbu, upw, n_chunks, ring, stage_bytes, off_w, off_s, off_z, red_off, bar_off,
        smem_bytes, rb, rtn, wf;
};

// per-(dtype, NT) launchers: each lives in its own translation unit (wq_gemv2_inst.cu, -DGV_INST_F16 / -DGV_INST_NT)
int launch_gemv2_variant(bool f16, int nt, int gw, int kbu, const G2Plan* gp, const G2Args& a, int grid, cudaStream_t stream);

template <int NT, int GW, int KBU, bool F16, bool ZP, bool RTN, bool EXT, int WF = 0>
static int launch_gemv2(const G2Plan* gp, const G2Args& a, int grid, cudaStream_t stream)
{
    auto kfn = wq_gemv2_kernel<NT, GW, KBU, F16, ZP, RTN, EXT, WF>;
    // the opt-in is per device (cudaFuncSetAttribute applies to the current device only): one flag per device ordinal
    static unsigned long long attr_mask = 0;
    int dev = 0;
    CKB_CUDA_CHECK(cudaGetDevice(&dev));
    if(dev >= 64 || !(__atomic_load_n(&attr_mask, __ATOMIC_ACQUIRE) >> dev & 1))
    {
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        // without this the driver sizes the L1/shared carve-out for ONE block and a second CTA never fits
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kfn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        if(dev < 64) __atomic_fetch_or(&attr_mask, 1ull << dev, __ATOMIC_RELEASE);
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim          = dim3(grid);
    cfg.blockDim         = dim3(g2_threads(NT));
    cfg.dynamicSmemBytes = gp->smem_bytes;
    cfg.stream           = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    cfg.numAttrs = a.pdl ? 1 : 0;
    CKB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kfn, gp->tmA, gp->tmW[0], gp->tmW[1], gp->tmS[0], gp->tmS[1], gp->tmZ[0],
                                      gp->tmZ[1], a));
    return CKB_OK;
}

} // namespace ckb
