// wq_gemv_kernel.cuh -- K2: W4A16 group-128 GEMM for decode shapes (M <= 64), HBM-bound.
//
//   E[M,N] = cde( A[M,K] * ((q[K,N] - zp) * s), D0, D1 )
//
// Why this kernel does NOT use tcgen05 (measured, tests/micro/{mma_latency,pipe_rates}.cu on B200):
//   * a tcgen05.mma with M=128, K=16 has a 46-cycle issue floor for every N <= 64, so with the weight as the
//     128-row operand one SM consumes at most 2048 weights / 46 clk = 44.5 weights/clk -- 95 % of the HBM rate
//     (6.5 TB/s / 148 SMs / 1.965 GHz = 22.4 B/clk = 44.8 int4 weights/clk) with no slack at all;
//   * mma.sync m16n8k16 issues every 8 clk per SM sub-partition = 128 weights/clk/SM (M <= 8), 64 (M <= 16);
//   * the weight never has to leave registers: ld.shared -> lop3/sub/mul -> A fragment -> HMMA, no TMEM staging,
//     no dequant-team -> MMA-warp -> team barrier chain (which made the previous TMEM-A kernel latency-bound).
// The int4 -> 16-bit conversion itself costs 15 instructions per 8 weights; at the HBM rate that is ~70 % of the
// SM's issue slots, so instruction count per weight, not tensor throughput, is what this kernel is designed around.
//
// Decomposition (no split-K, no cross-CTA reduction, one wave):
//   * channels are handled in groups of 16 (one m16 MMA row block).  The n16-groups are spread evenly over
//     grid * passes "slots"; a CTA runs its `passes` slots one after the other, each over the WHOLE K range.
//   * inside a CTA the 8 consumer warps split K: per 512-k chunk (one ring stage) warp w owns the 64-k block w and
//     loops over the slot's G <= 16/NT groups, so activations are read from shared memory once per warp and
//     the activation fragments are reused for all G groups.  Partial sums of the 8 warps meet in shared memory at
//     the end of the slot (8 KB .. 64 KB), then the CDE epilogue stores E.
//   * producer warp: ONE lane issues 3 TMA tensor copies per stage (weights 3-D box, scales 2-D box, activations
//     3-D box with 128B swizzle; + zero points when present).  Weights / scales do not depend on the previous
//     kernel of the stream and are issued before griddepcontrol.wait (PDL); activations after it.
// Replaces kernel_fpAintB_gemm_wmma / kernel_gemm_xdl_cshuffle_v3 (multi-ABD) for the [EXT] W4A16 row a20.
#pragma once
#include "common.cuh"
#include "dequant.cuh"
#include "epilogue.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace ckb {

constexpr int GV_PAIRS     = 8;                       // warp pairs; pair p owns the 128-k group p of every stage
constexpr int GV_CONSUMERS = 2 * GV_PAIRS;            // consumer warps
constexpr int GV_THREADS   = 32 * GV_CONSUMERS;       // lane 0 of warp 0 doubles as the TMA producer
constexpr int GV_CHUNK_KB  = 2 * GV_PAIRS;            // 64-k blocks per stage (1024 k) for one block per warp
// slots of <= 2 groups give a warp only 8 weight words per 64-k block: those variants take TWO blocks per warp and stage
// (2048 k per stage), halving the per-stage overhead (barrier wait, fragment loads, scale loads, release)
__host__ __device__ constexpr int gv_kbw(int nt, int gw) { return (gw <= 2 && nt <= 2) ? 2 : 1; }
constexpr int GV_MAX_STAGES = 12;

struct GVArgs
{
    int M, N, K;
    int n_groups;   // ceil(N / 16)
    int kb_total;   // ceil(K / 64)
    int passes;     // slots per CTA
    int n_slots;    // gridDim.x * passes
    int g_hi;       // groups in the larger slots (tensor-map box of tmW[0] / tmS[0] / tmZ[0])
    int stages;     // ring depth
    int stage_bytes;
    int off_w, off_s, off_z; // offsets inside a stage (activations at 0)
    int scratch_off;         // partial-sum scratch (after the ring)
    int bar_off;             // mbarriers (after the scratch)
    int zero_const;
    int has_zero;
    int act_rows;   // rows of the activation box (= M)
    int pdl;        // launch with programmatic stream serialization (0: plain launch)
    int dbg;        // debug knobs (CKB_GEMV_DBG): 1 = skip the math, 2 = weights by 1-D bulk copies
    const uint8_t* w;
    long long* gtrace;
    unsigned long long* gspan; // debug (CKB_GSPAN): per-launch [min entry, max exit, min first full, max pdl-wait done]
    EpiParams ep;
};

__host__ __device__ constexpr int gv_gmax(int nt) { return 16 / nt; }

#ifdef CKB_ENABLE_TRACE
#define GV_DBG(bit) (args.dbg & (bit))
#define GV_GT(k)                                                                                          \
    do {                                                                                                  \
        if(args.gtrace) args.gtrace[blockIdx.x * 8 + (k)] = (long long)ptx::global_timer_ns();            \
    } while(0)
// per-chunk timeline of CTA 0 / 74, warps 0 and 7: [after full wait, after the chunk's math] (globaltimer)
#define GV_CT(it, k)                                                                                      \
    do {                                                                                                  \
        if(args.gtrace && (blockIdx.x == 0 || blockIdx.x == 74) && (warp == 0 || warp == 7) && lane == 0 && (it) < 64)   \
            args.gtrace[4096 + (((blockIdx.x != 0) * 2 + (warp != 0)) * 64 + (it)) * 2 + (k)] = (long long)ptx::global_timer_ns(); \
    } while(0)
#define GV_SPAN_MIN(k) do { if(args.gspan) atomicMin(args.gspan + (k), (unsigned long long)ptx::global_timer_ns()); } while(0)
#define GV_SPAN_MAX(k) do { if(args.gspan) atomicMax(args.gspan + (k), (unsigned long long)ptx::global_timer_ns()); } while(0)
#else
#define GV_DBG(bit) false
#define GV_GT(k) do { } while(0)
#define GV_CT(it, k) do { } while(0)
#define GV_SPAN_MIN(k) do { } while(0)
#define GV_SPAN_MAX(k) do { } while(0)
#endif

static __device__ __noinline__ void gv_store_one(const EpiParams& ep, int m, int n, float acc)
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

// small variants are built for two resident CTAs per SM, so that under programmatic dependent launch the NEXT kernel's
// CTAs (prologue, tensor-map fetch, weight prefill) overlap the tail of the current one
template <int NT, int GW, bool F16, bool ZP>
__global__ void __launch_bounds__(GV_THREADS, (NT == 1 && GW <= 3) ? 2 : 1)
wq_gemv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW0,
               const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmS0,
               const __grid_constant__ CUtensorMap tmS1, const __grid_constant__ CUtensorMap tmZ0,
               const __grid_constant__ CUtensorMap tmZ1, const __grid_constant__ GVArgs args)
{
    // GW = groups of 16 channels per slot (= g_hi of the launch; slots with g_hi - 1 groups run one dead group)
    constexpr uint32_t MAGIC16 = F16 ? 0x6400u : 0x4300u;
    constexpr int KBW      = gv_kbw(NT, GW);       // 64-k blocks per warp and stage
    constexpr int CHUNK_KB = GV_CHUNK_KB * KBW;    // 64-k blocks per stage
    constexpr int SROWS    = GV_PAIRS * KBW;       // 128-k scale groups per stage

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint8_t* smem      = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
    const uint32_t bar_base  = smem_base + args.bar_off;
    auto full_bar  = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (GV_MAX_STAGES + s); };

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if(threadIdx.x == 0) GV_GT(0);
    if(threadIdx.x == 0) GV_SPAN_MIN(0);

    if(threadIdx.x == 0)
    {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmW0);
        ptx::prefetch_tmap(&tmS0);
        for(int s = 0; s < args.stages; ++s)
        {
            ptx::mbar_init(full_bar(s), 1);
            ptx::mbar_init(empty_bar(s), GV_CONSUMERS);
        }
        ptx::fence_mbar_init();
    }
    __syncthreads();
    ptx::pdl_launch_dependents();
    if(threadIdx.x == 0) GV_GT(1);
    if(GV_DBG(4)) return; // debug: launch-rate floor (prologue only)

    const int n_chunks = (args.kb_total + CHUNK_KB - 1) / CHUNK_KB;
    const int slot0    = blockIdx.x * args.passes;
    const int RB       = args.act_rows; // activation rows per 64-k block in a stage (= M)
    auto slot_range = [&](int slot, int& gbeg, int& G) {
        gbeg           = int((long long)args.n_groups * slot / args.n_slots);
        const int gend = int((long long)args.n_groups * (slot + 1) / args.n_slots);
        G              = gend - gbeg;
    };

    // ===================== producer state (lane 0 of warp 0 issues all TMA copies) =====================
    // (slot, chunk) cursor: the slot's group range is computed once per slot (64-bit divisions are ~150 clk each)
    struct Cur { int ps, c, gbeg, G; };
    const int total          = args.passes * n_chunks;
    const uint32_t act_bytes = CHUNK_KB * RB * 128;
    const bool is_producer   = (threadIdx.x == 0);
    uint64_t pol_stream      = 0;
    auto cur_next = [&](Cur& q) {
        if(++q.c == n_chunks)
        {
            q.c = 0;
            if(++q.ps < args.passes) slot_range(slot0 + q.ps, q.gbeg, q.G);
        }
    };
    auto issue_weights = [&](int s, const Cur& q) {
        const uint32_t st = smem_base + s * args.stage_bytes;
        const bool hi     = (q.G == args.g_hi);
        uint32_t bytes    = act_bytes + q.G * (CHUNK_KB * 512 + SROWS * 32);
        if(ZP) bytes += SROWS * q.G * 16;
        if(GV_DBG(2))
        {
            const int nkb = min(CHUNK_KB, args.kb_total - q.c * CHUNK_KB);
            bytes -= q.G * (CHUNK_KB - nkb) * 512;
            ptx::mbar_arrive_expect_tx(full_bar(s), bytes);
            for(int gi = 0; gi < q.G; ++gi)
                ptx::bulk_load_1d(st + args.off_w + gi * (CHUNK_KB * 512),
                                  args.w + ((size_t)(q.gbeg + gi) * args.kb_total + q.c * CHUNK_KB) * 512, nkb * 512, full_bar(s), pol_stream);
        }
        else
        {
        ptx::mbar_arrive_expect_tx(full_bar(s), bytes);
        ptx::tma_load_3d_hint(st + args.off_w, hi ? &tmW0 : &tmW1, full_bar(s), 0, q.c * CHUNK_KB, q.gbeg, pol_stream);
        }
        ptx::tma_load_2d(st + args.off_s, hi ? &tmS0 : &tmS1, full_bar(s), q.gbeg * 16, q.c * SROWS);
        if(ZP) ptx::tma_load_2d(st + args.off_z, hi ? &tmZ0 : &tmZ1, full_bar(s), q.gbeg * 16, q.c * SROWS);
    };
    auto issue_acts = [&](int s, int c) {
        ptx::tma_load_3d(smem_base + s * args.stage_bytes, &tmA, full_bar(s), 0, 0, c * CHUNK_KB);
    };
    Cur qw{0, 0, 0, 0};
    int p_issued = 0, p_s = 0;
    uint32_t p_ph = 0; // parity of the empty-barrier phase that frees slot p_s for its next fill
    bool pdl_waited = false;
    {
        // Ring prefill: stage i is issued by lane 0 of warp i (TMA issue costs ~100 ns per copy from one thread, so the
        // first fills go out in parallel).  Weights / scales do not depend on the previous kernel in the stream and are
        // issued before the programmatic-dependent-launch wait, activations only after it.
        const int pre = min(args.stages, total);
        if(lane == 0 && warp < pre && !GV_DBG(8))
        {
            pol_stream = ptx::policy_evict_first();
            Cur q;
            q.ps = warp / n_chunks;
            q.c  = warp - q.ps * n_chunks;
            slot_range(slot0 + q.ps, q.gbeg, q.G);
            issue_weights(warp, q);
            if(warp == 0) GV_GT(2);
            ptx::pdl_wait_prior_grids();
            if(warp == 0) GV_SPAN_MAX(3);
            pdl_waited = true;
            issue_acts(warp, q.c);
        }
        else
        {
            // nothing a consumer needs exists before the activations do: block in griddepcontrol.wait instead of
            // spinning on the first full barrier (a co-resident CTA of the previous kernel needs the issue slots)
            ptx::pdl_wait_prior_grids();
            pdl_waited = true;
        }
        if(is_producer)
        {
            // cursor of the first iteration that the prefill did not cover
            qw.ps = pre / n_chunks;
            qw.c  = pre - qw.ps * n_chunks;
            if(pre < total) slot_range(slot0 + qw.ps, qw.gbeg, qw.G);
            p_issued = pre;
        }
    }
    const int p_defer = args.stages >= 4 ? 1 : 0; // refill a slot one iteration after warp 0 released it: no stall on stragglers
    {
        // ===================== consumers: ld.shared -> dequant -> HMMA =====================
        // Warp w owns the 64-k block w of every stage and runs all G groups of the slot against it: the activation
        // fragments are read from shared memory once per warp (no redundancy across the CTA) and reused G times.
        // All shared-memory offsets that depend only on the thread are computed once, outside the loops.
        const int g = lane >> 2, t = lane & 3;
        const int tid  = threadIdx.x;
        const EpiParams& ep = args.ep;
        uint32_t a_addr[KBW][NT][2]; // activation fragment chunks (128B-swizzled rows [kb][m][64 k]) relative to the stage
        bool a_ok[NT];
#pragma unroll
        for(int nt = 0; nt < NT; ++nt)
        {
            const int r = 8 * nt + g;
            a_ok[nt]    = r < RB;
#pragma unroll
            for(int b = 0; b < KBW; ++b)
            {
                const uint32_t rix = (warp * KBW + b) * RB + r; // row of the stage's activation box
                a_addr[b][nt][0]   = rix * 128 + (((2 * t) ^ (rix & 7)) << 4);
                a_addr[b][nt][1]   = rix * 128 + (((2 * t + 1) ^ (rix & 7)) << 4);
            }
        }
        const uint32_t w_off = args.off_w + warp * KBW * 512 + lane * 16;
        const int srow       = (warp * KBW) >> 1; // this warp's 128-k scale group inside the stage
        uint32_t st   = smem_base;   // current stage
        uint32_t fbar = full_bar(0);
        int it = 0, s = 0;
        uint32_t ph = 0;
        for(int ps = 0; ps < args.passes; ++ps)
        {
            int gbeg, G;
            slot_range(slot0 + ps, gbeg, G);
            const uint32_t s_off = args.off_s + (srow * 16 * G + g) * 2; // this warp's 128-k group row: [SROWS][16 G]
            const uint32_t z_off = args.off_z + srow * 16 * G + g;
            float acc[GW][NT][4];
#pragma unroll
            for(int gi = 0; gi < GW; ++gi)
#pragma unroll
                for(int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for(int e = 0; e < 4; ++e) acc[gi][nt][e] = 0.f;

            for(int c = 0; c < n_chunks; ++c, ++it)
            {
                if(!GV_DBG(8)) ptx::mbar_wait(fbar, ph);
                if(tid == 0 && it == 0) GV_GT(3);
                if(tid == 0 && it == 0) GV_SPAN_MIN(2);
                GV_CT(it, 0);
                if(c * CHUNK_KB + warp * KBW < args.kb_total && !GV_DBG(1))
                {
                    // (a block past the end of K holds zero-filled activations and weights: computing it is harmless)
                    uint4 bf[KBW][NT][2];
#pragma unroll
                    for(int b = 0; b < KBW; ++b)
#pragma unroll
                        for(int nt = 0; nt < NT; ++nt)
                        {
                            if(a_ok[nt])
                            {
                                bf[b][nt][0] = ptx::lds128(st + a_addr[b][nt][0]);
                                bf[b][nt][1] = ptx::lds128(st + a_addr[b][nt][1]);
                            }
                            else
                                bf[b][nt][0] = bf[b][nt][1] = make_uint4(0u, 0u, 0u, 0u);
                        }
                    // All GW groups unconditionally (no branches between groups: the scheduler overlaps one group's
                    // HMMAs with the next group's unpack).  A slot with GW - 1 groups computes one dead group on whatever
                    // the stage holds there; its accumulators are never stored.
#pragma unroll
                    for(int gi = 0; gi < GW; ++gi)
                    {
                        const uint32_t sl = ptx::lds_u16(st + s_off + gi * 32), sh = ptx::lds_u16(st + s_off + gi * 32 + 16);
                        uint32_t zm_lo, zm_hi;
                        if(ZP)
                        {
                            zm_lo = (MAGIC16 + ptx::lds_u8(st + z_off + gi * 16)) * 0x10001u;
                            zm_hi = (MAGIC16 + ptx::lds_u8(st + z_off + gi * 16 + 8)) * 0x10001u;
                        }
                        else
                            zm_lo = zm_hi = (MAGIC16 + uint32_t(args.zero_const)) * 0x10001u;
                        // exact (q - zp) in A's type -> fp32 partial sums of this warp's 64 / 128 k (one scale group)
                        float part[NT][4];
#pragma unroll
                        for(int nt = 0; nt < NT; ++nt) part[nt][0] = part[nt][1] = part[nt][2] = part[nt][3] = 0.f;
#pragma unroll
                        for(int b = 0; b < KBW; ++b)
                        {
                            const uint4 wv       = ptx::lds128(st + w_off + b * 512 + gi * (CHUNK_KB * 512));
                            const uint32_t w4[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                            for(int ks = 0; ks < 4; ++ks)
                            {
                                uint32_t a[4];
                                dequant_w4_word_exact2<F16>(w4[ks], zm_lo, zm_hi, a);
#pragma unroll
                                for(int nt = 0; nt < NT; ++nt)
                                {
                                    const uint4 bb = bf[b][nt][ks >> 1];
                                    if(ks & 1) ptx::mma_m16n8k16<F16>(part[nt], a, bb.z, bb.w);
                                    else ptx::mma_m16n8k16<F16>(part[nt], a, bb.x, bb.y);
                                }
                            }
                        }
                        // apply the (group, channel) scale to the partial sums
                        const float s_lo = F16 ? __half2float(__ushort_as_half((unsigned short)sl)) : __uint_as_float(sl << 16);
                        const float s_hi = F16 ? __half2float(__ushort_as_half((unsigned short)sh)) : __uint_as_float(sh << 16);
#pragma unroll
                        for(int nt = 0; nt < NT; ++nt)
                        {
                            acc[gi][nt][0] = fmaf(s_lo, part[nt][0], acc[gi][nt][0]);
                            acc[gi][nt][1] = fmaf(s_lo, part[nt][1], acc[gi][nt][1]);
                            acc[gi][nt][2] = fmaf(s_hi, part[nt][2], acc[gi][nt][2]);
                            acc[gi][nt][3] = fmaf(s_hi, part[nt][3], acc[gi][nt][3]);
                        }
                    }
                }
                __syncwarp();
                GV_CT(it, 1);
                if(lane == 0) ptx::mbar_arrive(fbar + 8u * GV_MAX_STAGES); // the slot's empty barrier
                st += args.stage_bytes;
                fbar += 8;
                if(++s == args.stages) { s = 0; ph ^= 1; st = smem_base; fbar = full_bar(0); }
                if(is_producer && !GV_DBG(8))
                    while(p_issued < total && p_issued <= it - p_defer + args.stages)
                    {
                        ptx::mbar_wait(empty_bar(p_s), p_ph);
                        issue_weights(p_s, qw);
                        issue_acts(p_s, qw.c);
                        cur_next(qw);
                        ++p_issued;
                        if(++p_s == args.stages) { p_s = 0; p_ph ^= 1; }
                    }
            }

            // ---- slot end: partial sums of the 16 warps meet in shared memory (two rounds of 8), then the CDE epilogue ----
            if(tid == 0 && ps == args.passes - 1) GV_GT(6);
            ptx::named_bar_sync<1>(32 * GV_CONSUMERS); // previous slot's readers are done with the scratch
            float4* scratch = reinterpret_cast<float4*>(const_cast<uint8_t*>(smem) + args.scratch_off);
            const int n_vec = G * NT * 32;
            if(warp >= GV_PAIRS)
            {
#pragma unroll
                for(int gi = 0; gi < GW; ++gi)
                    if(gi < G)
#pragma unroll
                        for(int nt = 0; nt < NT; ++nt)
                            scratch[(warp - GV_PAIRS) * n_vec + (gi * NT + nt) * 32 + lane] =
                                make_float4(acc[gi][nt][0], acc[gi][nt][1], acc[gi][nt][2], acc[gi][nt][3]);
            }
            ptx::named_bar_sync<1>(32 * GV_CONSUMERS);
            if(warp < GV_PAIRS)
            {
#pragma unroll
                for(int gi = 0; gi < GW; ++gi)
                    if(gi < G)
#pragma unroll
                        for(int nt = 0; nt < NT; ++nt)
                        {
                            float4* q      = scratch + warp * n_vec + (gi * NT + nt) * 32 + lane;
                            const float4 o = *q;
                            *q = make_float4(acc[gi][nt][0] + o.x, acc[gi][nt][1] + o.y, acc[gi][nt][2] + o.z, acc[gi][nt][3] + o.w);
                        }
            }
            ptx::named_bar_sync<1>(32 * GV_CONSUMERS);
            if(!pdl_waited) { ptx::pdl_wait_prior_grids(); pdl_waited = true; }
            const bool fast = ep.cde_op == CKB_CDE_PASSTHROUGH && (ep.e_dtype == CKB_BF16 || ep.e_dtype == CKB_F16);
            for(int idx = tid; idx < n_vec; idx += 32 * GV_CONSUMERS)
            {
                float4 v = scratch[idx];
#pragma unroll
                for(int w = 1; w < GV_PAIRS; ++w)
                {
                    const float4 p = scratch[w * n_vec + idx];
                    v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
                }
                const int l2 = idx & 31, nt = (idx >> 5) % NT, gi = (idx >> 5) / NT;
                const int n_lo = 16 * (gbeg + gi) + (l2 >> 2), n_hi = n_lo + 8;
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
                            gv_store_one(ep, m, n, vv[r]);
                    }
                }
            }
        }
    }
    if(threadIdx.x == 0) GV_GT(7);
    if(threadIdx.x == 0) GV_SPAN_MAX(1);
}

struct GVPlan
{
    CUtensorMap tmA, tmW[2], tmS[2], tmZ[2];
    int32_t nt, passes, g_hi, stages, stage_bytes, off_w, off_s, off_z, scratch_off, bar_off, smem_bytes;
};

// per-(dtype, NT) launchers: each lives in its own translation unit (wq_gemv_inst.cu compiled with -DGV_INST_F16 / -DGV_INST_NT)
int launch_gemv_variant(bool f16, int nt, int gw, const GVPlan* gp, const GVArgs& a, int grid, cudaStream_t stream);

template <int NT, int GW, bool F16, bool ZP>
static int launch_gemv(const GVPlan* gp, const GVArgs& a, int grid, cudaStream_t stream)
{
    auto kfn = wq_gemv_kernel<NT, GW, F16, ZP>;
    static PerDeviceOnce attr_once;
    uint64_t dev_bit = 0;
    if(attr_once.need(dev_bit))
    {
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        // without this the driver sizes the L1/shared carve-out for ONE block and a second CTA never fits
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kfn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        attr_once.mark(dev_bit);
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim          = dim3(grid);
    cfg.blockDim         = dim3(GV_THREADS);
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
