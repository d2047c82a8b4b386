// wq_gemv.cu -- K2: W4A16 group-128 GEMM for decode shapes (M <= 64), HBM-bound.
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
#include "common.cuh"
#include "dequant.cuh"
#include "epilogue.cuh"
#include "kernels.h"
#include "ptx.cuh"

#include <vector>

namespace ckb {
namespace {

constexpr int GV_PAIRS     = 8;                       // warp pairs; pair p owns the 128-k group p of every stage
constexpr int GV_CONSUMERS = 2 * GV_PAIRS;            // consumer warps
constexpr int GV_THREADS   = 32 * GV_CONSUMERS;       // lane 0 of warp 0 doubles as the TMA producer
constexpr int GV_CHUNK_KB  = 2 * GV_PAIRS;            // 64-k blocks per stage (1024 k)
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
    int dbg;        // debug knobs (CKB_GEMV_DBG): 1 = skip the math, 2 = weights by 1-D bulk copies
    const uint8_t* w;
    long long* gtrace;
    EpiParams ep;
};

__host__ __device__ constexpr int gv_gmax(int nt) { return 16 / nt; }

#ifdef CKB_ENABLE_TRACE
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
#else
#define GV_GT(k) do { } while(0)
#define GV_CT(it, k) do { } while(0)
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

template <int NT, int GH, bool F16>
__global__ void __launch_bounds__(GV_THREADS, 1)
wq_gemv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW0,
               const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmS0,
               const __grid_constant__ CUtensorMap tmS1, const __grid_constant__ CUtensorMap tmZ0,
               const __grid_constant__ CUtensorMap tmZ1, const __grid_constant__ GVArgs args)
{
    // GH = groups of 16 channels per warp of a pair (compile-time bound: slots of this launch hold <= 2 GH groups)
    constexpr uint32_t MAGIC16 = F16 ? 0x6400u : 0x4300u;

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint8_t* smem      = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
    const uint32_t bar_base  = smem_base + args.bar_off;
    auto full_bar  = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (GV_MAX_STAGES + s); };

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if(threadIdx.x == 0) GV_GT(0);

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

    const int n_chunks = (args.kb_total + GV_CHUNK_KB - 1) / GV_CHUNK_KB;
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
    const uint32_t act_bytes = GV_CHUNK_KB * RB * 128;
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
        uint32_t bytes    = act_bytes + q.G * (GV_CHUNK_KB * 512 + GV_PAIRS * 32);
        if(args.has_zero) bytes += GV_PAIRS * q.G * 16;
        if(args.dbg & 2)
        {
            const int nkb = min(GV_CHUNK_KB, args.kb_total - q.c * GV_CHUNK_KB);
            bytes -= q.G * (GV_CHUNK_KB - nkb) * 512;
            ptx::mbar_arrive_expect_tx(full_bar(s), bytes);
            for(int gi = 0; gi < q.G; ++gi)
                ptx::bulk_load_1d(st + args.off_w + gi * (GV_CHUNK_KB * 512),
                                  args.w + ((size_t)(q.gbeg + gi) * args.kb_total + q.c * GV_CHUNK_KB) * 512, nkb * 512, full_bar(s), pol_stream);
        }
        else
        {
        ptx::mbar_arrive_expect_tx(full_bar(s), bytes);
        ptx::tma_load_3d_hint(st + args.off_w, hi ? &tmW0 : &tmW1, full_bar(s), 0, q.c * GV_CHUNK_KB, q.gbeg, pol_stream);
        }
        ptx::tma_load_2d(st + args.off_s, hi ? &tmS0 : &tmS1, full_bar(s), q.gbeg * 16, q.c * GV_PAIRS);
        if(args.has_zero) ptx::tma_load_2d(st + args.off_z, hi ? &tmZ0 : &tmZ1, full_bar(s), q.gbeg * 16, q.c * GV_PAIRS);
    };
    auto issue_acts = [&](int s, int c) {
        ptx::tma_load_3d(smem_base + s * args.stage_bytes, &tmA, full_bar(s), 0, 0, c * GV_CHUNK_KB);
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
        if(lane == 0 && warp < pre)
        {
            pol_stream = ptx::policy_evict_first();
            Cur q;
            q.ps = warp / n_chunks;
            q.c  = warp - q.ps * n_chunks;
            slot_range(slot0 + q.ps, q.gbeg, q.G);
            issue_weights(warp, q);
            if(warp == 0) GV_GT(2);
            ptx::pdl_wait_prior_grids();
            pdl_waited = true;
            issue_acts(warp, q.c);
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
    const int p_defer = args.stages >= 3 ? 1 : 0; // refill a slot one iteration after warp 0 released it: no stall on stragglers
    {
        // ===================== consumers: ld.shared -> dequant -> HMMA =====================
        // warp pair p owns the 128-k group p of every stage (two 64-k blocks); the pair's warps split the slot's groups
        // of 16 channels (half h takes groups h, h+2, ...).  With 4 warps per sub-partition sharing one tensor pipe
        // (8 clk per HMMA) a warp's own HMMA chain is never the limiter, so one accumulator per (group, token tile).
        const int g = lane >> 2, t = lane & 3;
        const int pair = warp >> 1, half = warp & 1;
        const int tid  = threadIdx.x;
        const EpiParams& ep = args.ep;
        int it = 0, s = 0;
        uint32_t ph = 0;
        for(int ps = 0; ps < args.passes; ++ps)
        {
            int gbeg, G;
            slot_range(slot0 + ps, gbeg, G);
            float acc[GH][NT][4];
#pragma unroll
            for(int li = 0; li < GH; ++li)
#pragma unroll
                for(int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for(int e = 0; e < 4; ++e) acc[li][nt][e] = 0.f;

            for(int c = 0; c < n_chunks; ++c, ++it)
            {
                ptx::mbar_wait(full_bar(s), ph);
                if(tid == 0 && it == 0) GV_GT(3);
                GV_CT(it, 0);
                const int kb0 = c * GV_CHUNK_KB + 2 * pair;
                if(kb0 < args.kb_total && !(args.dbg & 1))
                {
                    const uint8_t* st = smem + s * args.stage_bytes;
                    // this pair's scale / zero-point row of the stage: [pair][16 G]
                    const uint8_t* srow = st + args.off_s + (pair * 16 * G + g) * 2;
                    const uint8_t* zrow = st + args.off_z + pair * 16 * G + g;
                    uint32_t sc_lo[GH], sc_hi[GH], zm_lo[GH], zm_hi[GH];
#pragma unroll
                    for(int li = 0; li < GH; ++li)
                    {
                        const int gi = 2 * li + half;
                        if(gi < G)
                        {
                            sc_lo[li] = uint32_t(*reinterpret_cast<const uint16_t*>(srow + gi * 32)) * 0x10001u;
                            sc_hi[li] = uint32_t(*reinterpret_cast<const uint16_t*>(srow + gi * 32 + 16)) * 0x10001u;
                            if(args.has_zero)
                            {
                                zm_lo[li] = (MAGIC16 + zrow[gi * 16]) * 0x10001u;
                                zm_hi[li] = (MAGIC16 + zrow[gi * 16 + 8]) * 0x10001u;
                            }
                            else
                                zm_lo[li] = zm_hi[li] = (MAGIC16 + uint32_t(args.zero_const)) * 0x10001u;
                        }
                    }
#pragma unroll
                    for(int blk = 0; blk < 2; ++blk)
                    {
                        if(kb0 + blk >= args.kb_total) break;
                        // activation fragments of this 64-k block: token row 8 nt + g, k = 16 t .. 16 t + 15
                        uint4 bf[NT][2];
#pragma unroll
                        for(int nt = 0; nt < NT; ++nt)
                        {
                            const int r = 8 * nt + g;
                            if(r < RB)
                            {
                                const uint32_t rix  = (2 * pair + blk) * RB + r; // row of the stage's [kb][m][128 B] box
                                const uint32_t roff = rix * 128;
                                const uint32_t sw   = rix & 7; // 128B swizzle phase (stages are 1024-byte aligned)
#pragma unroll
                                for(int j = 0; j < 2; ++j)
                                    bf[nt][j] = *reinterpret_cast<const uint4*>(st + roff + (((2 * t + j) ^ sw) << 4));
                            }
                            else
                                bf[nt][0] = bf[nt][1] = make_uint4(0u, 0u, 0u, 0u);
                        }
                        const uint8_t* wrow = st + args.off_w + (2 * pair + blk) * 512 + lane * 16;
#pragma unroll
                        for(int li = 0; li < GH; ++li)
                        {
                            const int gi = 2 * li + half;
                            if(gi < G)
                            {
                                const uint4 wv       = *reinterpret_cast<const uint4*>(wrow + gi * (GV_CHUNK_KB * 512));
                                const uint32_t w4[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                                for(int ks = 0; ks < 4; ++ks)
                                {
                                    uint32_t a[4];
                                    dequant_w4_word_scaled2<F16>(w4[ks], zm_lo[li], zm_hi[li], sc_lo[li], sc_hi[li], a);
#pragma unroll
                                    for(int nt = 0; nt < NT; ++nt)
                                    {
                                        const uint4 b = bf[nt][ks >> 1];
                                        if(ks & 1) ptx::mma_m16n8k16<F16>(acc[li][nt], a, b.z, b.w);
                                        else ptx::mma_m16n8k16<F16>(acc[li][nt], a, b.x, b.y);
                                    }
                                }
                            }
                        }
                    }
                }
                __syncwarp();
                GV_CT(it, 1);
                if(lane == 0) ptx::mbar_arrive(empty_bar(s));
                if(++s == args.stages) { s = 0; ph ^= 1; }
                if(is_producer)
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

            // ---- slot end: partial sums of the 8 warp pairs meet in shared memory, then the CDE epilogue ----
            if(tid == 0 && ps == args.passes - 1) GV_GT(6);
            ptx::named_bar_sync<1>(32 * GV_CONSUMERS); // previous slot's readers are done with the scratch
            float4* scratch = reinterpret_cast<float4*>(const_cast<uint8_t*>(smem) + args.scratch_off);
#pragma unroll
            for(int li = 0; li < GH; ++li)
            {
                const int gi = 2 * li + half;
                if(gi < G)
#pragma unroll
                    for(int nt = 0; nt < NT; ++nt)
                        scratch[((pair * G + gi) * NT + nt) * 32 + lane] =
                            make_float4(acc[li][nt][0], acc[li][nt][1], acc[li][nt][2], acc[li][nt][3]);
            }
            ptx::named_bar_sync<1>(32 * GV_CONSUMERS);
            if(!pdl_waited) { ptx::pdl_wait_prior_grids(); pdl_waited = true; }
            const bool fast = ep.cde_op == CKB_CDE_PASSTHROUGH && (ep.e_dtype == CKB_BF16 || ep.e_dtype == CKB_F16);
            const int n_vec = G * NT * 32;
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
}

struct GVPlan
{
    CUtensorMap tmA, tmW[2], tmS[2], tmZ[2];
    int32_t nt, passes, g_hi, stages, stage_bytes, off_w, off_s, off_z, scratch_off, bar_off, smem_bytes;
};

template <int NT, int GH, bool F16>
int launch_gemv(const GVPlan* gp, const GVArgs& a, int grid, cudaStream_t stream)
{
    auto kfn = wq_gemv_kernel<NT, GH, F16>;
    static bool attr_set = false;
    if(!attr_set)
    {
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set = true;
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
    cfg.numAttrs = 1;
    CKB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kfn, gp->tmA, gp->tmW[0], gp->tmW[1], gp->tmS[0], gp->tmS[1], gp->tmZ[0],
                                      gp->tmZ[1], a));
    return CKB_OK;
}

// stage layout: [activations 16 x M x 128 B][weights G x 8192][scales 8 x 16 G x 2][zero points 8 x 16 G]
struct GVLayout { int off_w, off_s, off_z, stage_bytes, scratch; };
GVLayout gv_layout(int M, int nt, int g, bool zp)
{
    GVLayout l;
    l.off_w       = (GV_CHUNK_KB * M * 128 + 1023) & ~1023;
    l.off_s       = l.off_w + g * GV_CHUNK_KB * 512;
    l.off_z       = l.off_s + GV_PAIRS * g * 32;
    l.stage_bytes = (l.off_z + (zp ? GV_PAIRS * g * 16 : 0) + 1023) & ~1023;
    l.scratch     = GV_PAIRS * g * nt * 512;
    return l;
}

} // namespace

int plan_wq_gemv(const ckb_gemm_problem& p, ckb_gemm_plan& plan)
{
    static_assert(sizeof(GVPlan) <= sizeof(plan.opaque), "opaque too small");
    GVPlan* gp     = reinterpret_cast<GVPlan*>(plan.opaque);
    const bool f16 = p.a_dtype == CKB_F16;
    const bool zp  = p.zero != nullptr;
    const int nt   = p.M <= 8 ? 1 : (p.M <= 16 ? 2 : (p.M <= 32 ? 4 : 8));
    const int n_groups = (p.N + 15) / 16, kb_total = (p.K + 63) / 64, kg_total = (p.K + 127) / 128;
    const int n_chunks = (kb_total + GV_CHUNK_KB - 1) / GV_CHUNK_KB;
    const int sms      = device_info().sm_count > 0 ? device_info().sm_count : 148;
    int grid           = std::min(sms, n_groups);
    if(const char* e = getenv("CKB_GEMV_GRID")) grid = std::max(1, std::min(atoi(e), n_groups));
    int budget = 224 * 1024 - 1024 /*align*/ - 256 /*barriers*/;
    if(const char* e = getenv("CKB_GEMV_SMEM_KB")) budget = std::min(budget, atoi(e) * 1024 - 1280);
    // largest group count per slot whose ring (>= 2 stages, or the whole K range) + scratch fit
    int gmax = gv_gmax(nt);
    while(gmax > 1)
    {
        const GVLayout l = gv_layout(p.M, nt, gmax, zp);
        if(std::min(2, n_chunks) * l.stage_bytes + l.scratch <= budget) break;
        --gmax;
    }
    const int per_cta = (n_groups + grid - 1) / grid;
    const int passes  = (per_cta + gmax - 1) / gmax;
    const int n_slots = grid * passes;
    const int g_hi    = (n_groups + n_slots - 1) / n_slots;
    const int g_lo    = std::max(1, n_groups / n_slots);
    const GVLayout l  = gv_layout(p.M, nt, g_hi, zp);
    gp->off_w = l.off_w; gp->off_s = l.off_s; gp->off_z = l.off_z; gp->stage_bytes = l.stage_bytes;
    gp->stages = std::min({GV_MAX_STAGES, (budget - l.scratch) / l.stage_bytes, passes * n_chunks});
    if(gp->stages < 1) return set_error(CKB_ERR_UNSUPPORTED, "W4 decode kernel: stage does not fit in shared memory");
    gp->scratch_off = gp->stages * gp->stage_bytes;
    gp->bar_off     = gp->scratch_off + l.scratch;
    gp->smem_bytes  = gp->bar_off + 256 + 1024;
    gp->nt = nt; gp->passes = passes; gp->g_hi = g_hi;

    const CUtensorMapDataType adt = f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    {   // activations as [K/64][M][64]: box = 16 k-blocks x M rows x 64 k, 128B swizzle
        const uint64_t dims[3] = {64, uint64_t(p.M), uint64_t(kb_total)};
        const uint64_t str[2]  = {uint64_t(p.lda) * 2, 128};
        const uint32_t box[3]  = {64, uint32_t(p.M), GV_CHUNK_KB};
        if(int rc = encode_tmap_nd(&gp->tmA, adt, 3, p.a, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    }
    for(int i = 0; i < 2; ++i)
    {
        const int gb = i == 0 ? g_hi : g_lo;
        {   // packed weights [n16 group][k64 block][128 words]
            const uint64_t dims[3] = {128, uint64_t(kb_total), uint64_t(n_groups)};
            const uint64_t str[2]  = {512, uint64_t(kb_total) * 512};
            const uint32_t box[3]  = {128, GV_CHUNK_KB, uint32_t(gb)};
            if(int rc = encode_tmap_nd(&gp->tmW[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, p.b, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        {   // group scales [K/128][N]
            const uint64_t dims[2] = {uint64_t(p.N), uint64_t(kg_total)};
            const uint64_t str[1]  = {uint64_t(p.N) * 2};
            const uint32_t box[2]  = {uint32_t(16 * gb), GV_PAIRS};
            if(int rc = encode_tmap_nd(&gp->tmS[i], adt, 2, p.scale, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        if(zp)
        {
            const uint64_t dims[2] = {uint64_t(p.N), uint64_t(kg_total)};
            const uint64_t str[1]  = {uint64_t(p.N)};
            const uint32_t box[2]  = {uint32_t(16 * gb), GV_PAIRS};
            if(int rc = encode_tmap_nd(&gp->tmZ[i], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, p.zero, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        else
            gp->tmZ[i] = gp->tmS[i];
    }
    plan.grid            = grid;
    plan.block           = GV_THREADS;
    plan.smem_bytes      = gp->smem_bytes;
    plan.split_units     = 1;
    plan.workspace_bytes = 0;
    snprintf(plan.kernel_name, sizeof plan.kernel_name, "wq_gemv_w4g128_%s_hmma_nsplit_g%d_p%d_m%d_s%d", f16 ? "f16" : "bf16",
             g_hi, passes, 8 * nt, gp->stages);
    return CKB_OK;
}

int run_wq_gemv(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream)
{
    const GVPlan* gp = reinterpret_cast<const GVPlan*>(plan.opaque);
    GVArgs a{};
    a.M = p.M; a.N = p.N; a.K = p.K;
    a.n_groups = (p.N + 15) / 16;
    a.kb_total = (p.K + 63) / 64;
    a.passes   = gp->passes;
    a.n_slots  = plan.grid * gp->passes;
    a.g_hi     = gp->g_hi;
    a.stages   = gp->stages;
    a.stage_bytes = gp->stage_bytes;
    a.off_w = gp->off_w; a.off_s = gp->off_s; a.off_z = gp->off_z;
    a.scratch_off = gp->scratch_off;
    a.bar_off     = gp->bar_off;
    a.zero_const  = p.zero_const;
    a.has_zero    = p.zero != nullptr;
    a.act_rows    = p.M;
    a.dbg         = getenv("CKB_GEMV_DBG") ? atoi(getenv("CKB_GEMV_DBG")) : 0;
    a.w           = static_cast<const uint8_t*>(p.b);
    a.gtrace      = nullptr;
    fill_epi_params(p, a.ep);
    const bool f16 = p.a_dtype == CKB_F16;
    // kernel variant: NT token tiles x GH groups per warp (2 GH >= g_hi, 2 GH NT <= 16)
    const int gh = gp->g_hi <= 2 ? 1 : (gp->g_hi <= 4 ? 2 : (gp->g_hi <= 8 ? 4 : 8));
    auto go = [&]() -> int {
#define GV_CASE(NT_, GH_)                                                                                      \
    if(gp->nt == NT_ && gh == GH_)                                                                             \
        return f16 ? launch_gemv<NT_, GH_, true>(gp, a, plan.grid, stream) : launch_gemv<NT_, GH_, false>(gp, a, plan.grid, stream);
        GV_CASE(1, 1) GV_CASE(1, 2) GV_CASE(1, 4) GV_CASE(1, 8)
        GV_CASE(2, 1) GV_CASE(2, 2) GV_CASE(2, 4)
        GV_CASE(4, 1) GV_CASE(4, 2)
        GV_CASE(8, 1)
#undef GV_CASE
        return set_error(CKB_ERR_INVALID, "no W4 decode kernel variant for nt %d g_hi %d", gp->nt, gp->g_hi);
    };
#ifdef CKB_ENABLE_TRACE
    if(const char* gf = getenv("CKB_GTRACE"))
    {
        static long long* gbuf = nullptr;
        const int NW = 8 * 1024;
        if(!gbuf) cudaMalloc(&gbuf, NW * sizeof(long long));
        cudaDeviceSynchronize();
        cudaMemset(gbuf, 0, NW * sizeof(long long));
        cudaDeviceSynchronize();
        a.gtrace = gbuf;
        const int rc = go();
        cudaDeviceSynchronize();
        std::vector<long long> hb(NW);
        cudaMemcpy(hb.data(), gbuf, NW * sizeof(long long), cudaMemcpyDeviceToHost);
        if(FILE* f = fopen(gf, "a"))
        {
            long long t0 = 0;
            for(int i = 0; i < NW; ++i) if(hb[i] && (!t0 || hb[i] < t0)) t0 = hb[i];
            fprintf(f, "# M%d N%d K%d grid %d\n", p.M, p.N, p.K, plan.grid);
            for(int b = 0; b < plan.grid && b < 512; ++b)
            {
                fprintf(f, "%d", b);
                for(int k = 0; k < 8; ++k) fprintf(f, " %lld", hb[b * 8 + k] ? hb[b * 8 + k] - t0 : -1);
                fprintf(f, "\n");
            }
            for(int w = 0; w < 4; ++w)
            {
                fprintf(f, "# chunks cta%d warp%d:", (w >> 1) ? 74 : 0, (w & 1) ? 7 : 0);
                for(int i = 0; i < 64; ++i)
                {
                    const long long a0 = hb[4096 + (w * 64 + i) * 2], a1 = hb[4096 + (w * 64 + i) * 2 + 1];
                    if(a0) fprintf(f, " [%lld %lld]", a0 - t0, a1 - t0);
                }
                fprintf(f, "\n");
            }
            fclose(f);
        }
        return rc;
    }
#endif
    return go();
}

} // namespace ckb
