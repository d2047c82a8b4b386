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
//
// Warp roles (one CTA per SM):
//   warp 0        producer: bulk/TMA copies of weight units and activation boxes into the smem ring
//   warp 1        MMA issuer (one lane): 8 x tcgen05.mma (128 x N_MMA x 16) per unit, one commit
//   warp 2        TMEM allocator
//   warps 4-7     epilogue: per unit tcgen05.ld of the partial accumulator, acc += scale * partial
//   warps 8..     dequant teams (4 warps each, TEAMS teams): team t owns units t, t+TEAMS, ...
// Barriers per unit: full[s] (bytes landed) -> team: wread[s] (smem read), afull[a] (TMEM A written)
//   -> MMA: mma_done[s] (commit) -> producer (slot free), team (A stage free), epilogue (D ready)
//   -> epilogue: dempty[b].
#include "common.cuh"
#include "dequant.cuh"
#include "epilogue.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace ckb {
namespace {

constexpr int DQ_UNIT_N = 128;
constexpr int DQ_UNIT_K = 128;
constexpr int DQ_DBUF   = 4; // TMEM accumulator buffers
constexpr int DQ_TEAMS  = 3; // dequant teams
constexpr int DQ_THREADS = 256 + 128 * DQ_TEAMS;
// Warp numbering matters: the SM's warp arbiter favours HIGH warp ids, so the latency-critical
// single-thread roles (MMA issuer, producer) sit above the ALU-heavy dequant teams.
constexpr int DQ_WARP_DQ0  = 0;
constexpr int DQ_WARP_EPI0 = 4 * DQ_TEAMS;
constexpr int DQ_WARP_MMA  = DQ_WARP_EPI0 + 4;
constexpr int DQ_WARP_PROD = DQ_WARP_EPI0 + 5;
constexpr int DQ_WARP_TMEM = DQ_WARP_EPI0 + 6;

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
    long long* trace; // debug timeline (CKB_TRACE), null in production
    EpiParams ep;
};

__host__ __device__ constexpr int dq_w_bytes(int wf) { return wf == WF_W4_GROUP ? 8192 : 16384; }
__host__ __device__ constexpr int dq_ast(int n_mma) { return n_mma <= 32 ? 6 : 4; }
// two smem rings: weights (released as soon as the dequant team has read them) and activations
// (released when the unit's MMAs have completed).  The weight ring is what keeps HBM busy.
__host__ __device__ constexpr int dq_wst(int wf) { return wf == WF_W4_GROUP ? 16 : 8; }
__host__ __device__ constexpr int dq_actst(int n_mma) { return n_mma <= 16 ? 8 : 4; }
constexpr int DQ_MD = 8; // ring of per-unit "MMAs done" barriers
__host__ __device__ constexpr int dq_smem_bytes(int n_mma, int wf)
{
    return dq_actst(n_mma) * n_mma * 256 + dq_wst(wf) * dq_w_bytes(wf) + 1024 /*barriers*/ + 1024 /*align*/;
}

__device__ __forceinline__ void store_one(const EpiParams& ep, int m, int n, float acc)
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
#define CKB_TR(r, i, k)                                                                   \
    do {                                                                                  \
        if(args.trace && blockIdx.x == 0 && (i) < 64) args.trace[((r)*64 + (i)) * 4 + (k)] = clock64(); \
    } while(0)

// ring position helper: advance by `step` without integer division
struct Ring
{
    int s;
    uint32_t ph;
    __device__ __forceinline__ void advance(int step, int size)
    {
        s += step;
        if(s >= size) { s -= size; ph ^= 1; }
    }
};

template <int N_MMA, int WF, bool F16>
__global__ void __launch_bounds__(DQ_THREADS, 1)
wq_decode_kernel(const __grid_constant__ CUtensorMap tmAct, const __grid_constant__ CUtensorMap tmW,
                 const DQArgs args)
{
    constexpr int ACT_BYTES   = N_MMA * 256; // two boxes of [N_MMA][64 k] 16-bit
    constexpr int W_BYTES     = dq_w_bytes(WF);
    constexpr int WST         = dq_wst(WF);      // weight ring stages
    constexpr int ACT_ST      = dq_actst(N_MMA); // activation ring stages
    constexpr int AST         = dq_ast(N_MMA);   // TMEM A-operand stages (64 columns each)
    constexpr int TMEM_COLS   = 512;
    constexpr int D_COL0      = AST * 64;
    constexpr bool GROUP      = (WF == WF_W4_GROUP);
    static_assert(WST >= DQ_TEAMS && DQ_MD >= AST && DQ_MD % ACT_ST == 0 && AST * 64 + DQ_DBUF * N_MMA <= 512,
                  "ring/TMEM budget");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem            = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t smem_base = ptx::smem_u32(smem);
    const uint32_t act_base  = smem_base;                       // [ACT_ST][ACT_BYTES]
    const uint32_t w_base    = smem_base + ACT_ST * ACT_BYTES;  // [WST][W_BYTES]
    constexpr int BAR_OFF    = ACT_ST * ACT_BYTES + WST * W_BYTES;
    const uint32_t bar_base  = smem_base + BAR_OFF;
    auto wfull_bar   = [&](int s) { return bar_base + 8u * s; };          // weights landed
    auto wread_bar   = [&](int s) { return bar_base + 8u * (16 + s); };   // weights read by the team
    auto actfull_bar = [&](int t) { return bar_base + 8u * (32 + t); };   // activations landed
    auto mdone_bar   = [&](int m) { return bar_base + 8u * (40 + m); };   // unit's MMAs complete
    auto afull_bar   = [&](int a) { return bar_base + 8u * (48 + a); };   // TMEM A stage written
    auto dempty_bar  = [&](int b) { return bar_base + 8u * (56 + b); };   // TMEM D buffer drained
    const uint32_t tmem_slot = bar_base + 8u * 60;
    volatile int* flag_smem  = reinterpret_cast<volatile int*>(smem + BAR_OFF + 8 * 61);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if(warp == DQ_WARP_PROD && lane == 0) ptx::prefetch_tmap(&tmAct);
    if(warp == DQ_WARP_MMA && lane == 0)
    {
        for(int s = 0; s < WST; ++s)
        {
            ptx::mbar_init(wfull_bar(s), 1);
            ptx::mbar_init(wread_bar(s), 4);
        }
        for(int t = 0; t < ACT_ST; ++t) ptx::mbar_init(actfull_bar(t), 1);
        for(int m = 0; m < DQ_MD; ++m) ptx::mbar_init(mdone_bar(m), 1);
        for(int a = 0; a < AST; ++a) ptx::mbar_init(afull_bar(a), 4);
        for(int b = 0; b < DQ_DBUF; ++b) ptx::mbar_init(dempty_bar(b), 4);
        ptx::fence_mbar_init();
    }
    if(warp == DQ_WARP_TMEM)
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
    const int nt0           = int(u_begin / k_units); // first tile / k-unit of this CTA (one division)
    const int kg0           = int(u_begin - (long long)nt0 * k_units);

    if(warp == DQ_WARP_PROD)
    {
        // ===================== producer =====================
        if(lane == 0)
        {
            const uint64_t pol_stream = ptx::policy_evict_first();
            int ntw = nt0, kgw = kg0; // weight cursor
            int nta = nt0, kga = kg0; // activation cursor
            auto adv = [&](int& nt_, int& kg_) { if(++kg_ == k_units) { kg_ = 0; ++nt_; } };
            auto issue_w = [&](int s) {
                ptx::mbar_arrive_expect_tx(wfull_bar(s), W_BYTES);
                if(GROUP)
                    ptx::bulk_load_1d(w_base + s * W_BYTES, args.w_packed + ((size_t)ntw * k_units + kgw) * W_BYTES,
                                      W_BYTES, wfull_bar(s), pol_stream);
                else
                    ptx::tma_load_2d_hint(w_base + s * W_BYTES, &tmW, wfull_bar(s), kgw * DQ_UNIT_K, ntw * DQ_UNIT_N, pol_stream);
                adv(ntw, kgw);
            };
            auto issue_a = [&](int t) {
                const uint32_t st = act_base + t * ACT_BYTES;
                ptx::mbar_arrive_expect_tx(actfull_bar(t), ACT_BYTES);
                ptx::tma_load_2d(st, &tmAct, actfull_bar(t), kga * DQ_UNIT_K, 0);
                ptx::tma_load_2d(st + ACT_BYTES / 2, &tmAct, actfull_bar(t), kga * DQ_UNIT_K + 64, 0);
                adv(nta, kga);
            };
            // weights do not depend on the previous kernel in the stream: fill the weight ring before the
            // programmatic-dependent-launch wait, activations only after it.
            int iw = 0, ia = 0;
            for(; iw < min(WST, n_my); ++iw) issue_w(iw);
            ptx::pdl_wait_prior_grids();
            for(; ia < min(ACT_ST, n_my); ++ia) issue_a(ia);
            Ring rw{0, 0};  // weight slot of unit iw and the parity of its PREVIOUS use's wread phase
            Ring rt{0, 0};  // activation slot of unit ia
            Ring rm{0, 0};  // mdone slot of unit ia - ACT_ST
            // Non-blocking polling of both rings: blocking on one could starve the other (the MMA that
            // frees an activation slot may be waiting for activations this thread has not issued yet).
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
                if(!progress) __nanosleep(32);
            }
        }
    }
    else if(warp == DQ_WARP_MMA)
    {
        // ===================== MMA issuer =====================
        // Warp-uniform loop, one elected lane issues.  afull[a] implies full[s] (the dequant team waited
        // on it before arriving).  Per-iteration overhead (barrier probe ~100 clk, tcgen05 fence ~140 clk)
        // is amortised over every unit whose operands are already there (up to 4 per iteration).
        Ring r{0, 0}, ra{0, 0}, rb{0, 0}, rt{0, 0}; // mdone slot, TMEM A stage, D buffer, activation slot
        int i = 0;
        while(i < n_my)
        {
            ptx::mbar_wait3(afull_bar(ra.s), ra.ph, dempty_bar(rb.s), rb.ph ^ 1, actfull_bar(rt.s), rt.ph);
            int nb = 1;
            if(i + 3 < n_my)
            {
                Ring xa = ra, xb = rb;
                uint32_t bars[6], pars[6];
#pragma unroll
                for(int u = 0; u < 3; ++u)
                {
                    xa.advance(1, AST);
                    xb.advance(1, DQ_DBUF);
                    bars[2 * u] = afull_bar(xa.s);       pars[2 * u] = xa.ph;
                    bars[2 * u + 1] = dempty_bar(xb.s);  pars[2 * u + 1] = xb.ph ^ 1;
                }
                const uint32_t ok = ptx::mbar_test_wait6(bars, pars);
                if((ok & 3) == 3) { nb = 2; if((ok & 12) == 12) { nb = 3; if((ok & 48) == 48) nb = 4; } }
            }
            if(lane == 0) CKB_TR(1, i, 0);
            ptx::tcgen05_fence_after_sync();
            for(int u = 0; u < nb; ++u)
            {
                if(u > 0) ptx::mbar_wait(actfull_bar(rt.s), rt.ph);
                const uint32_t act    = act_base + rt.s * ACT_BYTES;
                const uint32_t d_tmem = tmem_base + D_COL0 + rb.s * N_MMA;
                const uint32_t a_tmem = tmem_base + ra.s * 64;
                const uint64_t db0    = ptx::make_smem_desc_sw128(act, 1024);
                if(ptx::elect_one())
                {
#pragma unroll
                    for(int j = 0; j < DQ_UNIT_K / 16; ++j)
                    {
                        // descriptor start address is in 16-byte units: +32 B per k-step, second box after ACT_BYTES/2
                        const uint64_t db = db0 + uint64_t(((j / 4) * (ACT_BYTES / 2) + (j % 4) * 32) >> 4);
                        ptx::mma_f16_ts(d_tmem, a_tmem + j * 8, db, args.idesc, j != 0);
                    }
                    ptx::mma_commit(mdone_bar(r.s));
                }
                __syncwarp();
                r.advance(1, DQ_MD);
                ra.advance(1, AST);
                rb.advance(1, DQ_DBUF);
                rt.advance(1, ACT_ST);
            }
            if(lane == 0) CKB_TR(1, i, 1);
            i += nb;
        }
    }
    else if(warp < DQ_WARP_EPI0)
    {
        // ===================== dequant teams: smem -> registers -> TMEM (A operand) =====================
        const int q    = warp & 3;                        // TMEM lane quarter
        const int team = (warp - DQ_WARP_DQ0) >> 2;
        const int row  = q * 32 + lane;                   // channel inside the tile
        const bool tr  = (q == 0 && lane == 0);
        float ch_scale_f   = 1.f;
        uint32_t ch_scale2 = 0;
        int cur_nt = -1;
        int nt = nt0, kg = kg0 + team;
        while(kg >= k_units) { kg -= k_units; ++nt; }
        auto load_zero = [&](int nt_, int kg_) -> int {
            return int(__ldg(args.zero + (size_t)kg_ * args.N + min(nt_ * DQ_UNIT_N + row, args.N - 1)));
        };
        int z = args.zero_const;
        if(GROUP && args.zero && team < n_my) z = load_zero(nt, kg);
        Ring rw{team, 0};            // weight ring slot of unit i
        Ring rm{team % DQ_MD, 0};    // mdone ring slot of unit i
        int a_s = team % AST;
        for(int i = team; i < n_my; i += DQ_TEAMS)
        {
            if((WF == WF_W8_MUL || WF == WF_W8_FPAINTB) && nt != cur_nt)
            {
                cur_nt            = nt;
                const int n       = min(nt * DQ_UNIT_N + row, args.N - 1);
                const uint16_t sb = static_cast<const uint16_t*>(args.scale)[n];
                ch_scale_f        = bf16_bits_to_f32(sb);
                ch_scale2         = uint32_t(sb) | (uint32_t(sb) << 16);
            }
            // zero point of this team's NEXT unit: in flight while the current one is processed
            int nt_n = nt, kg_n = kg + DQ_TEAMS;
            while(kg_n >= k_units) { kg_n -= k_units; ++nt_n; }
            int z_next = args.zero_const;
            if(GROUP && args.zero && i + DQ_TEAMS < n_my) z_next = load_zero(nt_n, kg_n);
            // previous user of TMEM A stage a_s is unit i - AST: ring slot (r.s - AST) mod STAGES
            int sp = rm.s - AST;
            uint32_t php = rm.ph;
            if(sp < 0) { sp += DQ_MD; php ^= 1; }
            const bool need_prev = i >= AST;
            if(tr) CKB_TR(2 + team, i, 0);
            ptx::mbar_wait(wfull_bar(rw.s), rw.ph);
            if(tr) CKB_TR(2 + team, i, 1);
            const uint8_t* st    = smem + ACT_ST * ACT_BYTES + rw.s * W_BYTES;
            const uint32_t taddr = tmem_base + a_s * 64 + (uint32_t(q * 32) << 16);
            if(GROUP)
            {
                // tile layout [chunk j = 0..3][row][16 B]: chunk j = k 32j .. 32j+31 of this channel
                uint4 c[4];
#pragma unroll
                for(int j = 0; j < 4; ++j) c[j] = *reinterpret_cast<const uint4*>(st + j * 2048 + row * 16);
                __syncwarp();
                if(lane == 0) ptx::mbar_arrive(wread_bar(rw.s));
                const uint32_t zm = F16 ? (0x6400u + z) * 0x10001u : (0x4300u + z) * 0x10001u;
                if(need_prev) ptx::mbar_wait(mdone_bar(sp), php);
                if(tr) CKB_TR(2 + team, i, 2);
                ptx::tcgen05_fence_after_sync();
#pragma unroll
                for(int j = 0; j < 4; ++j)
                {
                    uint32_t out[16];
                    dequant_w4_word<F16>(c[j].x, zm, out + 0);
                    dequant_w4_word<F16>(c[j].y, zm, out + 4);
                    dequant_w4_word<F16>(c[j].z, zm, out + 8);
                    dequant_w4_word<F16>(c[j].w, zm, out + 12);
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
                if(lane == 0) ptx::mbar_arrive(wread_bar(rw.s));
                if(need_prev) ptx::mbar_wait(mdone_bar(sp), php);
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
            rw.advance(DQ_TEAMS, WST);
            rm.advance(DQ_TEAMS, DQ_MD);
            a_s += DQ_TEAMS % AST;
            if(a_s >= AST) a_s -= AST;
            nt = nt_n; kg = kg_n; z = z_next;
        }
    }
    else if(warp < DQ_WARP_MMA)
    {
        // ===================== epilogue: TMEM -> registers, group scale, stream-K reduction =====================
        ptx::pdl_wait_prior_grids();
        const int q   = warp & 3;
        const int row = q * 32 + lane;
        const bool tr = (q == 0 && lane == 0);
        const EpiParams& ep = args.ep;
        float acc[N_MMA];
#pragma unroll
        for(int j = 0; j < N_MMA; ++j) acc[j] = 0.f;
        int seg_units = 0;
        int nt = nt0, kg = kg0;     // unit being drained
        int ntp = nt0, kgp = kg0;   // unit whose scale is being prefetched (4 ahead)
        Ring r{0, 0}, rb{0, 0};
        auto load_scale = [&](int nt_, int kg_) -> uint16_t {
            const int n_ = min(nt_ * DQ_UNIT_N + row, args.N - 1);
            return __ldg(static_cast<const uint16_t*>(args.scale) + (size_t)kg_ * args.N + n_);
        };
        auto adv = [&](int& nt_, int& kg_) { if(++kg_ == k_units) { kg_ = 0; ++nt_; } };
        // group scales come straight from global memory through a 4-deep register ring
        uint16_t sc_ring[4] = {0, 0, 0, 0};
#pragma unroll
        for(int j = 0; j < 4; ++j)
            if(GROUP && j < n_my) { sc_ring[j] = load_scale(ntp, kgp); adv(ntp, kgp); }

        auto flush = [&]() {
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
                float* ws = args.ws + ((size_t)nt * N_MMA) * DQ_UNIT_N + row;
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
        };
        auto to_scale = [&](uint16_t bits) -> float {
            if(!GROUP) return 1.f;
            return F16 ? __half2float(__ushort_as_half(bits)) : bf16_bits_to_f32(bits);
        };

        for(int i0 = 0; i0 < n_my; i0 += 4)
        {
#pragma unroll
            for(int pr = 0; pr < 2; ++pr)
            {
                const int ia = i0 + 2 * pr;
                if(ia >= n_my) break;
                // two units per barrier wait when both belong to the same tile segment (N_MMA == 16 only)
                const bool pair = (N_MMA == 16) && (ia + 1 < n_my) && (kg + 1 < k_units);
                Ring r2 = r, rb2 = rb;
                r2.advance(1, DQ_MD);
                rb2.advance(1, DQ_DBUF);
                if(tr) CKB_TR(5, ia, 0);
                if(pair) ptx::mbar_wait(mdone_bar(r2.s), r2.ph); // MMAs complete in order: unit ia is done too
                else ptx::mbar_wait(mdone_bar(r.s), r.ph);
                if(tr) CKB_TR(5, ia, 1);
                ptx::tcgen05_fence_after_sync();
                const uint32_t lane_off = uint32_t(q * 32) << 16;
                if(pair)
                {
                    uint32_t v0[16], v1[16];
                    ptx::tmem_ld_x16(tmem_base + D_COL0 + rb.s * N_MMA + lane_off, v0);
                    ptx::tmem_ld_x16(tmem_base + D_COL0 + rb2.s * N_MMA + lane_off, v1);
                    ptx::tmem_ld_wait();
                    if(tr) CKB_TR(5, ia, 2);
                    ptx::tcgen05_fence_before_sync();
                    __syncwarp();
                    if(lane == 0) { ptx::mbar_arrive(dempty_bar(rb.s)); ptx::mbar_arrive(dempty_bar(rb2.s)); }
                    const float s0 = to_scale(sc_ring[2 * pr]), s1 = to_scale(sc_ring[2 * pr + 1]);
#pragma unroll
                    for(int j = 0; j < 16; ++j) acc[j] = fmaf(s0, __uint_as_float(v0[j]), acc[j]);
#pragma unroll
                    for(int j = 0; j < 16; ++j) acc[j] = fmaf(s1, __uint_as_float(v1[j]), acc[j]);
                    seg_units += 2;
                    adv(nt, kg);                       // now at unit ia + 1
                    const bool seg_end = (ia + 2 == n_my) || (kg + 1 == k_units);
                    if(seg_end) flush();
                    adv(nt, kg);
                    r = r2; rb = rb2;
                    r.advance(1, DQ_MD);
                    rb.advance(1, DQ_DBUF);
                }
                else
                {
                    for(int h = 0; h < 2; ++h)
                    {
                        const int i = ia + h;
                        if(i >= n_my) break;
                        if(h == 1)
                        {
                            ptx::mbar_wait(mdone_bar(r.s), r.ph);
                            ptx::tcgen05_fence_after_sync();
                        }
                        const float s0 = to_scale(h == 0 ? sc_ring[2 * pr] : sc_ring[2 * pr + 1]);
#pragma unroll
                        for(int c = 0; c < N_MMA; c += 16)
                        {
                            uint32_t v[16];
                            ptx::tmem_ld_x16(tmem_base + D_COL0 + rb.s * N_MMA + c + lane_off, v);
                            ptx::tmem_ld_wait();
#pragma unroll
                            for(int j = 0; j < 16; ++j) acc[c + j] = fmaf(s0, __uint_as_float(v[j]), acc[c + j]);
                        }
                        ptx::tcgen05_fence_before_sync();
                        __syncwarp();
                        if(lane == 0) ptx::mbar_arrive(dempty_bar(rb.s));
                        ++seg_units;
                        const bool seg_end = (i + 1 == n_my) || (kg + 1 == k_units);
                        if(seg_end) flush();
                        adv(nt, kg);
                        r.advance(1, DQ_MD);
                        rb.advance(1, DQ_DBUF);
                    }
                }
                if(tr) CKB_TR(5, ia, 3);
                // refill the two ring entries just consumed with the scales of units ia+4, ia+5
#pragma unroll
                for(int h = 0; h < 2; ++h)
                    if(GROUP && ia + h + 4 < n_my) { sc_ring[2 * pr + h] = load_scale(ntp, kgp); adv(ntp, kgp); }
            }
        }
    }

    ptx::tcgen05_fence_before_sync();
    __syncthreads();
    if(warp == DQ_WARP_TMEM)
    {
        ptx::tcgen05_fence_after_sync();
        ptx::tmem_dealloc<TMEM_COLS>(tmem_base);
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
    static bool attr_set = false;
    if(!attr_set)
    {
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim          = dim3(plan.grid);
    cfg.blockDim         = dim3(DQ_THREADS);
    cfg.dynamicSmemBytes = smem;
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

int decode_smem(int n_mma, int wf) { return dq_smem_bytes(n_mma, wf); }

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
    plan.grid            = std::max(1, std::min(sms, wp->units_total));
    plan.block           = DQ_THREADS;
    plan.smem_bytes      = decode_smem(n_mma, wf);
    plan.split_units     = k_units;
    plan.workspace_bytes = uint64_t(n_tiles) * n_mma * DQ_UNIT_N * 4 + uint64_t(n_tiles) * 4;
    static const char* wfn[] = {"w4g128", "w8mul", "w8fpaintb", "w8pass"};
    snprintf(plan.kernel_name, sizeof plan.kernel_name, "wq_decode_%s_%s_tmemA_streamk_128x%dx128_t%d",
             wfn[wf], f16 ? "f16" : "bf16", n_mma, DQ_TEAMS);
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
    fill_epi_params(p, a.ep);
    a.trace = nullptr;
    if(const char* tf = getenv("CKB_TRACE"))
    {
        static long long* buf = nullptr;
        if(!buf) cudaMallocManaged(&buf, 6 * 64 * 4 * sizeof(long long));
        cudaDeviceSynchronize();
        memset(buf, 0, 6 * 64 * 4 * sizeof(long long));
        a.trace = buf;
        int rc = wp->n_mma == 16 ? dispatch_decode<16>(wf, f16, wp, a, plan, stream)
                                 : (wp->n_mma == 32 ? dispatch_decode<32>(wf, f16, wp, a, plan, stream)
                                                    : dispatch_decode<64>(wf, f16, wp, a, plan, stream));
        cudaDeviceSynchronize();
        if(FILE* f = fopen(tf, "w"))
        {
            long long t0 = 0;
            for(int i = 0; i < 6 * 64 * 4; ++i) if(buf[i] && (!t0 || buf[i] < t0)) t0 = buf[i];
            const char* names[6] = {"prod", "mma", "dq0", "dq1", "dq2", "epi"};
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
    switch(wp->n_mma)
    {
    case 16: return dispatch_decode<16>(wf, f16, wp, a, plan, stream);
    case 32: return dispatch_decode<32>(wf, f16, wp, a, plan, stream);
    default: return dispatch_decode<64>(wf, f16, wp, a, plan, stream);
    }
}

} // namespace ckb
