// wq_fused.cu -- K1 (round 2): prefill-shaped weight-quantized GEMM (M > 64) as ONE kernel, no workspace.
//
//   E[M][N] = cde( A[M][K] (bf16 | fp16)  x  dequant(Bq)[N][K] ),   fp32 accumulation in TMEM
//
// A cluster of two CTAs on one TPC owns a 256 (M) x 256 (N) output tile (tcgen05.mma.cta_group::2, M = 256).  Per 64-k block
// CTA r of the pair
//   * TMA-loads its 128 rows of A (128B-swizzled, K-major) into the operand stage            (warp 0, one thread),
//   * TMA-loads the PACKED weights of its 128 output channels -- 4 KB of int4 codes + group scales (+ zero points), or 8 KB of
//     int8 -- into a raw ring                                                                 (warp 3, one thread),
//   * eight transform warps unpack, dequantise into A's type with exactly the per-element arithmetic of dequant.cuh (the
//     reference's B-functor definitions: one rounding into A's type) and write the 128B-swizzled K-major operand stage that
//     the UMMA descriptor expects, then fence.proxy.async and arrive on the LEADER CTA's full barrier   (warps 8-15),
//   * the leader's elected thread issues the MMAs for both CTAs and multicasts the commits     (warp 1 of CTA 0),
//   * four epilogue warps drain the CTA's 128 accumulator rows: tcgen05.ld -> CDE functor -> 16-bit -> 128B-swizzled staging
//     in shared memory -> TMA store (cp.async.bulk.tensor ... bulk_group), double buffered per warp    (warps 4-7).
// The quantised weight is read from HBM once per M-band (the raster keeps a band of A in L2), never expanded in memory.
//
// Replaces the B path of kernel_gemm_xdl_cshuffle_v3 (multi-ABD): the B copy with the Multiply functor feeding the block
// pipeline (grid/gridwise_gemm_xdl_cshuffle_v3_multi_abd.hpp:1532-1597), the dequantising threadwise transfer
// (thread/threadwise_tensor_slice_transfer_v3r1_dequant.hpp:509-547) and the fpAintB pipeline
// (grid/gridwise_fpAintB_gemm_wmma.hpp:735-771) for prefill shapes.
#include <algorithm>
#include <atomic>
#include <cstdlib>

#include "common.cuh"
#include "dequant.cuh"
#include "epilogue.cuh"
#include "kernels.h"
#include "ptx.cuh"
#include "tile_map.cuh"

namespace ckb {

namespace {

constexpr int FU_BM         = 256; // pair tile rows (128 per CTA)
constexpr int FU_BN         = 256; // pair tile columns (each CTA dequantises 128 of them)
constexpr int FU_BK         = 64;  // elements of K per stage: one 128-byte swizzle atom of a 16-bit type
#ifndef FU_STAGES_N
#define FU_STAGES_N 4
#endif
#ifndef FU_TG
#define FU_TG 4 // transform groups (k-block interleave of the transform warps)
#endif
constexpr int FU_STAGES     = FU_STAGES_N;
constexpr int FU_HALF       = 128 * 128; // one operand half (128 rows x 128 B) per stage and CTA
constexpr int FU_STAGE_BYTES = 2 * FU_HALF;
constexpr int FU_RAW_STAGES = 4;
constexpr int FU_TW         = 8;   // transform warps
// a transform group waits with parity on barriers it last saw FU_TG k-blocks ago: it must never be a whole ring ahead
static_assert(FU_TG <= FU_STAGES_N && FU_TG <= 4, "transform groups must not outnumber the ring stages (parity waits would alias)");
constexpr int FU_THREADS    = 32 * (8 + FU_TW);
#ifndef FU_EPI_BUFS
#define FU_EPI_BUFS 2
#endif
constexpr int FU_EPI_WARP   = FU_EPI_BUFS * 32 * 128; // 32-row x 64-column 16-bit staging buffers per epilogue warp
constexpr int FU_EPI_BYTES  = 4 * FU_EPI_WARP;
// packed weights of one raw stage: W4 = 2 k-blocks (one 128-k scale group) x 8 channel groups x 512 B, then 128 scales and
// 128 zero points; W8 = one k-block of 128 channels x 64 B
constexpr int FU_W4_SCALE_OFF = 8192, FU_W4_ZERO_OFF = 8192 + 256;
__host__ __device__ constexpr int fu_raw_kb(int wf) { return wf == WF_W4_GROUP ? 2 : 1; }
__host__ __device__ constexpr int fu_raw_bytes(int wf) { return wf == WF_W4_GROUP ? 8192 + 512 : 8192; }
__host__ __device__ constexpr int fu_smem(int wf)
{
    return FU_STAGES * FU_STAGE_BYTES + FU_RAW_STAGES * fu_raw_bytes(wf) + FU_EPI_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
}

struct FUArgs
{
    int M, N, K;
    int num_kb;
    int num_m_blk, num_n_blk;
    int band;
    uint32_t idesc;
    int zero_const, has_zero;
    int tma_store;
    const uint16_t* scale; // W8: per-channel scale in A's type
    EpiParams ep;
};

__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

template <int WF, bool F16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FU_THREADS, 1)
wq_fused_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                const __grid_constant__ CUtensorMap tmS, const __grid_constant__ CUtensorMap tmZ,
                const __grid_constant__ CUtensorMap tmE, const FUArgs args)
{
    constexpr int RKB = fu_raw_kb(WF), RAW_BYTES = fu_raw_bytes(WF);
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t raw_base  = smem_base + FU_STAGES * FU_STAGE_BYTES;
    const uint32_t epi_base  = raw_base + FU_RAW_STAGES * RAW_BYTES; // 1024-aligned: RAW_BYTES * 4 is a multiple of 1024
    const uint32_t bar_base  = epi_base + FU_EPI_BYTES;
    auto full_bar      = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar     = [&](int s) { return bar_base + 8u * (FU_STAGES + s); };
    auto raw_full_bar  = [&](int r) { return bar_base + 8u * (2 * FU_STAGES + r); };
    auto raw_empty_bar = [&](int r) { return bar_base + 8u * (2 * FU_STAGES + FU_RAW_STAGES + r); };
    auto tfull_bar     = [&](int a) { return bar_base + 8u * (2 * FU_STAGES + 2 * FU_RAW_STAGES + a); };
    auto tempty_bar    = [&](int a) { return bar_base + 8u * (2 * FU_STAGES + 2 * FU_RAW_STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * FU_STAGES + 2 * FU_RAW_STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

    if(warp == 0 && lane == 0)
    {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmW);
        if(WF == WF_W4_GROUP) ptx::prefetch_tmap(&tmS);
        if(args.tma_store) ptx::prefetch_tmap(&tmE);
    }
    if(warp == 1 && lane == 0)
    {
        for(int s = 0; s < FU_STAGES; ++s)
        {
            // leader only: its A producer's arrive (+ both CTAs' A bytes) and one arrive per transform warp of BOTH CTAs
            ptx::mbar_init(full_bar(s), 1 + 2 * (FU_TW / FU_TG));
            ptx::mbar_init(empty_bar(s), 1); // one multicast commit
        }
        for(int r = 0; r < FU_RAW_STAGES; ++r)
        {
            ptx::mbar_init(raw_full_bar(r), 1);
            ptx::mbar_init(raw_empty_bar(r), RKB * (FU_TW / FU_TG)); // one arrive per reading warp and k-block of the stage
        }
        for(int a = 0; a < 2; ++a)
        {
            ptx::mbar_init(tfull_bar(a), 1);
            ptx::mbar_init(tempty_bar(a), 8); // leader only: 4 epilogue warps of each CTA
        }
        ptx::fence_mbar_init();
    }
    if(warp == 2)
    {
        ptx::tmem_alloc_pair<512>(tmem_slot);
        ptx::tmem_relinquish_pair();
    }
    ptx::tcgen05_fence_before_sync();
    ptx::cluster_sync(); // barriers of both CTAs are initialised before any remote arrive / multicast commit
    ptx::tcgen05_fence_after_sync();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    const int num_tiles    = args.num_m_blk * args.num_n_blk;
    const int num_kb       = args.num_kb;
    const int raw_per_tile = (num_kb + RKB - 1) / RKB;

    if(warp == 0)
    {
        // ===================== A producer: this CTA's 128 rows, bytes credited to the leader's full barrier =====================
        if(lane == 0)
        {
            int s = 0;
            uint32_t ph = 0;
            for(int t = pair; t < num_tiles; t += num_pairs)
            {
                int mb, nb;
                tile_coord(t, args.num_m_blk, args.num_n_blk, mb, nb, args.band);
                for(int kb = 0; kb < num_kb; ++kb)
                {
                    ptx::mbar_wait(empty_bar(s), ph ^ 1);
                    if(rank == 0) ptx::mbar_arrive_expect_tx(full_bar(s), 2 * FU_HALF);
                    ptx::tma_load_2d_pair(smem_base + s * FU_STAGE_BYTES, &tmA, full_bar(s), kb * FU_BK, mb * FU_BM + int(rank) * 128);
                    if(++s == FU_STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    }
    else if(warp == 3)
    {
        // ===================== packed-weight producer: this CTA's 128 channels into the local raw ring =====================
        if(lane == 0)
        {
            int r = 0;
            uint32_t rph = 0;
            for(int t = pair; t < num_tiles; t += num_pairs)
            {
                int mb, nb;
                tile_coord(t, args.num_m_blk, args.num_n_blk, mb, nb, args.band);
                const int n0 = nb * FU_BN + int(rank) * 128;
                for(int rk = 0; rk < raw_per_tile; ++rk)
                {
                    ptx::mbar_wait(raw_empty_bar(r), rph ^ 1);
                    const uint32_t dst = raw_base + r * RAW_BYTES;
                    if(WF == WF_W4_GROUP)
                    {
                        ptx::mbar_arrive_expect_tx(raw_full_bar(r), 8192 + 256 + (args.has_zero ? 128 : 0));
                        ptx::tma_load_3d(dst, &tmW, raw_full_bar(r), 0, rk * 2, n0 >> 4);
                        ptx::tma_load_2d(dst + FU_W4_SCALE_OFF, &tmS, raw_full_bar(r), n0, rk);
                        if(args.has_zero) ptx::tma_load_2d(dst + FU_W4_ZERO_OFF, &tmZ, raw_full_bar(r), n0, rk);
                    }
                    else
                    {
                        ptx::mbar_arrive_expect_tx(raw_full_bar(r), 8192);
                        ptx::tma_load_2d(dst, &tmW, raw_full_bar(r), rk * FU_BK, n0);
                    }
                    if(++r == FU_RAW_STAGES) { r = 0; rph ^= 1; }
                }
            }
        }
    }
    else if(warp == 1)
    {
        // ===================== MMA issuer (leader CTA only) =====================
        if(rank == 0)
        {
            int s = 0;
            uint32_t ph = 0;
            int as = 0;
            uint32_t aph = 0;
            for(int t = pair; t < num_tiles; t += num_pairs)
            {
                ptx::mbar_wait(tempty_bar(as), aph ^ 1);
                ptx::tcgen05_fence_after_sync();
                const uint32_t d_tmem = tmem_base + as * FU_BN;
                for(int kb = 0; kb < num_kb; ++kb)
                {
                    ptx::mbar_wait(full_bar(s), ph);
                    ptx::tcgen05_fence_after_sync();
                    if(lane == 0)
                    {
                        const uint32_t sa = smem_base + s * FU_STAGE_BYTES;
                        const uint64_t da = ptx::make_smem_desc_sw128(sa, 1024);
                        const uint64_t db = ptx::make_smem_desc_sw128(sa + FU_HALF, 1024);
#pragma unroll
                        for(int j = 0; j < 4; ++j) // +32 bytes of K inside the swizzle atom = +2 in the (addr >> 4) field
                            ptx::mma_f16_ss_pair(d_tmem, da + 2 * j, db + 2 * j, args.idesc, (kb | j) != 0);
                        ptx::mma_commit_pair(empty_bar(s));
                        if(kb == num_kb - 1) ptx::mma_commit_pair(tfull_bar(as));
                    }
                    __syncwarp();
                    if(++s == FU_STAGES) { s = 0; ph ^= 1; }
                }
                if(++as == 2) { as = 0; aph ^= 1; }
            }
        }
    }
    else if(warp >= 8)
    {
        // ===================== transform: raw ring -> dequantised, swizzled K-major operand stage =====================
        // The transform warps form FU_TG groups; group j takes the k-blocks whose running index is j (mod FU_TG), so one warp's
        // chain (wait raw -> LDS -> dequant -> wait empty -> STS -> proxy fence -> remote arrive) spans FU_TG k-blocks of MMA
        // time instead of one (measured: with every warp on every k-block the kernel is bound by that chain's latency).
        constexpr int WPG = FU_TW / FU_TG;
        const int tw = warp - 8, grp = tw / WPG, wig = tw % WPG;
        int s = 0, r = 0, sel = 0;
        uint32_t ph = 0, rph = 0;
        const uint32_t full_leader = ptx::leader_addr(full_bar(0));
        if(WF == WF_W4_GROUP)
        {
            // a warp converts channel groups cg = wig, wig + WPG, ...: lane (g, t4) holds 16 consecutive k of channels g, g + 8
            constexpr int NI = 8 / WPG;
            const int g = lane >> 2, t4 = lane & 3;
            const uint32_t x = uint32_t(g & 7);
            const uint32_t c0 = ((2u * t4) ^ x) << 4, c1 = ((2u * t4 + 1u) ^ x) << 4;
            constexpr uint32_t MAGIC = F16 ? 0x6400u : 0x4300u;
            for(int t = pair; t < num_tiles; t += num_pairs)
            {
                for(int kb = 0; kb < num_kb; ++kb)
                {
                    const int sub = kb & 1;
                    const bool last_of_raw = sub == 1 || kb == num_kb - 1;
                    if(sel == grp)
                    {
                        const uint32_t raw = raw_base + r * RAW_BYTES;
                        ptx::mbar_wait(raw_full_bar(r), rph);
                        uint4 v[NI];
                        uint32_t zm_lo[NI], zm_hi[NI], sc_lo[NI], sc_hi[NI];
#pragma unroll
                        for(int i = 0; i < NI; ++i)
                        {
                            const int cg = wig + WPG * i, row_lo = cg * 16 + g;
                            v[i] = ptx::lds128(raw + uint32_t((cg * 2 + sub) * 512 + lane * 16));
                            const uint32_t sl = ptx::lds_u16(raw + FU_W4_SCALE_OFF + 2 * row_lo);
                            const uint32_t sh = ptx::lds_u16(raw + FU_W4_SCALE_OFF + 2 * (row_lo + 8));
                            const uint32_t zl = args.has_zero ? ptx::lds_u8(raw + FU_W4_ZERO_OFF + row_lo) : uint32_t(args.zero_const);
                            const uint32_t zh = args.has_zero ? ptx::lds_u8(raw + FU_W4_ZERO_OFF + row_lo + 8) : uint32_t(args.zero_const);
                            zm_lo[i] = (MAGIC + zl) * 0x10001u; zm_hi[i] = (MAGIC + zh) * 0x10001u;
                            sc_lo[i] = sl * 0x10001u; sc_hi[i] = sh * 0x10001u;
                        }
                        __syncwarp();
                        if(lane == 0)
                        {
                            ptx::mbar_arrive(raw_empty_bar(r));
                            if(sub == 0 && kb == num_kb - 1) ptx::mbar_arrive(raw_empty_bar(r)); // odd tail: stands in for the absent k-block
                        }
                        ptx::mbar_wait(empty_bar(s), ph ^ 1);
#pragma unroll
                        for(int i = 0; i < NI; ++i)
                        {
                            const uint32_t w4[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
                            uint32_t lo[8], hi[8];
#pragma unroll
                            for(int j = 0; j < 4; ++j)
                            {
                                uint32_t o[4];
                                dequant_w4_word_scaled2<F16>(w4[j], zm_lo[i], zm_hi[i], sc_lo[i], sc_hi[i], o);
                                lo[2 * j] = o[0]; lo[2 * j + 1] = o[2];
                                hi[2 * j] = o[1]; hi[2 * j + 1] = o[3];
                            }
                            const int row_lo   = (wig + WPG * i) * 16 + g;
                            const uint32_t dlo = smem_base + s * FU_STAGE_BYTES + FU_HALF + uint32_t(row_lo) * 128u;
                            sts128(dlo + c0, lo[0], lo[1], lo[2], lo[3]);
                            sts128(dlo + c1, lo[4], lo[5], lo[6], lo[7]);
                            sts128(dlo + 1024u + c0, hi[0], hi[1], hi[2], hi[3]);
                            sts128(dlo + 1024u + c1, hi[4], hi[5], hi[6], hi[7]);
                        }
                        ptx::fence_proxy_async_smem();
                        __syncwarp();
                        if(lane == 0) ptx::mbar_arrive_remote(full_leader + 8u * s);
                    }
                    if(last_of_raw) { if(++r == FU_RAW_STAGES) { r = 0; rph ^= 1; } }
                    if(++s == FU_STAGES) { s = 0; ph ^= 1; }
                    if(++sel == FU_TG) sel = 0;
                }
            }
        }
        else
        {
            // thread gt of the group converts 16 consecutive k (16 raw bytes) of channel row0 + (8 * WPG) * i, i < NI
            constexpr int NI = 16 / WPG;
            const int gt   = wig * 32 + lane;
            const int row0 = gt >> 2, q = gt & 3;
            const uint32_t x = uint32_t(row0 & 7);
            const uint32_t c0 = ((2u * q) ^ x) << 4, c1 = ((2u * q + 1u) ^ x) << 4;
            for(int t = pair; t < num_tiles; t += num_pairs)
            {
                int mb, nb;
                tile_coord(t, args.num_m_blk, args.num_n_blk, mb, nb, args.band);
                uint32_t sc16[NI];
#pragma unroll
                for(int i = 0; i < NI; ++i)
                {
                    const int n = nb * FU_BN + int(rank) * 128 + row0 + 8 * WPG * i;
                    sc16[i]     = (WF != WF_W8_PASS && n < args.N) ? uint32_t(args.scale[n]) : 0u;
                }
                for(int kb = 0; kb < num_kb; ++kb)
                {
                    if(sel == grp)
                    {
                        const uint32_t raw = raw_base + r * RAW_BYTES;
                        ptx::mbar_wait(raw_full_bar(r), rph);
                        uint4 v[NI];
#pragma unroll
                        for(int i = 0; i < NI; ++i) v[i] = ptx::lds128(raw + uint32_t((row0 + 8 * WPG * i) * 64 + q * 16));
                        __syncwarp();
                        if(lane == 0) ptx::mbar_arrive(raw_empty_bar(r));
                        ptx::mbar_wait(empty_bar(s), ph ^ 1);
#pragma unroll
                        for(int i = 0; i < NI; ++i)
                        {
                            const uint32_t w[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
                            uint32_t o[8];
#pragma unroll
                            for(int j = 0; j < 4; ++j)
                            {
                                if(WF == WF_W8_MUL) dequant_s8x4_mul_bf16(w[j], __uint_as_float(sc16[i] << 16), o + 2 * j);
                                else if(WF == WF_W8_FPAINTB) dequant_u8x4_fpaintb(w[j], sc16[i] * 0x10001u, o + 2 * j);
                                else dequant_s8x4_pass<F16>(w[j], o + 2 * j);
                            }
                            const uint32_t d = smem_base + s * FU_STAGE_BYTES + FU_HALF + uint32_t(row0 + 8 * WPG * i) * 128u;
                            sts128(d + c0, o[0], o[1], o[2], o[3]);
                            sts128(d + c1, o[4], o[5], o[6], o[7]);
                        }
                        ptx::fence_proxy_async_smem();
                        __syncwarp();
                        if(lane == 0) ptx::mbar_arrive_remote(full_leader + 8u * s);
                    }
                    if(++r == FU_RAW_STAGES) { r = 0; rph ^= 1; }
                    if(++s == FU_STAGES) { s = 0; ph ^= 1; }
                    if(++sel == FU_TG) sel = 0;
                }
            }
        }
    }
    else if(warp >= 4)
    {
        // ===================== epilogue (both CTAs, own 128 rows) =====================
        const EpiParams& ep = args.ep;
        const int q = warp & 3;
        const uint32_t stg = epi_base + q * FU_EPI_WARP;
        int buf = 0;
        int as = 0;
        uint32_t aph = 0;
        for(int t = pair; t < num_tiles; t += num_pairs)
        {
            int mb, nb;
            tile_coord(t, args.num_m_blk, args.num_n_blk, mb, nb, args.band);
            ptx::mbar_wait(tfull_bar(as), aph);
            ptx::tcgen05_fence_after_sync();
            const int m0         = mb * FU_BM + int(rank) * 128 + q * 32;
            const int m          = m0 + lane;
            const int m_ld       = min(m, ep.M - 1); // D rows are read with a clamped index; rows >= M are never stored
            const uint32_t taddr = tmem_base + as * FU_BN + (uint32_t(q * 32) << 16);
#pragma unroll 1
            for(int c2 = 0; c2 < FU_BN / 64; ++c2)
            {
                const int n0 = nb * FU_BN + c2 * 64;
                if(n0 >= ep.N) break;
                uint32_t pk[32];
#pragma unroll
                for(int h = 0; h < 2; ++h)
                {
                    uint32_t r[32];
                    float v[32];
                    ptx::tmem_ld_x32(taddr + c2 * 64 + h * 32, r);
                    ptx::tmem_ld_wait();
                    if(n0 + h * 32 < ep.N) cde_chunk32_f32(ep, m_ld, n0 + h * 32, r, v);
                    else
                    {
#pragma unroll
                        for(int i = 0; i < 32; ++i) v[i] = 0.f;
                    }
                    if(!args.tma_store) { store_row_f32<32>(ep, m, n0 + h * 32, v); continue; }
#pragma unroll
                    for(int i = 0; i < 16; ++i)
                    {
                        const uint32_t a = ep.e_dtype == CKB_BF16 ? f32_to_bf16_bits(v[2 * i], ep.flags)
                                                                  : __half_as_ushort(__float2half_rn(v[2 * i]));
                        const uint32_t b = ep.e_dtype == CKB_BF16 ? f32_to_bf16_bits(v[2 * i + 1], ep.flags)
                                                                  : __half_as_ushort(__float2half_rn(v[2 * i + 1]));
                        pk[h * 16 + i]   = a | (b << 16);
                    }
                }
                if(args.tma_store)
                {
                    // the store issued from this buffer two chunks ago must have finished READING shared memory
                    if(lane == 0) ptx::tma_store_wait_read<FU_EPI_BUFS - 1>();
                    __syncwarp();
                    const uint32_t row = stg + buf * (32 * 128) + lane * 128;
#pragma unroll
                    for(int c = 0; c < 8; ++c)
                        sts128(row + ((uint32_t(c) ^ uint32_t(lane & 7)) << 4), pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
                    ptx::fence_proxy_async_smem();
                    __syncwarp();
                    if(lane == 0 && m0 < ep.M)
                    {
                        ptx::tma_store_2d(&tmE, stg + buf * (32 * 128), n0, m0); // rows >= M and columns >= N are clipped
                        ptx::tma_store_commit();
                    }
                    if(++buf == FU_EPI_BUFS) buf = 0;
                }
            }
            ptx::tcgen05_fence_before_sync();
            __syncwarp();
            if(lane == 0) ptx::mbar_arrive_remote(ptx::leader_addr(tempty_bar(as)));
            if(++as == 2) { as = 0; aph ^= 1; }
        }
        if(args.tma_store && lane == 0) ptx::tma_store_wait<0>();
    }

    ptx::tcgen05_fence_before_sync();
    ptx::cluster_sync(); // neither CTA may leave (or free TMEM) while the other can still touch its barriers / TMEM
    if(warp == 2)
    {
        ptx::tcgen05_fence_after_sync();
        ptx::tmem_dealloc_pair<512>(tmem_base);
    }
}

struct FUPlan
{
    CUtensorMap tmA, tmW, tmS, tmZ, tmE;
    int32_t wf, tma_store, band;
};

int weight_format(const ckb_gemm_problem& p)
{
    return p.b_op == CKB_BOP_GROUP_RTN ? WF_W4_GROUP : (p.b_op == CKB_BOP_MULTIPLY ? WF_W8_MUL : (p.b_op == CKB_BOP_FPAINTB ? WF_W8_FPAINTB : WF_W8_PASS));
}

template <int WF, bool F16>
int launch_fused(const FUPlan* fp, const FUArgs& a, int grid, cudaStream_t stream)
{
    // opt in to > 48 KB of dynamic shared memory once per device (the attribute is per device, and the call is not free)
    static std::atomic<uint64_t> done{0};
    int dev = 0;
    CKB_CUDA_CHECK(cudaGetDevice(&dev));
    const uint64_t bit = uint64_t(1) << (dev & 63);
    if(!(done.load(std::memory_order_acquire) & bit))
    {
        CKB_CUDA_CHECK(cudaFuncSetAttribute(wq_fused_kernel<WF, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, fu_smem(WF)));
        done.fetch_or(bit, std::memory_order_release);
    }
    wq_fused_kernel<WF, F16><<<grid, FU_THREADS, fu_smem(WF), stream>>>(fp->tmA, fp->tmW, fp->tmS, fp->tmZ, fp->tmE, a);
    CKB_CUDA_CHECK(cudaGetLastError());
    return CKB_OK;
}

} // namespace

int plan_wq_fused(const ckb_gemm_problem& p, ckb_gemm_plan& plan)
{
    static_assert(sizeof(FUPlan) <= sizeof(plan.opaque), "opaque too small");
    FUPlan* fp     = reinterpret_cast<FUPlan*>(plan.opaque);
    const bool f16 = p.a_dtype == CKB_F16;
    const int wf   = weight_format(p);
    fp->wf         = wf;
    const CUtensorMapDataType adt = f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    if(int rc = encode_tmap_2d(&fp->tmA, adt, 2, p.a, p.M, p.K, p.lda, 128, FU_BK, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    if(wf == WF_W4_GROUP)
    {
        const int n_groups = (p.N + 15) / 16, kb_total = (p.K + 63) / 64, kg_total = (p.K + 127) / 128;
        {   // packed weights [n16 group][k64 block][128 words]: box = 8 groups x 2 k-blocks
            const uint64_t dims[3] = {128, uint64_t(kb_total), uint64_t(n_groups)};
            const uint64_t str[2]  = {512, uint64_t(kb_total) * 512};
            const uint32_t box[3]  = {128, 2, 8};
            if(int rc = encode_tmap_nd(&fp->tmW, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, p.b, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        {   // group scales [K/128][N] in A's type: box = one group row x 128 channels
            const uint64_t dims[2] = {uint64_t(p.N), uint64_t(kg_total)};
            const uint64_t str[1]  = {uint64_t(p.N) * 2};
            const uint32_t box[2]  = {128, 1};
            if(int rc = encode_tmap_nd(&fp->tmS, adt, 2, p.scale, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        if(p.zero)
        {
            const uint64_t dims[2] = {uint64_t(p.N), uint64_t(kg_total)};
            const uint64_t str[1]  = {uint64_t(p.N)};
            const uint32_t box[2]  = {128, 1};
            if(int rc = encode_tmap_nd(&fp->tmZ, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, p.zero, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        else
            fp->tmZ = fp->tmS;
    }
    else
    {
        if(int rc = encode_tmap_2d(&fp->tmW, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p.b, p.N, p.K, p.ldb, 128, FU_BK, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        fp->tmS = fp->tmW;
        fp->tmZ = fp->tmW;
    }
    // TMA-store epilogue needs a 16-bit E with a 16-byte-aligned base and row pitch; anything else takes the direct stores
    fp->tma_store = (p.e_dtype == CKB_BF16 || p.e_dtype == CKB_F16) && (reinterpret_cast<uintptr_t>(p.e) & 15) == 0 && (p.lde * 2) % 16 == 0;
    if(const char* e = getenv("CKB_FUSED_TMA_STORE")) fp->tma_store = fp->tma_store && atoi(e) != 0;
    if(fp->tma_store)
    {
        const CUtensorMapDataType edt = p.e_dtype == CKB_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
        if(int rc = encode_tmap_2d(&fp->tmE, edt, 2, p.e, p.M, p.N, p.lde, 32, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    }
    else
        fp->tmE = fp->tmA;
    const int num_m_blk = (p.M + FU_BM - 1) / FU_BM, num_n_blk = (p.N + FU_BN - 1) / FU_BN;
    // raster: bands of M-blocks whose A rows (band x 256 x K x 2 bytes) stay L2-resident while the packed weights stream
    // through once per band; equalised so the last band is not a sliver
    const long long a_blk_bytes = 2LL * FU_BM * p.K;
    int band = int(std::max<long long>(1, (32LL << 20) / a_blk_bytes));
    if(const char* e = getenv("CKB_SS_BAND")) band = std::max(1, atoi(e));
    const int n_bands = (num_m_blk + band - 1) / band;
    fp->band = (num_m_blk + n_bands - 1) / n_bands;
    const int sms = device_info().sm_count > 0 ? device_info().sm_count : 148;
    plan.grid     = 2 * std::max(1, std::min(sms / 2, num_m_blk * num_n_blk));
    plan.block    = FU_THREADS;
    plan.smem_bytes      = fu_smem(wf);
    plan.workspace_bytes = 0;
    plan.split_units     = 1;
    static const char* wfn[] = {"w4g128", "w8mul", "w8fpaintb", "w8pass"};
    snprintf(plan.kernel_name, sizeof plan.kernel_name, "wq_prefill_fused_%s_%s_tcgen05_2cta_256x256x64_s%d_%s", wfn[wf], f16 ? "f16" : "bf16", FU_STAGES,
             fp->tma_store ? "tmastore" : "ststore");
    return CKB_OK;
}

int run_wq_fused(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream)
{
    const FUPlan* fp = reinterpret_cast<const FUPlan*>(plan.opaque);
    const bool f16   = p.a_dtype == CKB_F16;
    FUArgs a{};
    a.M = p.M; a.N = p.N; a.K = p.K;
    a.num_kb    = (p.K + FU_BK - 1) / FU_BK;
    a.num_m_blk = (p.M + FU_BM - 1) / FU_BM;
    a.num_n_blk = (p.N + FU_BN - 1) / FU_BN;
    a.band      = fp->band;
    a.idesc     = ptx::make_idesc(1, f16 ? 0 : 1, f16 ? 0 : 1, FU_BM, FU_BN);
    a.zero_const = p.zero_const;
    a.has_zero   = p.zero != nullptr;
    a.tma_store  = fp->tma_store;
    a.scale      = static_cast<const uint16_t*>(p.scale);
    fill_epi_params(p, a.ep);
    switch(fp->wf)
    {
    case WF_W4_GROUP: return f16 ? launch_fused<WF_W4_GROUP, true>(fp, a, plan.grid, stream) : launch_fused<WF_W4_GROUP, false>(fp, a, plan.grid, stream);
    case WF_W8_MUL: return launch_fused<WF_W8_MUL, false>(fp, a, plan.grid, stream);
    case WF_W8_FPAINTB: return launch_fused<WF_W8_FPAINTB, true>(fp, a, plan.grid, stream);
    default: return f16 ? launch_fused<WF_W8_PASS, true>(fp, a, plan.grid, stream) : launch_fused<WF_W8_PASS, false>(fp, a, plan.grid, stream);
    }
}

} // namespace ckb
