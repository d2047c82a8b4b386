// gemm_ss.cu -- K3: both operands streamed by TMA into a 128B-swizzled smem ring and consumed by
// tcgen05.mma straight from shared memory (SS form).  kind::i8 (s8 x s8 -> s32 in TMEM) serves
// CK's int8 quantized GEMM (example/14_gemm_quantization, DeviceGemmMultipleD_Xdl_CShuffle) and
// the int8 MultiplyMultiply composition; kind::f16 (bf16/fp16 -> fp32) is the dense comparator.
//
// Replaces: kernel_gemm_multiple_d_xdl_cshuffle (grid/gridwise_gemm_multiple_d_xdl_cshuffle.hpp:77,499)
//           and the multi-D flavour of kernel_gemm_xdl_cshuffle_v3 (grid/gridwise_gemm_xdl_cshuffle_v3_multi_d.hpp).
//
// CTA = 8 warps, persistent over output tiles of 128 (M) x 256 (N):
//   warp 0    TMA producer          (empty[s] -> full[s])
//   warp 1    tcgen05.mma issuer    (full[s] -> empty[s]; tmem_empty[a] -> tmem_full[a])
//   warp 2    TMEM allocator
//   warp 4-7  epilogue              (tmem_full[a] -> tmem_empty[a]); warp w owns TMEM lanes 32*(w%4)..
// TMEM: two 256-column fp32/s32 accumulators (all 512 columns) so the epilogue of tile i overlaps
// the main loop of tile i+1.
#include "common.cuh"
#include "epilogue.cuh"
#include "kernels.h"
#include "ptx.cuh"
#include "tile_map.cuh"

namespace ckb {

namespace {

constexpr int SS_BM       = 128;
constexpr int SS_BN       = 256;
constexpr int SS_BK_BYTES = 128; // one swizzle atom of K per stage row
constexpr int SS_STAGES   = 4;
constexpr int SS_A_BYTES  = SS_BM * SS_BK_BYTES;
constexpr int SS_B_BYTES  = SS_BN * SS_BK_BYTES;
constexpr int SS_STAGE_BYTES = SS_A_BYTES + SS_B_BYTES;
constexpr int SS_THREADS  = 256;
constexpr int SS_SMEM     = SS_STAGES * SS_STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

struct SSArgs
{
    int M, N;
    int num_kb;     // K blocks of 128 bytes
    int kb_elems;   // elements per K block (128 for i8, 64 for 16-bit)
    int num_m_blk, num_n_blk;
    uint32_t idesc;
    int is_i8;
    int band;       // M-blocks per raster band (tile_coord)
    EpiParams ep;
};

// One accumulator row chunk (32 columns) through the CDE functor.  The functor is selected ONCE per chunk so the
// element loops are branch-free; the int32 bias of Add_Activation_* is added in integer arithmetic as the reference does.
__device__ __forceinline__ void
ss_epilogue_chunk(const SSArgs& a, int m, int n0, const uint32_t* r)
{
    const EpiParams& ep = a.ep;
    if(m >= ep.M || n0 >= ep.N) return;
    float v[32], d0[32], d1[32];
    (void)d0;
    if(a.is_i8)
    {
        if(ep.cde_op == CKB_CDE_PASSTHROUGH && (ep.e_dtype == CKB_I32 || ep.e_dtype == CKB_I8))
        {
            const long long base = (long long)m * ep.lde + n0;
            if(ep.e_dtype == CKB_I32)
            {
                int32_t* e = static_cast<int32_t*>(ep.e) + base;
                if(n0 + 32 <= ep.N && ((reinterpret_cast<uintptr_t>(e) & 15) == 0))
                {
#pragma unroll
                    for(int i = 0; i < 32; i += 4)
                        *reinterpret_cast<uint4*>(e + i) = make_uint4(r[i], r[i + 1], r[i + 2], r[i + 3]);
                }
                else
                {
#pragma unroll
                    for(int i = 0; i < 32; ++i)
                        if(n0 + i < ep.N) e[i] = int32_t(r[i]);
                }
            }
            else
            {
                int8_t* e = static_cast<int8_t*>(ep.e) + base;
#pragma unroll
                for(int i = 0; i < 32; ++i)
                    if(n0 + i < ep.N) e[i] = static_cast<int8_t>(int32_t(r[i]));
            }
            return;
        }
        const bool relu = ep.act == CKB_ACT_RELU;
        auto act = [&](float x) { return relu ? fmaxf(x, 0.f) : x; };
        auto clampq = [](float x) { return fminf(fmaxf(x, -128.f), 127.f); };
        switch(ep.cde_op)
        {
        case CKB_CDE_ACT_MUL_CLAMP:
#pragma unroll
            for(int i = 0; i < 32; ++i) v[i] = clampq(__fmul_rn(ep.scale, act(float(int32_t(r[i])))));
            break;
        case CKB_CDE_ADD_ACT_MUL_CLAMP:
        case CKB_CDE_ADD_ACT_MUL2_CLAMP: {
            // bias is int32 and is added BEFORE the float conversion (quantization_operation.hpp:153-160)
            const int32_t* bp = static_cast<const int32_t*>(ep.d[0]);
            if(ep.cde_op == CKB_CDE_ADD_ACT_MUL2_CLAMP) load_d_chunk32(ep, 1, m, n0, d1);
#pragma unroll
            for(int i = 0; i < 32; ++i)
            {
                const int n     = min(n0 + i, ep.N - 1);
                const int32_t b = bp[(long long)m * ep.d_sm[0] + (long long)n * ep.d_sn[0]];
                const float s   = ep.cde_op == CKB_CDE_ADD_ACT_MUL2_CLAMP ? d1[i] : ep.scale;
                v[i]            = clampq(__fmul_rn(s, act(float(int32_t(r[i]) + b))));
            }
            break;
        }
        case CKB_CDE_ACT_MUL2_CLAMP:
            load_d_chunk32(ep, 0, m, n0, d0);
#pragma unroll
            for(int i = 0; i < 32; ++i) v[i] = clampq(__fmul_rn(d0[i], act(float(int32_t(r[i])))));
            break;
        case CKB_CDE_MULTIPLY_MULTIPLY:
            load_d_chunk32(ep, 0, m, n0, d0);
            load_d_chunk32(ep, 1, m, n0, d1);
#pragma unroll
            for(int i = 0; i < 32; ++i) v[i] = __fmul_rn(__fmul_rn(float(int32_t(r[i])), d0[i]), d1[i]);
            break;
        default:
#pragma unroll
            for(int i = 0; i < 32; ++i) v[i] = float(int32_t(r[i]));
        }
        store_row_f32<32>(ep, m, n0, v);
        return;
    }
    cde_chunk32_f32(ep, m, n0, r, v);
    store_row_f32<32>(ep, m, n0, v);
}

__global__ void __launch_bounds__(SS_THREADS, 1)
gemm_ss_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const SSArgs args)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base  = smem_base + SS_STAGES * SS_STAGE_BYTES;
    auto full_bar   = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar  = [&](int s) { return bar_base + 8u * (SS_STAGES + s); };
    auto tfull_bar  = [&](int a) { return bar_base + 8u * (2 * SS_STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * SS_STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * SS_STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if(warp == 0 && lane == 0)
    {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
    }
    if(warp == 1 && lane == 0)
    {
        for(int s = 0; s < SS_STAGES; ++s)
        {
            ptx::mbar_init(full_bar(s), 1);
            ptx::mbar_init(empty_bar(s), 1);
        }
        for(int a = 0; a < 2; ++a)
        {
            ptx::mbar_init(tfull_bar(a), 1);
            ptx::mbar_init(tempty_bar(a), 4);
        }
        ptx::fence_mbar_init();
    }
    if(warp == 2)
    {
        ptx::tmem_alloc<512>(tmem_slot);
        ptx::tmem_relinquish();
    }
    ptx::tcgen05_fence_before_sync();
    __syncthreads();
    ptx::tcgen05_fence_after_sync();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    const int num_tiles = args.num_m_blk * args.num_n_blk;
    const int num_kb    = args.num_kb;

    if(warp == 0 || warp == 3)
    {
        // ===================== TMA producers =====================
        // two issuing threads: warp 0 arms the stage and loads A, warp 3 loads B.  One thread needs ~100 ns per TMA
        // issue, and a k-block is only ~300 ns of MMA work, so a single producer sits right at the issue limit.
        // (B's bytes may land before A's thread has armed the barrier: its pending arrival keeps the phase open.)
        if(lane == 0)
        {
            const bool is_a = warp == 0;
            int s = 0;
            uint32_t ph = 0;
            for(int t = blockIdx.x; t < num_tiles; t += gridDim.x)
            {
                int mb, nb;
                tile_coord(t, args.num_m_blk, args.num_n_blk, mb, nb, args.band);
                for(int kb = 0; kb < num_kb; ++kb)
                {
                    ptx::mbar_wait(empty_bar(s), ph ^ 1);
                    const uint32_t sa = smem_base + s * SS_STAGE_BYTES;
                    if(is_a)
                    {
                        ptx::mbar_arrive_expect_tx(full_bar(s), SS_STAGE_BYTES);
                        ptx::tma_load_2d(sa, &tmA, full_bar(s), kb * args.kb_elems, mb * SS_BM);
                    }
                    else
                        ptx::tma_load_2d(sa + SS_A_BYTES, &tmB, full_bar(s), kb * args.kb_elems, nb * SS_BN);
                    if(++s == SS_STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    }
    else if(warp == 1)
    {
        // ===================== MMA issuer =====================
        int s = 0;
        uint32_t ph = 0;
        int as = 0;
        uint32_t aph = 0;
        for(int t = blockIdx.x; t < num_tiles; t += gridDim.x)
        {
            ptx::mbar_wait(tempty_bar(as), aph ^ 1);
            ptx::tcgen05_fence_after_sync();
            const uint32_t d_tmem = tmem_base + as * SS_BN;
            for(int kb = 0; kb < num_kb; ++kb)
            {
                ptx::mbar_wait(full_bar(s), ph);
                ptx::tcgen05_fence_after_sync();
                if(lane == 0)
                {
                    const uint32_t sa = smem_base + s * SS_STAGE_BYTES;
                    const uint64_t da = ptx::make_smem_desc_sw128(sa, 1024);
                    const uint64_t db = ptx::make_smem_desc_sw128(sa + SS_A_BYTES, 1024);
#pragma unroll
                    for(int j = 0; j < SS_BK_BYTES / 32; ++j)
                    {
                        // +32 bytes of K inside the swizzle atom = +2 in the (addr >> 4) field
                        if(args.is_i8)
                            ptx::mma_i8_ss(d_tmem, da + 2 * j, db + 2 * j, args.idesc, (kb | j) != 0);
                        else
                            ptx::mma_f16_ss(d_tmem, da + 2 * j, db + 2 * j, args.idesc, (kb | j) != 0);
                    }
                    ptx::mma_commit(empty_bar(s));
                    if(kb == num_kb - 1) ptx::mma_commit(tfull_bar(as));
                }
                __syncwarp();
                if(++s == SS_STAGES) { s = 0; ph ^= 1; }
            }
            if(++as == 2) { as = 0; aph ^= 1; }
        }
    }
    else if(warp >= 4)
    {
        // ===================== epilogue =====================
        const int q = warp & 3;
        int as = 0;
        uint32_t aph = 0;
        for(int t = blockIdx.x; t < num_tiles; t += gridDim.x)
        {
            int mb, nb;
            tile_coord(t, args.num_m_blk, args.num_n_blk, mb, nb, args.band);
            ptx::mbar_wait(tfull_bar(as), aph);
            ptx::tcgen05_fence_after_sync();
            const int m          = mb * SS_BM + q * 32 + lane;
            const uint32_t taddr = tmem_base + as * SS_BN + (uint32_t(q * 32) << 16);
#pragma unroll 1
            for(int c = 0; c < SS_BN / 32; ++c)
            {
                uint32_t r[32];
                ptx::tmem_ld_x32(taddr + c * 32, r);
                ptx::tmem_ld_wait();
                ss_epilogue_chunk(args, m, nb * SS_BN + c * 32, r);
            }
            ptx::tcgen05_fence_before_sync();
            __syncwarp();
            if(lane == 0) ptx::mbar_arrive(tempty_bar(as));
            if(++as == 2) { as = 0; aph ^= 1; }
        }
    }

    ptx::tcgen05_fence_before_sync();
    __syncthreads();
    if(warp == 2)
    {
        ptx::tcgen05_fence_after_sync();
        ptx::tmem_dealloc<512>(tmem_base);
    }
}


// ---------------------------------------------------------------------------------------------
// CTA-pair variant (cta_group::2): a cluster of two CTAs on one TPC owns a 256 (M) x 256 (N) tile.  CTA r loads A rows
// [128r, 128r+128) and B rows (= output columns) [128r, 128r+128) of every k-block into its OWN shared memory -- half
// the B traffic per SM of the 1-CTA kernel -- and both credit the LEADER's full barrier; the leader's elected thread
// issues tcgen05.mma.cta_group::2 (M = 256: 128 accumulator lanes in each CTA's TMEM) and multicasts the commits to the
// barriers of both CTAs.  Each CTA's epilogue warps drain their own 128 rows and release the accumulator on the
// leader's barrier.  32 KB per stage and CTA -> a 6-stage ring.
constexpr int SS2_BM       = 256;
constexpr int SS2_STAGES   = 6;
constexpr int SS2_HALF     = 128 * SS_BK_BYTES; // one operand half per stage and CTA
constexpr int SS2_STAGE_BYTES = 2 * SS2_HALF;
constexpr int SS2_SMEM     = SS2_STAGES * SS2_STAGE_BYTES + 1024 + 256;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(SS_THREADS, 1)
gemm_ss2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBh, const SSArgs args)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base  = smem_base + SS2_STAGES * SS2_STAGE_BYTES;
    auto full_bar   = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar  = [&](int s) { return bar_base + 8u * (SS2_STAGES + s); };
    auto tfull_bar  = [&](int a) { return bar_base + 8u * (2 * SS2_STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * SS2_STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * SS2_STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

    if(warp == 0 && lane == 0)
    {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmBh);
    }
    if(warp == 1 && lane == 0)
    {
        for(int s = 0; s < SS2_STAGES; ++s)
        {
            ptx::mbar_init(full_bar(s), 1);  // used in the leader only: its producer's arrive + both CTAs' bytes
            ptx::mbar_init(empty_bar(s), 1); // one multicast commit
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

    const int num_m_blk = (args.M + SS2_BM - 1) / SS2_BM;
    const int num_tiles = num_m_blk * args.num_n_blk;
    const int num_kb    = args.num_kb;

    if(warp == 0 || warp == 3)
    {
        // ===================== TMA producers (both CTAs; warp 0 loads A, warp 3 loads B) =====================
        if(lane == 0)
        {
            const bool is_a = warp == 0;
            int s = 0;
            uint32_t ph = 0;
            for(int t = pair; t < num_tiles; t += num_pairs)
            {
                int mb, nb;
                tile_coord(t, num_m_blk, args.num_n_blk, mb, nb, args.band);
                for(int kb = 0; kb < num_kb; ++kb)
                {
                    ptx::mbar_wait(empty_bar(s), ph ^ 1);
                    const uint32_t sa = smem_base + s * SS2_STAGE_BYTES;
                    if(is_a)
                    {
                        if(rank == 0) ptx::mbar_arrive_expect_tx(full_bar(s), 2 * SS2_STAGE_BYTES);
                        ptx::tma_load_2d_pair(sa, &tmA, full_bar(s), kb * args.kb_elems, mb * SS2_BM + int(rank) * 128);
                    }
                    else
                        ptx::tma_load_2d_pair(sa + SS2_HALF, &tmBh, full_bar(s), kb * args.kb_elems, nb * SS_BN + int(rank) * 128);
                    if(++s == SS2_STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    }
    else if(warp == 1 && rank == 0)
    {
        // ===================== MMA issuer (leader CTA only) =====================
        int s = 0;
        uint32_t ph = 0;
        int as = 0;
        uint32_t aph = 0;
        for(int t = pair; t < num_tiles; t += num_pairs)
        {
            ptx::mbar_wait(tempty_bar(as), aph ^ 1);
            ptx::tcgen05_fence_after_sync();
            const uint32_t d_tmem = tmem_base + as * SS_BN;
            for(int kb = 0; kb < num_kb; ++kb)
            {
                ptx::mbar_wait(full_bar(s), ph);
                ptx::tcgen05_fence_after_sync();
                if(lane == 0)
                {
                    const uint32_t sa = smem_base + s * SS2_STAGE_BYTES;
                    const uint64_t da = ptx::make_smem_desc_sw128(sa, 1024);
                    const uint64_t db = ptx::make_smem_desc_sw128(sa + SS2_HALF, 1024);
#pragma unroll
                    for(int j = 0; j < SS_BK_BYTES / 32; ++j)
                    {
                        if(args.is_i8)
                            ptx::mma_i8_ss_pair(d_tmem, da + 2 * j, db + 2 * j, args.idesc, (kb | j) != 0);
                        else
                            ptx::mma_f16_ss_pair(d_tmem, da + 2 * j, db + 2 * j, args.idesc, (kb | j) != 0);
                    }
                    ptx::mma_commit_pair(empty_bar(s));
                    if(kb == num_kb - 1) ptx::mma_commit_pair(tfull_bar(as));
                }
                __syncwarp();
                if(++s == SS2_STAGES) { s = 0; ph ^= 1; }
            }
            if(++as == 2) { as = 0; aph ^= 1; }
        }
    }
    else if(warp >= 4)
    {
        // ===================== epilogue (both CTAs, own 128 rows) =====================
        const int q = warp & 3;
        int as = 0;
        uint32_t aph = 0;
        for(int t = pair; t < num_tiles; t += num_pairs)
        {
            int mb, nb;
            tile_coord(t, num_m_blk, args.num_n_blk, mb, nb, args.band);
            ptx::mbar_wait(tfull_bar(as), aph);
            ptx::tcgen05_fence_after_sync();
            const int m          = mb * SS2_BM + int(rank) * 128 + q * 32 + lane;
            const uint32_t taddr = tmem_base + as * SS_BN + (uint32_t(q * 32) << 16);
#pragma unroll 1
            for(int c = 0; c < SS_BN / 32; ++c)
            {
                uint32_t r[32];
                ptx::tmem_ld_x32(taddr + c * 32, r);
                ptx::tmem_ld_wait();
                ss_epilogue_chunk(args, m, nb * SS_BN + c * 32, r);
            }
            ptx::tcgen05_fence_before_sync();
            __syncwarp();
            if(lane == 0) ptx::mbar_arrive_remote(ptx::leader_addr(tempty_bar(as)));
            if(++as == 2) { as = 0; aph ^= 1; }
        }
    }

    ptx::tcgen05_fence_before_sync();
    ptx::cluster_sync(); // neither CTA may leave (or free TMEM) while the other can still touch its barriers / TMEM
    if(warp == 2)
    {
        ptx::tcgen05_fence_after_sync();
        ptx::tmem_dealloc_pair<512>(tmem_base);
    }
}

} // namespace

// ---------------------------------------------------------------------------------------------
int plan_gemm_ss(const ckb_gemm_problem& p, ckb_gemm_plan& plan)
{
    const bool i8 = p.a_dtype == CKB_I8;
    const int eb  = dtype_bytes(p.a_dtype);
    SSPlan* sp    = reinterpret_cast<SSPlan*>(plan.opaque);
    static_assert(sizeof(SSPlan) <= sizeof(plan.opaque), "opaque too small");
    const CUtensorMapDataType dt = i8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                      : (p.a_dtype == CKB_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                               : CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
    const uint32_t box_k = SS_BK_BYTES / eb;
    if(int rc = encode_tmap_2d(&sp->tmA, dt, eb, p.a, p.M, p.K, p.lda, SS_BM, box_k, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    if(int rc = encode_tmap_2d(&sp->tmB, dt, eb, p.b, p.N, p.K, p.ldb, SS_BN, box_k, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    if(int rc = encode_tmap_2d(&sp->tmBh, dt, eb, p.b, p.N, p.K, p.ldb, 128, box_k, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    const int num_m_blk = (p.M + SS_BM - 1) / SS_BM, num_n_blk = (p.N + SS_BN - 1) / SS_BN;
    const int sms       = device_info().sm_count > 0 ? device_info().sm_count : 148;
    // CTA pairs (256x256 tiles) once there are enough pair tiles to fill the machine; CKB_SS_2CTA=0/1 overrides
    const int pair_tiles = ((p.M + SS2_BM - 1) / SS2_BM) * num_n_blk;
    // (measured on B200, tests/ss_pair_probe.py: int8 gains 6-12 % from the pair, bf16/fp16 is level or slower, so 16-bit stays 1-CTA)
    sp->pair = i8 && p.M >= 2 * SS2_BM && pair_tiles >= sms / 2;
    if(p.flags & CKB_FLAG_SS_1CTA) sp->pair = 0;
    if(p.flags & CKB_FLAG_SS_2CTA) sp->pair = 1;
    if(const char* e = getenv("CKB_SS_2CTA")) sp->pair = atoi(e) != 0;
    if(sp->pair) plan.grid = 2 * std::max(1, std::min(sms / 2, pair_tiles));
    else plan.grid = std::max(1, std::min(sms, num_m_blk * num_n_blk));
    plan.block      = SS_THREADS;
    plan.smem_bytes = sp->pair ? SS2_SMEM : SS_SMEM;
    plan.workspace_bytes = 0;
    plan.split_units     = 1;
    snprintf(plan.kernel_name, sizeof plan.kernel_name, sp->pair ? "gemm_ss_%s_tcgen05_2cta_256x256x%d_s%d" : "gemm_ss_%s_tcgen05_128x256x%d_s%d",
             i8 ? "i8" : (p.a_dtype == CKB_BF16 ? "bf16" : "f16"), int(box_k), sp->pair ? SS2_STAGES : SS_STAGES);
    return CKB_OK;
}

static void fill_epi(const ckb_gemm_problem& p, EpiParams& ep)
{
    ep.cde_op = p.cde_op; ep.act = p.act; ep.scale = p.cde_scale; ep.flags = p.flags;
    ep.num_d = p.num_d;
    for(int i = 0; i < CKB_MAX_D; ++i)
    {
        ep.d_dtype[i] = p.d_dtype[i]; ep.d[i] = p.d[i]; ep.d_sm[i] = p.d_sm[i]; ep.d_sn[i] = p.d_sn[i];
    }
    ep.e_dtype = p.e_dtype; ep.e = p.e; ep.lde = p.lde; ep.M = p.M; ep.N = p.N;
}
void fill_epi_params(const ckb_gemm_problem& p, EpiParams& ep) { fill_epi(p, ep); }

int run_gemm_ss(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream)
{
    const SSPlan* sp = reinterpret_cast<const SSPlan*>(plan.opaque);
    const bool i8    = p.a_dtype == CKB_I8;
    const int eb     = dtype_bytes(p.a_dtype);
    SSArgs a{};
    a.M = p.M; a.N = p.N;
    a.kb_elems  = SS_BK_BYTES / eb;
    a.num_kb    = (p.K + a.kb_elems - 1) / a.kb_elems;
    a.num_m_blk = (p.M + SS_BM - 1) / SS_BM;
    a.num_n_blk = (p.N + SS_BN - 1) / SS_BN;
    a.is_i8     = i8;
    a.band      = 8;
    if(const char* e = getenv("CKB_SS_BAND")) a.band = std::max(1, atoi(e));
    if(sp->pair) a.band = std::max(1, a.band / 2); // pair tiles are 256 rows tall
    a.idesc     = i8 ? ptx::make_idesc(2, 1, 1, SS_BM, SS_BN)
                     : ptx::make_idesc(1, p.a_dtype == CKB_BF16 ? 1 : 0, p.a_dtype == CKB_BF16 ? 1 : 0, SS_BM, SS_BN);
    fill_epi(p, a.ep);
    static PerDeviceOnce attr_once;
    uint64_t dev_bit = 0;
    if(attr_once.need(dev_bit))
    {
        CKB_CUDA_CHECK(cudaFuncSetAttribute(gemm_ss_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SS_SMEM));
        CKB_CUDA_CHECK(cudaFuncSetAttribute(gemm_ss2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SS2_SMEM));
        attr_once.mark(dev_bit);
    }
    if(sp->pair)
    {
        // same instruction descriptor with M = 256 (both CTAs' 128 accumulator lanes); cluster dims are compiled in
        a.idesc = i8 ? ptx::make_idesc(2, 1, 1, SS2_BM, SS_BN)
                     : ptx::make_idesc(1, p.a_dtype == CKB_BF16 ? 1 : 0, p.a_dtype == CKB_BF16 ? 1 : 0, SS2_BM, SS_BN);
        gemm_ss2_kernel<<<plan.grid, SS_THREADS, SS2_SMEM, stream>>>(sp->tmA, sp->tmBh, a);
        CKB_CUDA_CHECK(cudaGetLastError());
        return CKB_OK;
    }
    gemm_ss_kernel<<<plan.grid, SS_THREADS, SS_SMEM, stream>>>(sp->tmA, sp->tmB, a);
    CKB_CUDA_CHECK(cudaGetLastError());
    return CKB_OK;
}

} // namespace ckb
