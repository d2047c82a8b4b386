// grouped_gemm.cu -- K4: grouped GEMM with fixed N,K and a DEVICE-resident argument array.
// Replaces kernel_grouped_gemm_xdl_fixed_nk (device/impl/device_grouped_gemm_multi_abd_xdl_fixed_nk.hpp:37-145):
// G independent GEMMs  E_g[M_g,N] = cde( A_g[M_g,K] * bop(B0_g[K,N], B1_g[N]), D0_g ), M_g read on the device.
// The reference gives every group the same number of workgroups (sized for the average M, :510-517) and lets each
// loop over its group's tiles; here ONE persistent grid walks the concatenated tile list of all groups, so a group
// with few rows does not hold idle CTAs.  Pointers live in device memory, so no host-encoded TMA descriptor can
// describe them: tiles are fetched with 16-byte global loads (register prefetch of the next k-tile under the
// current tile's MMAs), the int8 weight is expanded with the Multiply B-functor
// (binary_element_wise_operation.hpp:184-190: bf16( float(int8) * float(bf16 scale) ), truncating) on its way
// into shared memory, fragments come from ldmatrix, products accumulate in fp32 registers (HMMA m16n8k16).
// The client-31 problem (16 groups x ~48 rows x 1024 x 512) is ~10 MB / 0.8 GFLOP: launch- and HBM-bound.
#include <array>
#include <cstddef>
#include "common.cuh"
#include "epilogue.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace ckb {
namespace {

constexpr int GG_BM = 64, GG_BN = 64, GG_BK = 128, GG_THREADS = 256;
// row pitches (elements) = row + 8: 272- / 144-byte rows, so the 8 rows of one ldmatrix hit 8 distinct 16-byte bank groups
constexpr int GG_PITCH_K = GG_BK + 8; // As[m][k], Bs[n][k]
constexpr int GG_PITCH_N = GG_BN + 8; // Bs[k][n] (row-major B)
constexpr int GG_A_ELEMS = GG_BM * GG_PITCH_K;
constexpr int GG_B_ELEMS = (GG_BK * GG_PITCH_N > GG_BN * GG_PITCH_K) ? GG_BK * GG_PITCH_N : GG_BN * GG_PITCH_K;
constexpr int GG_RED_PITCH = GG_BN + 4; // fp32 partial sums [2 k-halves][64][68] = exactly the two A stages
static_assert(2 * GG_BM * GG_RED_PITCH * 4 <= 2 * GG_A_ELEMS * 2, "reduction area must fit in the A stages");
constexpr int GG_SMEM    = 2 * (GG_A_ELEMS + GG_B_ELEMS) * 2 + GG_BN * 4;

// byte offsets into one element of the caller's argument array
// (GroupedGemmMultiABDKernelArgument<1,NumB,NumD>, device_grouped_gemm_multi_abd_fixed_nk.hpp:15-31)
struct GroupedLayout
{
    int stride, off_a, off_b0, off_b1, off_d0, off_e, off_m, off_sa, off_sb0, off_sd0, off_se;
};

struct GroupedArgs
{
    const unsigned char* args;
    GroupedLayout L;
    int G, N, K;
    int has_scale, num_d, d_dtype;
    int a_col; // A is column-major (km_*): A[m,k] = a[k * StrideA + m]
    int cde_op;
    uint32_t flags;
};

template <typename T>
__device__ __forceinline__ T arg_field(const unsigned char* base, int off)
{
    return *reinterpret_cast<const T*>(base + off);
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t* r)
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t* r)
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}

// two int8 codes (bytes i, i+1 of wx = w ^ 0x80808080, i.e. q + 128) times their scales -> packed truncated bf16 pair.
// 0x4B000000 | (q + 128) is the float 2^23 + q + 128, so one PRMT and one exact FADD give float(q) without the
// quarter-rate I2F; the product with the bf16 scale is exact in fp32 (8-bit x 8-bit significands).
// Round 2: one PRMT + one FFMA per code (dequant.cuh, s8_biased_byte_times_scale: fma(2^15 + q + 128, s, -(2^15 + 128) s) = q s
// exactly); n0 / n1 = -32896 * s0 / s1.
template <int I>
__device__ __forceinline__ uint32_t mul_pair_bf16(uint32_t wx, float s0, float s1, float n0, float n1)
{
    const float f0 = fmaf(__uint_as_float(__byte_perm(wx, 0x47000000u, 0x7404 | (I << 4))), s0, n0);
    const float f1 = fmaf(__uint_as_float(__byte_perm(wx, 0x47000000u, 0x7404 | ((I + 1) << 4))), s1, n1);
    return __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632); // {hi16(f1), hi16(f0)}: truncation
}

// B_ROW = true : B0[k][n] (mk_kn_mn), shared tile Bs[k][n], fragments through ldmatrix.trans
// B_ROW = false: B0[n][k] (mk_nk_mn), shared tile Bs[n][k]
// 8 warps: warp w owns output rows 16*(w&3).. of the 64x64 tile and the k16 steps of half (w>>2) of every 128-k stage;
// the two k-halves meet in shared memory before the epilogue (an in-CTA split-K: twice the warps per byte in flight).
template <bool B_ROW, int CDE>
__global__ void __launch_bounds__(GG_THREADS, 2) grouped_gemm_kernel(const GroupedArgs ga)
{
    extern __shared__ __align__(16) unsigned char gg_smem[];
    uint16_t* const As_base = reinterpret_cast<uint16_t*>(gg_smem);
    uint16_t* const Bs_base = As_base + 2 * GG_A_ELEMS;
    float* const s_scale    = reinterpret_cast<float*>(Bs_base + 2 * GG_B_ELEMS);
    float* const red        = reinterpret_cast<float*>(gg_smem); // [2 k-halves][64][GG_RED_PITCH], aliases the A stages
    constexpr int BP        = B_ROW ? GG_PITCH_N : GG_PITCH_K;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp & 3, wk = warp >> 2;
    const int NT = (ga.N + GG_BN - 1) / GG_BN;
    const int KT = (ga.K + GG_BK - 1) / GG_BK;
    int buf = 0;

    for(int t = blockIdx.x;; t += gridDim.x)
    {
        // ---- locate tile t in the concatenated (group, m-tile, n-tile) list -------------------------------
        // (lane l reads M of group c+l, a warp scan gives the tile prefix: one round trip to L2 per 32 groups
        //  instead of one per group)
        int g = ga.G, base = 0, Mg = 0;
        for(int c = 0; c < ga.G; c += 32)
        {
            const int gi  = c + lane;
            const int Ml  = gi < ga.G ? arg_field<int>(ga.args + size_t(gi) * ga.L.stride, ga.L.off_m) : 0;
            const int cnt = ((Ml + GG_BM - 1) / GG_BM) * NT;
            int incl      = cnt;
#pragma unroll
            for(int d = 1; d < 32; d <<= 1)
            {
                const int v = __shfl_up_sync(0xffffffffu, incl, d);
                if(lane >= d) incl += v;
            }
            const unsigned hit = __ballot_sync(0xffffffffu, gi < ga.G && t < base + incl);
            if(hit)
            {
                const int src = __ffs(hit) - 1;
                g             = c + src;
                Mg            = __shfl_sync(0xffffffffu, Ml, src);
                base += __shfl_sync(0xffffffffu, incl - cnt, src);
                break;
            }
            base += __shfl_sync(0xffffffffu, incl, 31);
        }
        if(g >= ga.G) break;
        const unsigned char* ap = ga.args + size_t(g) * ga.L.stride;
        const int m0 = ((t - base) / NT) * GG_BM, n0 = ((t - base) % NT) * GG_BN;
        const uint16_t* A   = arg_field<const uint16_t*>(ap, ga.L.off_a);
        const int8_t* B     = arg_field<const int8_t*>(ap, ga.L.off_b0);
        const long long lda = arg_field<int>(ap, ga.L.off_sa), ldb = arg_field<int>(ap, ga.L.off_sb0);

        uint4 ra[4], rb[2];
        auto load_regs = [&](int kt) {
            const int k0 = kt * GG_BK;
            if(ga.a_col)
            {
                // column-major A: vectors run along m (8 rows at one k); 16-byte loads when the pitch and the pointer allow it,
                // element loads otherwise and at the ragged end of the group (rows past M_g repeat the last row)
                const bool vec_ok = (lda % 8 == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);
#pragma unroll
                for(int i = 0; i < 4; ++i)
                {
                    const int c = tid + i * GG_THREADS, k = c >> 3, mc = (c & 7) * 8;
                    if(k0 + k >= ga.K) { ra[i] = make_uint4(0, 0, 0, 0); continue; }
                    const uint16_t* col = A + (long long)(k0 + k) * lda;
                    if(vec_ok && m0 + mc + 8 <= Mg) ra[i] = *reinterpret_cast<const uint4*>(col + m0 + mc);
                    else
                    {
                        uint16_t e[8];
#pragma unroll
                        for(int j = 0; j < 8; ++j) e[j] = col[min(m0 + mc + j, Mg - 1)];
                        ra[i] = make_uint4(e[0] | uint32_t(e[1]) << 16, e[2] | uint32_t(e[3]) << 16, e[4] | uint32_t(e[5]) << 16, e[6] | uint32_t(e[7]) << 16);
                    }
                }
            }
            else
#pragma unroll
            for(int i = 0; i < 4; ++i)
            {
                const int c = tid + i * GG_THREADS, r = c >> 4, kc = (c & 15) * 8;
                const int m = min(m0 + r, Mg - 1); // rows past M_g repeat the last row; they are never stored
                ra[i] = (k0 + kc < ga.K) ? *reinterpret_cast<const uint4*>(A + m * lda + k0 + kc) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for(int i = 0; i < 2; ++i)
            {
                const int c = tid + i * GG_THREADS;
                const int r = B_ROW ? c >> 2 : c >> 3, cc = B_ROW ? (c & 3) * 16 : (c & 7) * 16;
                const int k = B_ROW ? k0 + r : k0 + cc, n = B_ROW ? n0 + cc : n0 + r;
                const bool ok = k < ga.K && n < ga.N;
                rb[i] = ok ? *reinterpret_cast<const uint4*>(B_ROW ? B + k * ldb + n : B + n * ldb + k) : make_uint4(0, 0, 0, 0);
            }
        };
        load_regs(0); // in flight while the scales arrive
        if(tid < GG_BN)
        {
            float s = 1.f;
            if(ga.has_scale && n0 + tid < ga.N) s = bf16_bits_to_f32(arg_field<const uint16_t*>(ap, ga.L.off_b1)[n0 + tid]);
            s_scale[tid] = s;
        }
        auto store_smem = [&](int b) {
            uint16_t* As = As_base + b * GG_A_ELEMS;
            uint16_t* Bs = Bs_base + b * GG_B_ELEMS;
            if(ga.a_col)
            {
                // transpose on the way in: the register holds 8 rows at one k
#pragma unroll
                for(int i = 0; i < 4; ++i)
                {
                    const int c = tid + i * GG_THREADS, k = c >> 3, mc = (c & 7) * 8;
                    const uint32_t w[4] = {ra[i].x, ra[i].y, ra[i].z, ra[i].w};
#pragma unroll
                    for(int j = 0; j < 8; ++j) As[(mc + j) * GG_PITCH_K + k] = uint16_t(w[j >> 1] >> (16 * (j & 1)));
                }
            }
            else
#pragma unroll
            for(int i = 0; i < 4; ++i)
            {
                const int c = tid + i * GG_THREADS, r = c >> 4, kc = (c & 15) * 8;
                *reinterpret_cast<uint4*>(As + r * GG_PITCH_K + kc) = ra[i];
            }
#pragma unroll
            for(int i = 0; i < 2; ++i)
            {
                const int c = tid + i * GG_THREADS;
                const int r = B_ROW ? c >> 2 : c >> 3, cc = B_ROW ? (c & 3) * 16 : (c & 7) * 16;
                const uint32_t w[4] = {rb[i].x ^ 0x80808080u, rb[i].y ^ 0x80808080u, rb[i].z ^ 0x80808080u, rb[i].w ^ 0x80808080u};
                uint32_t o[8];
#pragma unroll
                for(int j = 0; j < 8; j += 2)
                {
                    // 16 consecutive codes: along n at fixed k (B_ROW: one scale each) or along k at fixed n
                    const float s0 = B_ROW ? s_scale[cc + 2 * j] : s_scale[r], s1 = B_ROW ? s_scale[cc + 2 * j + 1] : s0;
                    const float s2 = B_ROW ? s_scale[cc + 2 * j + 2] : s0, s3 = B_ROW ? s_scale[cc + 2 * j + 3] : s0;
                    o[j]     = mul_pair_bf16<0>(w[j >> 1], s0, s1, s0 * -32896.f, s1 * -32896.f);
                    o[j + 1] = mul_pair_bf16<2>(w[j >> 1], s2, s3, s2 * -32896.f, s3 * -32896.f);
                }
                *reinterpret_cast<uint4*>(Bs + r * BP + cc)     = make_uint4(o[0], o[1], o[2], o[3]);
                *reinterpret_cast<uint4*>(Bs + r * BP + cc + 8) = make_uint4(o[4], o[5], o[6], o[7]);
            }
        };

        float acc[8][4];
#pragma unroll
        for(int i = 0; i < 8; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

        __syncthreads(); // s_scale visible; every warp is done with the previous tile's buffers / reduction area
        for(int kt = 0; kt < KT; ++kt)
        {
            store_smem(buf);
            __syncthreads();
            if(kt + 1 < KT) load_regs(kt + 1);
            const uint16_t* As = As_base + buf * GG_A_ELEMS;
            const uint16_t* Bs = Bs_base + buf * GG_B_ELEMS;
            if(m0 + wm * 16 < Mg) // a warp whose 16 rows lie past M_g only converts and synchronises (48-row groups: a quarter of the MMAs)
#pragma unroll
            for(int kq = 0; kq < GG_BK / 32; ++kq)
            {
                const int ks = wk * (GG_BK / 32) + kq;
                uint32_t af[4];
                ldmatrix_x4(ptx::smem_u32(As + (wm * 16 + (lane & 15)) * GG_PITCH_K + ks * 16 + (lane >> 4) * 8), af);
#pragma unroll
                for(int np = 0; np < 4; ++np)
                {
                    uint32_t bf[4]; // {b0,b1} of n8 tile 2np, {b0,b1} of n8 tile 2np+1
                    if(B_ROW)
                        ldmatrix_x4_trans(ptx::smem_u32(Bs + (ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * BP + np * 16 + (lane >> 4) * 8), bf);
                    else
                        ldmatrix_x4(ptx::smem_u32(Bs + (np * 16 + (lane & 7) + (lane >> 4) * 8) * BP + ks * 16 + ((lane >> 3) & 1) * 8), bf);
                    ptx::mma_m16n8k16<false>(acc[2 * np], af, bf[0], bf[1]);
                    ptx::mma_m16n8k16<false>(acc[2 * np + 1], af, bf[2], bf[3]);
                }
            }
            buf ^= 1;
        }

        // ---- both k-halves park their fp32 partial sums in shared memory (over the A stages); the epilogue is a
        //      compact rolled loop: thread -> 8 consecutive columns of one row, sum of the halves, CDE functor,
        //      one 16-byte store.  (Applying the functor on the MMA fragments unrolls FastGelu's expf + divide 32
        //      times: ~5600 SASS instructions, and ncu showed the warps starving on instruction fetch.)
        __syncthreads(); // all MMAs of the last stage are done: the A stages may be overwritten
        {
            float* my = red + wk * (GG_BM * GG_RED_PITCH) + (wm * 16 + (lane >> 2)) * GG_RED_PITCH + (lane & 3) * 2;
#pragma unroll
            for(int nt = 0; nt < 8; ++nt)
            {
                *reinterpret_cast<float2*>(my + nt * 8)                    = make_float2(acc[nt][0], acc[nt][1]);
                *reinterpret_cast<float2*>(my + 8 * GG_RED_PITCH + nt * 8) = make_float2(acc[nt][2], acc[nt][3]);
            }
        }
        __syncthreads();
        EpiParams ep{};
        ep.cde_op = CDE; ep.flags = ga.flags; ep.e_dtype = CKB_BF16; ep.d_dtype[0] = CKB_BF16;
        uint16_t* E         = arg_field<uint16_t*>(ap, ga.L.off_e);
        const long long lde = arg_field<int>(ap, ga.L.off_se);
        const void* D0      = ga.num_d ? arg_field<const void*>(ap, ga.L.off_d0) : nullptr;
        const long long ldd = ga.num_d ? arg_field<int>(ap, ga.L.off_sd0) : 0;
#pragma unroll 1
        for(int c = tid; c < GG_BM * GG_BN / 8; c += GG_THREADS)
        {
            const int r = c >> 3, cc = (c & 7) * 8, m = m0 + r, n = n0 + cc;
            if(m >= Mg || n >= ga.N) continue; // N % 16 == 0 (host check): the 8 columns are all inside
            const float* p0 = red + r * GG_RED_PITCH + cc;
            const float* p1 = p0 + GG_BM * GG_RED_PITCH;
            float v[8];
#pragma unroll
            for(int q = 0; q < 8; q += 4)
            {
                const float4 x = *reinterpret_cast<const float4*>(p0 + q), y = *reinterpret_cast<const float4*>(p1 + q);
                v[q] = x.x + y.x; v[q + 1] = x.y + y.y; v[q + 2] = x.z + y.z; v[q + 3] = x.w + y.w;
            }
            uint32_t o[4];
#pragma unroll
            for(int q = 0; q < 8; q += 2)
            {
                float d0 = 0.f, d1 = 0.f;
                if(D0)
                {
                    const long long off = m * ldd + n + q;
                    if(ga.d_dtype == CKB_F32) { d0 = static_cast<const float*>(D0)[off]; d1 = static_cast<const float*>(D0)[off + 1]; }
                    else { d0 = bf16_bits_to_f32(static_cast<const uint16_t*>(D0)[off]); d1 = bf16_bits_to_f32(static_cast<const uint16_t*>(D0)[off + 1]); }
                }
                const float e0 = cde_apply_f32(ep, v[q], d0, 0.f), e1 = cde_apply_f32(ep, v[q + 1], d1, 0.f);
                o[q >> 1] = uint32_t(f32_to_bf16_bits(e0, ga.flags)) | (uint32_t(f32_to_bf16_bits(e1, ga.flags)) << 16);
            }
            uint16_t* dst = E + m * lde + n;
            if((reinterpret_cast<uintptr_t>(dst) & 15) == 0) *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
            else
            {
#pragma unroll
                for(int q = 0; q < 4; ++q) *reinterpret_cast<uint32_t*>(dst + 2 * q) = o[q];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------------
// Round-2 form for row-major A (mk_nk_mn / mk_kn_mn): the same tiles, fragments and arithmetic as the kernel above, but the tile's
// loads are no longer one stage ahead of the math.  ncu on the kernel above (profiles/r1_grouped_ncu_summary.txt) shows a tile as
// KT serial round trips to HBM (each stage's loads are issued when the previous stage starts its MMAs, and a client-31 tile has
// only 4 stages); here G3_ST = 4 stages are in flight from the start of the tile: A through cp.async into a 4-slot ring (no
// registers), the int8 weight through 4 rotating register sets (it is converted on its way into shared memory, so it has to pass
// through registers anyway).  A client-31 tile (K = 512) therefore waits for HBM once.
constexpr int G3_ST   = 4;
constexpr int G3_SMEM = (G3_ST * GG_A_ELEMS + 2 * GG_B_ELEMS) * 2 + GG_BN * 4;
static_assert(2 * G3_SMEM + 2048 <= 228 * 1024, "two CTAs per SM");

__device__ __forceinline__ void cp_async_16(uint32_t dst, const void* src, int src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <bool B_ROW, int CDE>
__global__ void __launch_bounds__(GG_THREADS, 2) grouped_gemm_kernel_v3(const GroupedArgs ga)
{
    extern __shared__ __align__(16) unsigned char gg_smem[];
    uint16_t* const As_base = reinterpret_cast<uint16_t*>(gg_smem);
    uint16_t* const Bs_base = As_base + G3_ST * GG_A_ELEMS;
    float* const s_scale    = reinterpret_cast<float*>(Bs_base + 2 * GG_B_ELEMS);
    float* const red        = reinterpret_cast<float*>(gg_smem); // [2 k-halves][64][GG_RED_PITCH], aliases A slots 0 and 1
    constexpr int BP        = B_ROW ? GG_PITCH_N : GG_PITCH_K;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp & 3, wk = warp >> 2;
    const int NT = (ga.N + GG_BN - 1) / GG_BN;
    const int KT = (ga.K + GG_BK - 1) / GG_BK;
    int buf = 0;

    for(int t = blockIdx.x;; t += gridDim.x)
    {
        // ---- locate tile t (same walk as above) ------------------------------------------------------------
        int g = ga.G, base = 0, Mg = 0;
        for(int c = 0; c < ga.G; c += 32)
        {
            const int gi  = c + lane;
            const int Ml  = gi < ga.G ? arg_field<int>(ga.args + size_t(gi) * ga.L.stride, ga.L.off_m) : 0;
            const int cnt = ((Ml + GG_BM - 1) / GG_BM) * NT;
            int incl      = cnt;
#pragma unroll
            for(int d = 1; d < 32; d <<= 1)
            {
                const int v = __shfl_up_sync(0xffffffffu, incl, d);
                if(lane >= d) incl += v;
            }
            const unsigned hit = __ballot_sync(0xffffffffu, gi < ga.G && t < base + incl);
            if(hit)
            {
                const int src = __ffs(hit) - 1;
                g             = c + src;
                Mg            = __shfl_sync(0xffffffffu, Ml, src);
                base += __shfl_sync(0xffffffffu, incl - cnt, src);
                break;
            }
            base += __shfl_sync(0xffffffffu, incl, 31);
        }
        if(g >= ga.G) break;
        const unsigned char* ap = ga.args + size_t(g) * ga.L.stride;
        const int m0 = ((t - base) / NT) * GG_BM, n0 = ((t - base) % NT) * GG_BN;
        const uint16_t* A   = arg_field<const uint16_t*>(ap, ga.L.off_a);
        const int8_t* B     = arg_field<const int8_t*>(ap, ga.L.off_b0);
        const long long lda = arg_field<int>(ap, ga.L.off_sa), ldb = arg_field<int>(ap, ga.L.off_sb0);

        // this thread's two 16-byte pieces of a weight stage: (row r, 16 codes from cc); the same (r, cc) in every stage
        const int b_r0 = B_ROW ? tid >> 2 : tid >> 3, b_cc = B_ROW ? (tid & 3) * 16 : (tid & 7) * 16;
        constexpr int B_DR = B_ROW ? GG_THREADS / 4 : GG_THREADS / 8; // row step between the two pieces
        auto load_b = [&](int kt, uint4* rb) {
            const int k0 = kt * GG_BK;
#pragma unroll
            for(int i = 0; i < 2; ++i)
            {
                const int r = b_r0 + i * B_DR;
                const int k = B_ROW ? k0 + r : k0 + b_cc, n = B_ROW ? n0 + b_cc : n0 + r;
                const bool ok = k < ga.K && n < ga.N;
                rb[i] = ok ? *reinterpret_cast<const uint4*>(B_ROW ? B + k * ldb + n : B + n * ldb + k) : make_uint4(0, 0, 0, 0);
            }
        };
        // A stage kt -> ring slot `slot`: rows past M_g repeat the last row (never stored), k past K is zero-filled
        const int a_r0 = tid >> 4, a_kc = (tid & 15) * 8;
        auto issue_a = [&](int kt, int slot) {
            const int k0       = kt * GG_BK;
            const bool ok      = k0 + a_kc < ga.K;
            const uint32_t dst = ptx::smem_u32(As_base + slot * GG_A_ELEMS + a_r0 * GG_PITCH_K + a_kc);
#pragma unroll
            for(int i = 0; i < 4; ++i)
            {
                const int m = min(m0 + a_r0 + 16 * i, Mg - 1);
                cp_async_16(dst + i * (16 * GG_PITCH_K * 2), ok ? static_cast<const void*>(A + m * lda + k0 + a_kc) : static_cast<const void*>(A), ok ? 16 : 0);
            }
        };
        auto convert_b = [&](const uint4* rb, int b) {
            uint16_t* Bs = Bs_base + b * GG_B_ELEMS;
#pragma unroll
            for(int i = 0; i < 2; ++i)
            {
                const int r = b_r0 + i * B_DR;
                const uint32_t w[4] = {rb[i].x ^ 0x80808080u, rb[i].y ^ 0x80808080u, rb[i].z ^ 0x80808080u, rb[i].w ^ 0x80808080u};
                uint32_t o[8];
                if(B_ROW)
                {
#pragma unroll
                    for(int j = 0; j < 8; j += 2)
                    {
                        const float4 s = *reinterpret_cast<const float4*>(s_scale + b_cc + 2 * j); // 16 codes along n: one scale each
                        o[j]     = mul_pair_bf16<0>(w[j >> 1], s.x, s.y, s.x * -32896.f, s.y * -32896.f);
                        o[j + 1] = mul_pair_bf16<2>(w[j >> 1], s.z, s.w, s.z * -32896.f, s.w * -32896.f);
                    }
                }
                else
                {
                    const float s = s_scale[r], ns = s * -32896.f; // 16 codes along k of channel r
#pragma unroll
                    for(int j = 0; j < 8; j += 2)
                    {
                        o[j]     = mul_pair_bf16<0>(w[j >> 1], s, s, ns, ns);
                        o[j + 1] = mul_pair_bf16<2>(w[j >> 1], s, s, ns, ns);
                    }
                }
                *reinterpret_cast<uint4*>(Bs + r * BP + b_cc)     = make_uint4(o[0], o[1], o[2], o[3]);
                *reinterpret_cast<uint4*>(Bs + r * BP + b_cc + 8) = make_uint4(o[4], o[5], o[6], o[7]);
            }
        };

        uint4 rb[G3_ST][2];
#pragma unroll
        for(int j = 0; j < G3_ST; ++j)
            if(j < KT) load_b(j, rb[j]);
        if(tid < GG_BN)
        {
            float s = 1.f;
            if(ga.has_scale && n0 + tid < ga.N) s = bf16_bits_to_f32(arg_field<const uint16_t*>(ap, ga.L.off_b1)[n0 + tid]);
            s_scale[tid] = s;
        }
        float acc[8][4];
#pragma unroll
        for(int i = 0; i < 8; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

        __syncthreads(); // s_scale visible; every warp is done with the previous tile's reduction area (A slots 0, 1)
#pragma unroll
        for(int j = 0; j < G3_ST; ++j)
        {
            if(j < KT) issue_a(j, j);
            cp_async_commit();
        }
        // cp.async groups: stage s is group s (the prologue commits stages 0..3; iteration kt >= 1 commits stage kt + 3, possibly
        // empty), so at iteration kt at most G3_ST - 2 younger groups may still be pending when stage kt is needed
        for(int kt0 = 0; kt0 < KT; kt0 += G3_ST)
        {
#pragma unroll
            for(int j = 0; j < G3_ST; ++j)
            {
                const int kt = kt0 + j; // ring slot of stage kt = j
                if(kt < KT)
                {
                    convert_b(rb[j], buf);
                    if(kt + G3_ST < KT) load_b(kt + G3_ST, rb[j]);
                    cp_async_wait<G3_ST - 2>();
                    __syncthreads(); // stage kt (A slot j, Bs[buf]) visible; every warp is done with the MMAs of stage kt - 1
                    if(kt >= 1)
                    {
                        if(kt - 1 + G3_ST < KT) issue_a(kt - 1 + G3_ST, (j + G3_ST - 1) % G3_ST);
                        cp_async_commit();
                    }
                    const uint16_t* As = As_base + j * GG_A_ELEMS;
                    const uint16_t* Bs = Bs_base + buf * GG_B_ELEMS;
                    if(m0 + wm * 16 < Mg)
#pragma unroll
                    for(int kq = 0; kq < GG_BK / 32; ++kq)
                    {
                        const int ks = wk * (GG_BK / 32) + kq;
                        uint32_t af[4];
                        ldmatrix_x4(ptx::smem_u32(As + (wm * 16 + (lane & 15)) * GG_PITCH_K + ks * 16 + (lane >> 4) * 8), af);
#pragma unroll
                        for(int np = 0; np < 4; ++np)
                        {
                            uint32_t bf[4];
                            if(B_ROW)
                                ldmatrix_x4_trans(ptx::smem_u32(Bs + (ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * BP + np * 16 + (lane >> 4) * 8), bf);
                            else
                                ldmatrix_x4(ptx::smem_u32(Bs + (np * 16 + (lane & 7) + (lane >> 4) * 8) * BP + ks * 16 + ((lane >> 3) & 1) * 8), bf);
                            ptx::mma_m16n8k16<false>(acc[2 * np], af, bf[0], bf[1]);
                            ptx::mma_m16n8k16<false>(acc[2 * np + 1], af, bf[2], bf[3]);
                        }
                    }
                    buf ^= 1;
                }
            }
        }
        cp_async_wait<0>();
        __syncthreads(); // all MMAs of the last stage are done: A slots 0 and 1 may be overwritten
        {
            float* my = red + wk * (GG_BM * GG_RED_PITCH) + (wm * 16 + (lane >> 2)) * GG_RED_PITCH + (lane & 3) * 2;
#pragma unroll
            for(int nt = 0; nt < 8; ++nt)
            {
                *reinterpret_cast<float2*>(my + nt * 8)                    = make_float2(acc[nt][0], acc[nt][1]);
                *reinterpret_cast<float2*>(my + 8 * GG_RED_PITCH + nt * 8) = make_float2(acc[nt][2], acc[nt][3]);
            }
        }
        __syncthreads();
        EpiParams ep{};
        ep.cde_op = CDE; ep.flags = ga.flags; ep.e_dtype = CKB_BF16; ep.d_dtype[0] = CKB_BF16;
        uint16_t* E         = arg_field<uint16_t*>(ap, ga.L.off_e);
        const long long lde = arg_field<int>(ap, ga.L.off_se);
        const void* D0      = ga.num_d ? arg_field<const void*>(ap, ga.L.off_d0) : nullptr;
        const long long ldd = ga.num_d ? arg_field<int>(ap, ga.L.off_sd0) : 0;
#pragma unroll 1
        for(int c = tid; c < GG_BM * GG_BN / 8; c += GG_THREADS)
        {
            const int r = c >> 3, cc = (c & 7) * 8, m = m0 + r, n = n0 + cc;
            if(m >= Mg || n >= ga.N) continue;
            const float* p0 = red + r * GG_RED_PITCH + cc;
            const float* p1 = p0 + GG_BM * GG_RED_PITCH;
            float v[8];
#pragma unroll
            for(int q = 0; q < 8; q += 4)
            {
                const float4 x = *reinterpret_cast<const float4*>(p0 + q), y = *reinterpret_cast<const float4*>(p1 + q);
                v[q] = x.x + y.x; v[q + 1] = x.y + y.y; v[q + 2] = x.z + y.z; v[q + 3] = x.w + y.w;
            }
            uint32_t o[4];
#pragma unroll
            for(int q = 0; q < 8; q += 2)
            {
                float d0 = 0.f, d1 = 0.f;
                if(D0)
                {
                    const long long off = m * ldd + n + q;
                    if(ga.d_dtype == CKB_F32) { d0 = static_cast<const float*>(D0)[off]; d1 = static_cast<const float*>(D0)[off + 1]; }
                    else { d0 = bf16_bits_to_f32(static_cast<const uint16_t*>(D0)[off]); d1 = bf16_bits_to_f32(static_cast<const uint16_t*>(D0)[off + 1]); }
                }
                const float e0 = cde_apply_f32(ep, v[q], d0, 0.f), e1 = cde_apply_f32(ep, v[q + 1], d1, 0.f);
                o[q >> 1] = uint32_t(f32_to_bf16_bits(e0, ga.flags)) | (uint32_t(f32_to_bf16_bits(e1, ga.flags)) << 16);
            }
            uint16_t* dst = E + m * lde + n;
            if((reinterpret_cast<uintptr_t>(dst) & 15) == 0) *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
            else
            {
#pragma unroll
                for(int q = 0; q < 4; ++q) *reinterpret_cast<uint32_t*>(dst + 2 * q) = o[q];
            }
        }
    }
}

} // namespace

// Layout of GroupedGemmMultiABDKernelArgument<1, num_b, num_d> (device_grouped_gemm_multi_abd_fixed_nk.hpp:15-31 of the
// reference): taken with offsetof from a mirror of the struct, because a zero-length std::array member still occupies
// storage (one byte, then padding) -- the offsets of a NumDTensor == 0 instantiation are NOT "NumD * 8 bytes fewer".
namespace {
template <int NB, int ND>
struct KernelArgMirror
{
    std::array<const void*, 1> p_as_grid;
    std::array<const void*, NB> p_bs_grid;
    std::array<const void*, ND> p_ds_grid;
    void* p_e_grid;
    int M, N, K;
    std::array<int, 1> StrideAs;
    std::array<int, NB> StrideBs;
    std::array<int, ND> StrideDs;
    int StrideE;
};
template <int NB, int ND>
GroupedLayout layout_of()
{
    using T = KernelArgMirror<NB, ND>;
    GroupedLayout L{};
    L.off_a   = int(offsetof(T, p_as_grid));
    L.off_b0  = int(offsetof(T, p_bs_grid));
    L.off_b1  = NB > 1 ? L.off_b0 + 8 : L.off_b0;
    L.off_d0  = int(offsetof(T, p_ds_grid));
    L.off_e   = int(offsetof(T, p_e_grid));
    L.off_m   = int(offsetof(T, M));
    L.off_sa  = int(offsetof(T, StrideAs));
    L.off_sb0 = int(offsetof(T, StrideBs));
    L.off_sd0 = int(offsetof(T, StrideDs));
    L.off_se  = int(offsetof(T, StrideE));
    L.stride  = int(sizeof(T));
    return L;
}
// the instantiations the reference's grouped bf16 x int8 instances use (NumB 1|2, NumD 0|1), as the host compiler lays them out
using KA21 = KernelArgMirror<2, 1>;
using KA20 = KernelArgMirror<2, 0>;
using KA11 = KernelArgMirror<1, 1>;
using KA10 = KernelArgMirror<1, 0>;
static_assert(sizeof(KA21) == 72 && offsetof(KA21, p_e_grid) == 32 && offsetof(KA21, StrideE) == 68, "NumB 2, NumD 1");
static_assert(sizeof(KA20) == 72 && offsetof(KA20, p_e_grid) == 32 && offsetof(KA20, M) == 40 && offsetof(KA20, StrideE) == 68, "NumB 2, NumD 0");
static_assert(sizeof(KA10) == 64 && offsetof(KA10, p_e_grid) == 24 && offsetof(KA10, StrideE) == 56, "NumB 1, NumD 0");
static_assert(sizeof(KA11) == 64 && offsetof(KA11, p_e_grid) == 24, "NumB 1, NumD 1");
} // namespace

static GroupedLayout grouped_layout(int num_b, int num_d)
{
    if(num_b == 2) return num_d ? layout_of<2, 1>() : layout_of<2, 0>();
    return num_d ? layout_of<1, 1>() : layout_of<1, 0>();
}

int grouped_arg_bytes(int num_b, int num_d) { return grouped_layout(num_b, num_d).stride; }

// IsSupportedArgument of the grouped op: what can be checked from the host-side prototype
// (device_grouped_gemm_multi_abd_xdl_fixed_nk.hpp:580-610 checks the same things per gemm_desc).
int check_grouped(const ckb_gemm_problem& p, int group_count)
{
    if(group_count <= 0) return set_error(CKB_ERR_UNSUPPORTED, "group_count must be positive");
    if(p.a_dtype != CKB_BF16 || p.e_dtype != CKB_BF16) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: A and E must be bf16");
    if(p.a_layout != CKB_ROW && p.a_layout != CKB_COL) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: A must be row- or column-major");
    if(p.b_dtype != CKB_I8) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: B0 must be int8");
    if(p.b_layout != CKB_ROW && p.b_layout != CKB_COL) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: B0 must be row- or column-major");
    if(p.b_op != CKB_BOP_MULTIPLY && p.b_op != CKB_BOP_PASSTHROUGH) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: B op must be Multiply or PassThrough");
    if(p.b_op == CKB_BOP_MULTIPLY && p.s_dtype != CKB_BF16) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: B1 must be bf16");
    switch(p.cde_op)
    {
    case CKB_CDE_PASSTHROUGH: case CKB_CDE_FASTGELU:
        if(p.num_d != 0) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: this CDE op takes no D tensor");
        break;
    case CKB_CDE_ADD: case CKB_CDE_ADD_FASTGELU: case CKB_CDE_MULTIPLY: case CKB_CDE_MULTIPLY_FASTGELU:
        if(p.num_d != 1) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: this CDE op takes one D tensor");
        if(p.d_dtype[0] != CKB_BF16 && p.d_dtype[0] != CKB_F32) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: D0 must be bf16 or fp32");
        break;
    default: return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: cde_op %d is not built", p.cde_op);
    }
    if(p.N <= 0 || p.K <= 0) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: N and K must be positive");
    if(p.K % 16 || p.N % 16) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: N and K must be multiples of 16");
    if(p.a_layout == CKB_ROW && (p.lda % 8 || p.lda < p.K)) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: StrideA must be >= K and a multiple of 8");
    if(p.a_layout == CKB_COL && p.lda < 1) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: StrideA (pitch between k rows of column-major A) must be >= M");
    const int b_inner = p.b_layout == CKB_ROW ? p.N : p.K;
    if(p.ldb % 16 || p.ldb < b_inner) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: StrideB0 must cover the row and be a multiple of 16");
    if(p.lde % 2 || p.lde < p.N) return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: StrideE must be >= N and even");
    return CKB_OK;
}

int run_grouped(const ckb_gemm_problem& p, const void* dev_args, int group_count, int sum_m, cudaStream_t stream)
{
    if(int rc = check_grouped(p, group_count)) return rc;
    if(!dev_args) return set_error(CKB_ERR_INVALID, "grouped GEMM: the device argument array is null (SetDeviceKernelArgs)");
    if(sum_m <= 0) return CKB_OK;
    const int num_b = p.b_op == CKB_BOP_MULTIPLY ? 2 : 1;
    GroupedArgs ga{};
    ga.args = static_cast<const unsigned char*>(dev_args);
    ga.L    = grouped_layout(num_b, p.num_d);
    ga.G = group_count; ga.N = p.N; ga.K = p.K;
    ga.has_scale = num_b == 2; ga.num_d = p.num_d; ga.d_dtype = p.d_dtype[0];
    ga.cde_op = p.cde_op; ga.flags = p.flags;
    ga.a_col  = p.a_layout == CKB_COL;
    const long long NT    = (p.N + GG_BN - 1) / GG_BN;
    const long long tiles = ((long long)(sum_m + GG_BM - 1) / GG_BM + group_count) * NT; // upper bound on the tile count
    const int grid        = int(std::min<long long>(tiles, 3LL * device_info().sm_count));
    // row-major A: the form with four stages in flight (cp.async ring); column-major A (transposed on its way into shared memory)
    // and short K use the register-prefetch form
    // (CKB_GROUPED_KERNEL=1 / 3 forces one of them for A/B measurements)
    const char* force = getenv("CKB_GROUPED_KERNEL");
    // measured (profiles/r2_grouped_ab.txt): the deeper pipeline pays from K ~ 2048 up (5 - 14 %); below that its longer prologue costs
    // more than the overlap returns (K = 512: 16.4 vs 14.4 us), the per-stage cost being conversion + MMA issue, not HBM latency
    const bool v3     = !ga.a_col && ((force && *force) ? atoi(force) == 3 : p.K >= 2048);
    auto go = [&](auto kern, int smem) -> int {
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); // cheap, idempotent
        CKB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        kern<<<grid, GG_THREADS, smem, stream>>>(ga);
        return CKB_OK;
    };
    const bool row = p.b_layout == CKB_ROW;
    int rc = CKB_OK;
    switch(p.cde_op)
    {
#define GG_CASE(op)                                                                                                              \
    case op:                                                                                                                     \
        if(v3) rc = row ? go(grouped_gemm_kernel_v3<true, op>, G3_SMEM) : go(grouped_gemm_kernel_v3<false, op>, G3_SMEM);        \
        else rc = row ? go(grouped_gemm_kernel<true, op>, GG_SMEM) : go(grouped_gemm_kernel<false, op>, GG_SMEM);                \
        break;
        GG_CASE(CKB_CDE_PASSTHROUGH) GG_CASE(CKB_CDE_ADD) GG_CASE(CKB_CDE_FASTGELU) GG_CASE(CKB_CDE_ADD_FASTGELU)
        GG_CASE(CKB_CDE_MULTIPLY) GG_CASE(CKB_CDE_MULTIPLY_FASTGELU)
#undef GG_CASE
    default: return set_error(CKB_ERR_UNSUPPORTED, "grouped GEMM: cde_op %d is not built", p.cde_op);
    }
    if(rc) return rc;
    CKB_CUDA_CHECK(cudaGetLastError());
    return CKB_OK;
}

} // namespace ckb
