// wq_dequant.cu -- K1 (round-1 form): prefill-shaped weight-quantized GEMM (M > 64) as two launches:
//   1. this kernel expands the quantized weight ONCE into a K-major 16-bit matrix Bdq[N][K] in the workspace, applying
//      exactly the per-element arithmetic of the decode kernel / the oracle (same helpers, dequant.cuh);
//   2. the dense tcgen05 SS GEMM (gemm_ss.cu, kind::f16) consumes it with the fused CDE epilogue.
// The expansion is HBM-bound (reads N*K/2..N*K bytes, writes 2*N*K) and costs ~6 % of the GEMM at M = 4096; the
// fully fused form (dequant into the TMEM A operand inside the GEMM main loop, as K2 does) is the planned successor.
// Replaces the B path of kernel_gemm_xdl_cshuffle_v3 (multi-ABD), grid/gridwise_gemm_xdl_cshuffle_v3_multi_abd.hpp:1532-1555.
#include "common.cuh"
#include "dequant.cuh"
#include "epilogue.cuh"
#include "kernels.h"

namespace ckb {
namespace {

// one thread = 16 consecutive k of one channel n (16 source bytes for W8, 8 for W4 handled as 32 k per thread)
template <int WF, bool F16>
__global__ void __launch_bounds__(256)
wq_dequant_w8_kernel(const uint8_t* __restrict__ b, long long ldb, const uint16_t* __restrict__ scale, int N, int K,
                     uint16_t* __restrict__ out)
{
    const int chunks_per_row = (K + 15) / 16;
    const long long total    = (long long)N * chunks_per_row;
    for(long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x)
    {
        const int n  = int(t / chunks_per_row);
        const int k0 = int(t % chunks_per_row) * 16;
        const uint4 c = *reinterpret_cast<const uint4*>(b + (long long)n * ldb + k0); // ldb % 16 == 0, K padded reads stay in-row
        const uint16_t sb = (WF == WF_W8_PASS) ? uint16_t(0) : scale[n];
        const float sf    = bf16_bits_to_f32(sb);
        const uint32_t s2 = uint32_t(sb) | (uint32_t(sb) << 16);
        const uint32_t w4[4] = {c.x, c.y, c.z, c.w};
        uint32_t o[8];
#pragma unroll
        for(int i = 0; i < 4; ++i)
        {
            if(WF == WF_W8_MUL) dequant_s8x4_mul_bf16(w4[i], sf, o + 2 * i);
            else if(WF == WF_W8_FPAINTB) dequant_u8x4_fpaintb(w4[i], s2, o + 2 * i);
            else dequant_s8x4_pass<F16>(w4[i], o + 2 * i);
        }
        uint16_t* dst = out + (long long)n * K + k0;
        if(k0 + 16 <= K)
        {
            reinterpret_cast<uint4*>(dst)[0] = make_uint4(o[0], o[1], o[2], o[3]);
            reinterpret_cast<uint4*>(dst)[1] = make_uint4(o[4], o[5], o[6], o[7]);
        }
        else
        {
            for(int i = 0; i < K - k0; ++i) dst[i] = uint16_t(o[i / 2] >> (16 * (i & 1)));
        }
    }
}

// pre-packed int4 (layout of ckb_prepack_b_int4: [n16 group][k64 block][lane][4 words]): one thread = one 16-byte
// chunk = 16 consecutive k of channels r_lo and r_hi = r_lo + 8
template <bool F16>
__global__ void __launch_bounds__(256)
wq_dequant_w4_kernel(const uint8_t* __restrict__ packed, const uint16_t* __restrict__ scale, const uint8_t* __restrict__ zero,
                     int zero_const, int N, int K, int k_blocks, long long total_chunks, uint16_t* __restrict__ out)
{
    for(long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x; c < total_chunks; c += (long long)gridDim.x * blockDim.x)
    {
        const int lane     = int(c & 31);
        const long long bl = c >> 5;
        const int kb = int(bl % k_blocks), ng = int(bl / k_blocks);
        const int g = lane >> 2, t = lane & 3;
        const int r_lo = 16 * ng + g, r_hi = r_lo + 8;
        const int k0 = 64 * kb + 16 * t;
        if(r_lo >= N || k0 >= K) continue;
        const uint4 v        = *reinterpret_cast<const uint4*>(packed + c * 16);
        const long long srow = (long long)(kb >> 1) * N;
        const bool hi_ok     = r_hi < N;
        const uint32_t sl = scale[srow + r_lo], sh = hi_ok ? scale[srow + r_hi] : 0;
        const int zl = zero ? int(zero[srow + r_lo]) : zero_const;
        const int zh = (zero && hi_ok) ? int(zero[srow + r_hi]) : zero_const;
        const uint32_t magic = F16 ? 0x6400u : 0x4300u;
        const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
        uint32_t lo[8], hi[8]; // 16 k of each row as packed pairs
#pragma unroll
        for(int s = 0; s < 4; ++s)
        {
            uint32_t o[4];
            dequant_w4_word_scaled2<F16>(w4[s], (magic + zl) * 0x10001u, (magic + zh) * 0x10001u, sl | (sl << 16), sh | (sh << 16), o);
            lo[2 * s] = o[0]; lo[2 * s + 1] = o[2];
            hi[2 * s] = o[1]; hi[2 * s + 1] = o[3];
        }
        uint16_t* dlo = out + (long long)r_lo * K + k0;
        uint16_t* dhi = out + (long long)r_hi * K + k0;
        if(k0 + 16 <= K && (K % 8) == 0)
        {
            reinterpret_cast<uint4*>(dlo)[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            reinterpret_cast<uint4*>(dlo)[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
            if(hi_ok)
            {
                reinterpret_cast<uint4*>(dhi)[0] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                reinterpret_cast<uint4*>(dhi)[1] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
            }
        }
        else
        {
            for(int i = 0; i < min(16, K - k0); ++i)
            {
                dlo[i] = uint16_t(lo[i / 2] >> (16 * (i & 1)));
                if(hi_ok) dhi[i] = uint16_t(hi[i / 2] >> (16 * (i & 1)));
            }
        }
    }
}

} // namespace

int plan_wq_prefill(const ckb_gemm_problem& p, ckb_gemm_plan& plan)
{
    // dense operand in the workspace, then the SS GEMM; TMA descriptors are encoded at run time (they need the workspace)
    plan.workspace_bytes = (uint64_t(p.N) * p.K * 2 + 255) & ~uint64_t(255);
    const int num_m_blk = (p.M + 127) / 128, num_n_blk = (p.N + 255) / 256;
    const int sms       = device_info().sm_count > 0 ? device_info().sm_count : 148;
    plan.grid           = std::max(1, std::min(sms, num_m_blk * num_n_blk));
    plan.block          = 256;
    plan.split_units    = 1;
    static const char* wfn[] = {"w4g128", "w8mul", "w8fpaintb", "w8pass"};
    const int wf = p.b_op == CKB_BOP_GROUP_RTN ? 0 : (p.b_op == CKB_BOP_MULTIPLY ? 1 : (p.b_op == CKB_BOP_FPAINTB ? 2 : 3));
    snprintf(plan.kernel_name, sizeof plan.kernel_name, "wq_prefill_%s_%s_dequant+gemm_ss_tcgen05_128x256x64", wfn[wf],
             p.a_dtype == CKB_F16 ? "f16" : "bf16");
    return CKB_OK;
}

int run_wq_prefill(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, void* workspace, cudaStream_t stream)
{
    if(!workspace) return set_error(CKB_ERR_WORKSPACE, "prefill path needs %llu bytes of workspace (dequantised weight)",
                                    (unsigned long long)plan.workspace_bytes);
    uint16_t* bdq    = static_cast<uint16_t*>(workspace);
    const bool f16   = p.a_dtype == CKB_F16;
    const int blocks = 148 * 8;
    if(p.b_op == CKB_BOP_GROUP_RTN)
    {
        const int n_groups = (p.N + 15) / 16, k_tiles = (p.K + 63) / 64;
        const long long chunks = (long long)n_groups * k_tiles * 32;
        if(f16)
            wq_dequant_w4_kernel<true><<<blocks, 256, 0, stream>>>(static_cast<const uint8_t*>(p.b), static_cast<const uint16_t*>(p.scale),
                                                                   static_cast<const uint8_t*>(p.zero), p.zero_const, p.N, p.K, k_tiles, chunks, bdq);
        else
            wq_dequant_w4_kernel<false><<<blocks, 256, 0, stream>>>(static_cast<const uint8_t*>(p.b), static_cast<const uint16_t*>(p.scale),
                                                                    static_cast<const uint8_t*>(p.zero), p.zero_const, p.N, p.K, k_tiles, chunks, bdq);
    }
    else
    {
        const uint8_t* b = static_cast<const uint8_t*>(p.b);
        const uint16_t* s = static_cast<const uint16_t*>(p.scale);
        if(p.b_op == CKB_BOP_MULTIPLY) wq_dequant_w8_kernel<WF_W8_MUL, false><<<blocks, 256, 0, stream>>>(b, p.ldb, s, p.N, p.K, bdq);
        else if(p.b_op == CKB_BOP_FPAINTB) wq_dequant_w8_kernel<WF_W8_FPAINTB, true><<<blocks, 256, 0, stream>>>(b, p.ldb, s, p.N, p.K, bdq);
        else if(f16) wq_dequant_w8_kernel<WF_W8_PASS, true><<<blocks, 256, 0, stream>>>(b, p.ldb, s, p.N, p.K, bdq);
        else wq_dequant_w8_kernel<WF_W8_PASS, false><<<blocks, 256, 0, stream>>>(b, p.ldb, s, p.N, p.K, bdq);
    }
    CKB_CUDA_CHECK(cudaGetLastError());
    ckb_gemm_problem q = p;
    q.b = bdq; q.b_dtype = p.a_dtype; q.b_layout = CKB_COL; q.ldb = p.K; q.b_op = CKB_BOP_PASSTHROUGH;
    q.scale = nullptr; q.zero = nullptr;
    ckb_gemm_plan sub;
    memset(&sub, 0, sizeof sub);
    if(int rc = plan_gemm_ss(q, sub)) return rc;
    return run_gemm_ss(q, sub, stream);
}

} // namespace ckb
