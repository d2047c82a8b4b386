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

// pre-packed int4: one thread = one 16-byte chunk = 32 k of one channel (layout of ckb_prepack_b_int4)
template <bool F16>
__global__ void __launch_bounds__(256)
wq_dequant_w4_kernel(const uint8_t* __restrict__ packed, const uint16_t* __restrict__ scale, const uint8_t* __restrict__ zero,
                     int zero_const, int N, int K, int k_tiles, long long total_chunks, uint16_t* __restrict__ out)
{
    for(long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total_chunks; t += (long long)gridDim.x * blockDim.x)
    {
        const int r        = int(t & 127);
        const int j        = int((t >> 7) & 3);
        const long long tl = t >> 9;
        const int kt = int(tl % k_tiles), nt = int(tl / k_tiles);
        const int n = nt * 128 + r, k0 = kt * 128 + j * 32;
        if(n >= N || k0 >= K) continue;
        const uint4 c     = *reinterpret_cast<const uint4*>(packed + t * 16);
        const uint32_t sb = scale[(long long)kt * N + n];
        const uint32_t s2 = sb | (sb << 16);
        const int z       = zero ? int(zero[(long long)kt * N + n]) : zero_const;
        const uint32_t zm = F16 ? (0x6400u + z) * 0x10001u : (0x4300u + z) * 0x10001u;
        uint32_t o[16];
        dequant_w4_word_scaled<F16>(c.x, zm, s2, o + 0);
        dequant_w4_word_scaled<F16>(c.y, zm, s2, o + 4);
        dequant_w4_word_scaled<F16>(c.z, zm, s2, o + 8);
        dequant_w4_word_scaled<F16>(c.w, zm, s2, o + 12);
        uint16_t* dst = out + (long long)n * K + k0;
        if(k0 + 32 <= K && (K % 8) == 0)
        {
#pragma unroll
            for(int i = 0; i < 4; ++i) reinterpret_cast<uint4*>(dst)[i] = make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
        }
        else
        {
            for(int i = 0; i < min(32, K - k0); ++i) dst[i] = uint16_t(o[i / 2] >> (16 * (i & 1)));
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
    const int wf = p.b_op == CKB_BOP_GROUP ? 0 : (p.b_op == CKB_BOP_MULTIPLY ? 1 : (p.b_op == CKB_BOP_FPAINTB ? 2 : 3));
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
    if(p.b_op == CKB_BOP_GROUP)
    {
        const int n_tiles = (p.N + 127) / 128, k_tiles = (p.K + 127) / 128;
        const long long chunks = (long long)n_tiles * k_tiles * 512;
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
