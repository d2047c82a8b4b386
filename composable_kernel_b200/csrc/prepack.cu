// prepack.cu -- one-time weight layout transforms (excluded from GEMM timing, as the reference
// excludes UnsignedWeightPreprocessor, example/64_fpAintB_gemm/common.hpp:56-83).
#include "common.cuh"
#include "kernels.h"

namespace ckb {
namespace {

// int8 -> biased uint8: u8 = int8 + 128  ==  flip the sign bit
__global__ void prepack_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, long long n)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long n16    = n / 16;
    for(long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride)
    {
        uint4 v = reinterpret_cast<const uint4*>(src)[i];
        v.x ^= 0x80808080u; v.y ^= 0x80808080u; v.z ^= 0x80808080u; v.w ^= 0x80808080u;
        reinterpret_cast<uint4*>(dst)[i] = v;
    }
    for(long long i = n16 * 16 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        dst[i] = src[i] ^ 0x80;
}

// [EXT] 4-bit codes (one per byte, logical B[k,n]) -> fragment-major packed layout of the W4A16 kernels:
//   [n16 group ng][k64 block kb][lane = 4 g + t][word s = 0..3]   (512 B per (ng, kb); a warp reads it with one LDS.128)
//   word (ng, kb, lane, s) is the mma.sync m16n8k16 A fragment of rows r_lo = 16 ng + g, r_hi = r_lo + 8 and the four
//   k values k_i = 64 kb + 16 t + 4 s + i (the kernel pairs them with the same k of the activations):
//     nibble 0 = (r_lo,k0)  4 = (r_lo,k1)  1 = (r_hi,k0)  5 = (r_hi,k1)  2 = (r_lo,k2)  6 = (r_lo,k3)  3 = (r_hi,k2)  7 = (r_hi,k3)
//   so that (w >> 4p) & 0x000F000F, p = 0..3, yields the fragment registers {a0a1, a2a3, a4a5, a6a7}.
//   Codes outside N x K are 0 (their activations are zero-filled by TMA).
__global__ void prepack_int4_kernel(const uint8_t* __restrict__ src, long long sk, long long sn, int N, int K,
                                    int k_blocks, long long total_words, uint32_t* __restrict__ dst)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for(long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < total_words; w += stride)
    {
        const int s        = int(w & 3);
        const int lane     = int((w >> 2) & 31);
        const long long bl = w >> 7;
        const int kb       = int(bl % k_blocks);
        const int ng       = int(bl / k_blocks);
        const int g = lane >> 2, t = lane & 3;
        const int k0       = 64 * kb + 16 * t + 4 * s;
        const int r_lo = 16 * ng + g, r_hi = r_lo + 8;
        auto code = [&](int r, int k) -> uint32_t {
            return (r < N && k < K) ? uint32_t(src[k * sk + r * sn] & 0xF) : 0u;
        };
        uint32_t word = 0;
        word |= code(r_lo, k0 + 0) << 0;
        word |= code(r_lo, k0 + 1) << 16;
        word |= code(r_hi, k0 + 0) << 4;
        word |= code(r_hi, k0 + 1) << 20;
        word |= code(r_lo, k0 + 2) << 8;
        word |= code(r_lo, k0 + 3) << 24;
        word |= code(r_hi, k0 + 2) << 12;
        word |= code(r_hi, k0 + 3) << 28;
        dst[w] = word;
    }
}

__global__ void transpose_i8_kernel(const uint8_t* __restrict__ src, long long ld_src, int rows, int cols,
                                    uint8_t* __restrict__ dst, long long ld_dst)
{
    __shared__ uint8_t tile[64][65];
    const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
    for(int i = threadIdx.y; i < 64; i += blockDim.y)
    {
        const int r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < rows && c < cols) ? src[(long long)r * ld_src + c] : 0;
    }
    __syncthreads();
    for(int i = threadIdx.y; i < 64; i += blockDim.y)
    {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if(c < cols && r < rows) dst[(long long)c * ld_dst + r] = tile[threadIdx.x][i];
    }
}

// 16-bit elements: dst[c * ld_dst + r] = src[r * ld_src + c]  (column-major A -> row-major A, per call)
__global__ void transpose_16_kernel(const uint16_t* __restrict__ src, long long ld_src, int rows, int cols, uint16_t* __restrict__ dst,
                                    long long ld_dst)
{
    __shared__ uint16_t tile[64][66];
    const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
    for(int i = threadIdx.y; i < 64; i += blockDim.y)
    {
        const int r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < rows && c < cols) ? src[(long long)r * ld_src + c] : uint16_t(0);
    }
    __syncthreads();
    for(int i = threadIdx.y; i < 64; i += blockDim.y)
    {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if(c < cols && r < rows) dst[(long long)c * ld_dst + r] = tile[threadIdx.x][i];
    }
}

} // namespace

int launch_transpose_16(const void* src, long long ld_src, int rows, int cols, void* dst, long long ld_dst, cudaStream_t stream)
{
    if(rows <= 0 || cols <= 0) return CKB_OK;
    dim3 grid((cols + 63) / 64, (rows + 63) / 64), block(64, 8);
    transpose_16_kernel<<<grid, block, 0, stream>>>(static_cast<const uint16_t*>(src), ld_src, rows, cols, static_cast<uint16_t*>(dst), ld_dst);
    CKB_CUDA_CHECK(cudaGetLastError());
    return CKB_OK;
}

int launch_prepack_u8(const void* src, void* dst, long long n, cudaStream_t stream)
{
    if(n <= 0) return CKB_OK;
    const int blocks = int(std::min<long long>((n / 16 + 255) / 256 + 1, 148 * 8));
    prepack_u8_kernel<<<blocks, 256, 0, stream>>>(static_cast<const uint8_t*>(src), static_cast<uint8_t*>(dst), n);
    CKB_CUDA_CHECK(cudaGetLastError());
    return CKB_OK;
}

int launch_prepack_int4(const void* src, long long sk, long long sn, int N, int K, void* dst, cudaStream_t stream)
{
    const int n_groups = (N + 15) / 16, k_blocks = (K + 63) / 64;
    const long long words = (long long)n_groups * k_blocks * 128;
    if(words == 0) return CKB_OK;
    const int blocks = int(std::min<long long>((words + 255) / 256, 148 * 16));
    prepack_int4_kernel<<<blocks, 256, 0, stream>>>(static_cast<const uint8_t*>(src), sk, sn, N, K, k_blocks, words,
                                                    static_cast<uint32_t*>(dst));
    CKB_CUDA_CHECK(cudaGetLastError());
    return CKB_OK;
}

int launch_transpose_i8(const void* src, long long ld_src, int rows, int cols, void* dst, long long ld_dst,
                        cudaStream_t stream)
{
    if(rows <= 0 || cols <= 0) return CKB_OK;
    dim3 grid((cols + 63) / 64, (rows + 63) / 64), block(64, 8);
    transpose_i8_kernel<<<grid, block, 0, stream>>>(static_cast<const uint8_t*>(src), ld_src, rows, cols,
                                                    static_cast<uint8_t*>(dst), ld_dst);
    CKB_CUDA_CHECK(cudaGetLastError());
    return CKB_OK;
}

} // namespace ckb
