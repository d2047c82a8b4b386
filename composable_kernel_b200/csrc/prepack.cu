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

// [EXT] 4-bit codes (one per byte, logical B[k,n]) -> tile-major packed layout of the W4A16 kernels:
//   tile (nt, kt) of 128 channels x 128 k = 8192 bytes at ((nt * k_tiles) + kt) * 8192
//   inside:  [chunk j = 0..3][row r = 0..127][word i = 0..3]  (a warp reads 32 rows x 16 B = 512 B contiguous)
//   word (j,r,i) holds k = 32 j + 8 i + {0..7} of channel r with nibbles ordered
//   [k0,k2,k4,k6,k1,k3,k5,k7] so that (w >> 4p) & 0x000F000F yields the pair (k_2p, k_2p+1).
__global__ void prepack_int4_kernel(const uint8_t* __restrict__ src, long long sk, long long sn, int N, int K,
                                    int k_tiles, long long total_words, uint32_t* __restrict__ dst)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for(long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < total_words; w += stride)
    {
        const int i        = int(w & 3);
        const int r        = int((w >> 2) & 127);
        const int j        = int((w >> 9) & 3);
        const long long t  = w >> 11;
        const int kt       = int(t % k_tiles);
        const int nt       = int(t / k_tiles);
        const int n        = nt * 128 + r;
        const int k0       = kt * 128 + 32 * j + 8 * i;
        uint32_t word      = 0;
        if(n < N)
        {
            const int pos[8] = {0, 4, 1, 5, 2, 6, 3, 7}; // nibble position of k0..k7
#pragma unroll
            for(int e = 0; e < 8; ++e)
            {
                const int k = k0 + e;
                if(k < K) word |= uint32_t(src[k * sk + n * sn] & 0xF) << (4 * pos[e]);
            }
        }
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

} // namespace

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
    const int n_tiles = (N + 127) / 128, k_tiles = (K + 127) / 128;
    const long long words = (long long)n_tiles * k_tiles * 2048;
    if(words == 0) return CKB_OK;
    const int blocks = int(std::min<long long>((words + 255) / 256, 148 * 16));
    prepack_int4_kernel<<<blocks, 256, 0, stream>>>(static_cast<const uint8_t*>(src), sk, sn, N, K, k_tiles, words,
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
