// Microbenchmark: per-SM throughput of (a) legacy mma.sync m16n8k16 bf16, (b) the int4 -> bf16 dequant
// instruction mix (shift/lop3/sub.bf16x2/mul.bf16x2), as a function of resident warps per SM.
// Prints cycles per instruction group so kernel designs can be budgeted against the HBM rate.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma16816(float* d, const uint32_t* a, const uint32_t* b)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int NACC>
__global__ void k_mma(int iters, long long* out, float* sink)
{
    float acc[NACC][4];
    uint32_t a[4] = {threadIdx.x, threadIdx.x * 3u, 0x3f803f80u, 0x40004000u}, b[2] = {0x3f803f80u, threadIdx.x};
#pragma unroll
    for(int i = 0; i < NACC; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
    __syncthreads();
    long long t0 = clock64();
    for(int it = 0; it < iters; ++it)
    {
#pragma unroll
        for(int i = 0; i < NACC; ++i) mma16816(acc[i], a, b);
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for(int i = 0; i < NACC; ++i) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
    if(s == 12345.f) sink[0] = s;
    if(threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
    return r;
}
__device__ __forceinline__ uint32_t hsub2_bf16(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ uint32_t hmul2_bf16(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

// mode 0: shift+lop3+sub+mul (scale folded per weight); mode 1: shift+lop3+sub only (scale on partial sums)
template <int MODE>
__global__ void k_dq(int iters, long long* out, uint32_t* sink, uint32_t seed, uint32_t zm, uint32_t sc)
{
    uint32_t w[8];
#pragma unroll
    for(int i = 0; i < 8; ++i) w[i] = seed * (threadIdx.x + 1 + i);
    uint32_t x = 0;
    __syncthreads();
    long long t0 = clock64();
    for(int it = 0; it < iters; ++it)
    {
#pragma unroll
        for(int j = 0; j < 8; ++j)
        {
#pragma unroll
            for(int i = 0; i < 4; ++i)
            {
                uint32_t t = lop3<0xEA>(w[j] >> (4 * i), 0x000F000Fu, 0x43004300u);
                t          = hsub2_bf16(t, zm);
                if(MODE == 0) t = hmul2_bf16(t, sc);
                x ^= t; // stands in for the consumer (an MMA operand register); xor is alu-pipe
            }
            w[j] += x & 1; // keep the loop from being hoisted
        }
    }
    long long t1 = clock64();
    if(x == 0x12345u) sink[0] = x;
    if(threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

int main()
{
    long long* out; cudaMallocManaged(&out, 1024 * 8);
    float* sink; cudaMalloc(&sink, 64);
    const int iters = 2000;
    for(int warps : {4, 8, 16, 32})
    {
        k_mma<4><<<148, warps * 32>>>(iters, out, sink); cudaDeviceSynchronize();
        printf("mma.sync m16n8k16 bf16, %2d warps/SM, 4 indep acc: %.2f clk per MMA per SMSP (%.1f weights/clk/SM as A-operand)\n", warps,
               double(out[0]) / (iters * 4.0 * warps / 4.0), 256.0 * 4 / (double(out[0]) / (iters * 4.0 * warps / 4.0)));
        k_mma<8><<<148, warps * 32>>>(iters, out, sink); cudaDeviceSynchronize();
        printf("mma.sync m16n8k16 bf16, %2d warps/SM, 8 indep acc: %.2f clk per MMA per SMSP\n", warps,
               double(out[0]) / (iters * 8.0 * warps / 4.0));
    }
    for(int warps : {4, 8, 16, 32})
    {
        k_dq<0><<<148, warps * 32>>>(iters, out, (uint32_t*)sink, 0x9E3779B9u, 0x43084308u, 0x3c003c00u); cudaDeviceSynchronize();
        double clk_per_word = double(out[0]) / (iters * 8.0 * warps / 4.0); // per SMSP per warp-word (=256 weights)
        printf("dequant w4->bf16 (sub+mul), %2d warps/SM: %.2f clk per warp-word per SMSP -> %.0f clk per 128x128 unit per SM\n", warps,
               clk_per_word, clk_per_word * 64 / 4.0);
        k_dq<1><<<148, warps * 32>>>(iters, out, (uint32_t*)sink, 0x9E3779B9u, 0x43084308u, 0x3c003c00u); cudaDeviceSynchronize();
        clk_per_word = double(out[0]) / (iters * 8.0 * warps / 4.0);
        printf("dequant w4->bf16 (sub only), %2d warps/SM: %.2f clk per warp-word per SMSP -> %.0f clk per 128x128 unit per SM\n", warps,
               clk_per_word, clk_per_word * 64 / 4.0);
    }
    cudaError_t e = cudaGetLastError();
    if(e != cudaSuccess) printf("error %s\n", cudaGetErrorString(e));
    return 0;
}
