// Microbenchmark: legacy mma.sync int8 (IMMA.16832.U8.S8) throughput on sm_100a, and the cost of an int4 -> int8 unpack + IMMA
// loop (the candidate W4A16 decode path with block-float int8 activations).  clk per instruction per SM sub-partition.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ void imma(int* d, const uint32_t* a, uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <int NACC>
__global__ void k_imma(int iters, long long* out, int* sink)
{
    int acc[NACC][4];
    uint32_t a[4] = {threadIdx.x, threadIdx.x * 3u, 0x01020304u, 0x04030201u};
#pragma unroll
    for(int i = 0; i < NACC; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0;
    __syncthreads();
    long long t0 = clock64();
    for(int it = 0; it < iters; ++it)
#pragma unroll
        for(int i = 0; i < NACC; ++i) imma(acc[i], a, 0x01010101u, threadIdx.x);
    long long t1 = clock64();
    int s = 0;
#pragma unroll
    for(int i = 0; i < NACC; ++i) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
    if(s == 12345) sink[0] = s;
    if(threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}
// unpack loop: LDS.128 (4 words = 32 codes per lane) -> 8 registers of int8x4 (2 lop3 + 1 shf per word) -> 2 IMMA per limb
template <int G, int LIMBS>
__global__ void __launch_bounds__(512, 2) k_loop(int iters, long long* out, int* sink)
{
    extern __shared__ uint4 sm[];
    for(int i = threadIdx.x; i < 8 * 16 * 32; i += blockDim.x) sm[i] = make_uint4(i * 2654435761u, i * 40503u, i ^ 0x5bd1e995u, i + 77u);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int acc[G][LIMBS][4];
#pragma unroll
    for(int g = 0; g < G; ++g)
#pragma unroll
        for(int l = 0; l < LIMBS; ++l) acc[g][l][0] = acc[g][l][1] = acc[g][l][2] = acc[g][l][3] = 0;
    uint32_t b[LIMBS][4];
#pragma unroll
    for(int l = 0; l < LIMBS; ++l)
#pragma unroll
        for(int i = 0; i < 4; ++i) b[l][i] = 0x01020304u * (lane + i + l);
    long long t0 = clock64();
    for(int it = 0; it < iters; ++it)
    {
#pragma unroll
        for(int g = 0; g < G; ++g)
        {
            const uint4 wv = sm[(((g + it) & 7) * 16 + (warp & 15)) * 32 + lane];
            const uint32_t w4[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
            for(int h = 0; h < 2; ++h)
            {
                uint32_t a[4];
                a[0] = w4[2 * h] & 0x0F0F0F0Fu; a[1] = (w4[2 * h] >> 4) & 0x0F0F0F0Fu;
                a[2] = w4[2 * h + 1] & 0x0F0F0F0Fu; a[3] = (w4[2 * h + 1] >> 4) & 0x0F0F0F0Fu;
#pragma unroll
                for(int l = 0; l < LIMBS; ++l) imma(acc[g][l], a, b[l][2 * h], b[l][2 * h + 1]);
            }
        }
    }
    long long t1 = clock64();
    int s = 0;
#pragma unroll
    for(int g = 0; g < G; ++g)
#pragma unroll
        for(int l = 0; l < LIMBS; ++l) s += acc[g][l][0] + acc[g][l][1] + acc[g][l][2] + acc[g][l][3];
    if(s == 12345) sink[0] = s;
    if(threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}
template <int G, int LIMBS>
void loop(long long* out, int* sink)
{
    const int iters = 400;
    cudaFuncSetAttribute(k_loop<G, LIMBS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 16 * 32 * 16);
    for(int warps : {8, 16})
    {
        k_loop<G, LIMBS><<<148, warps * 32, 8 * 16 * 32 * 16>>>(iters, out, sink);
        cudaError_t e = cudaDeviceSynchronize();
        if(e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return; }
        printf("int4->int8 unpack + IMMA, G=%d limbs=%d %2d warps: %.1f clk per warp-word (8 weights x 32 lanes) per SMSP (HBM budget 22.9)\n", G, LIMBS, warps,
               double(out[0]) / (iters * (warps / 4.0) * G * 4.0));
    }
}
int main()
{
    long long* out; cudaMallocManaged(&out, 1024 * 8);
    int* sink; cudaMalloc(&sink, 64);
    const int iters = 2000;
    for(int warps : {4, 8, 16})
    {
        k_imma<4><<<148, warps * 32>>>(iters, out, sink); cudaDeviceSynchronize();
        printf("IMMA.16832.U8.S8 %2d warps/SM, 4 indep acc: %.2f clk per IMMA per SMSP\n", warps, double(out[0]) / (iters * 4.0 * warps / 4.0));
    }
    loop<2, 1>(out, sink); loop<2, 2>(out, sink); loop<2, 3>(out, sink); loop<4, 2>(out, sink);
    return 0;
}
