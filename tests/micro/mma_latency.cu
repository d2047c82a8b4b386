// Microbenchmark: cost of back-to-back tcgen05.mma (M=128, K=16, small N) on one SM.
//   mode bit0: 0 = A from smem (SS), 1 = A from TMEM (TS)
//   ndst: number of distinct accumulators rotated over (1 = fully dependent chain)
// prints cycles per MMA measured from first issue to commit-barrier completion.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../composable_kernel_b200/csrc/ptx.cuh"
using namespace ckb;

template <int N, int TS, int NDST>
__global__ void k(int nmma, long long* out)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar = base + 65536, slot = bar + 16;
    if(threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_mbar_init(); }
    if(threadIdx.x < 32) { ptx::tmem_alloc<512>(slot); ptx::tmem_relinquish(); }
    ptx::tcgen05_fence_before_sync(); __syncthreads(); ptx::tcgen05_fence_after_sync();
    uint32_t tb; asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tb) : "r"(slot));
    if(threadIdx.x == 0)
    {
        const uint32_t idesc = ptx::make_idesc(1, 1, 1, 128, N);
        const uint64_t da = ptx::make_smem_desc_sw128(base, 1024);
        const uint64_t db = ptx::make_smem_desc_sw128(base + 32768, 1024);
        for(int rep = 0; rep < 3; ++rep)
        {
            long long t0 = clock64();
            for(int i0 = 0; i0 < nmma; i0 += 8)
            {
#pragma unroll
                for(int i = 0; i < 8; ++i)
                {
                    const uint32_t d = tb + 256 + (i % NDST) * 32 * ((N + 31) / 32);
                    if(TS) ptx::mma_f16_ts(d, tb + i * 8, db + 2 * (i % 4), idesc, 1);
                    else   ptx::mma_f16_ss(d, da + 2 * (i % 4), db + 2 * (i % 4), idesc, 1);
                }
            }
            long long t1 = clock64();
            ptx::mma_commit(bar);
            ptx::mbar_wait(bar, rep & 1);
            long long t2 = clock64();
            out[0] = t1 - t0; out[1] = t2 - t0;
        }
    }
    __syncthreads();
    if(threadIdx.x < 32) ptx::tmem_dealloc<512>(tb);
}

template <int N, int TS, int NDST>
void run(long long* out)
{
    const int nmma = 256;
    cudaFuncSetAttribute(k<N, TS, NDST>, cudaFuncAttributeMaxDynamicSharedMemorySize, 80000);
    k<N, TS, NDST><<<1, 128, 80000>>>(nmma, out);
    cudaError_t e = cudaDeviceSynchronize();
    if(e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
    printf("%s N=%3d ndst=%d: issue %.1f clk/mma, complete %.1f clk/mma\n", TS ? "TS" : "SS", N, NDST,
           double(out[0]) / nmma, double(out[1]) / nmma);
}

int main()
{
    long long* out; cudaMallocManaged(&out, 16);
    run<16, 0, 1>(out); run<16, 0, 4>(out); run<64, 0, 1>(out); run<128, 0, 1>(out); run<256, 0, 1>(out);
    run<16, 1, 1>(out); run<16, 1, 4>(out); run<32, 1, 1>(out); run<64, 1, 1>(out); run<128, 1, 1>(out); run<256, 1, 1>(out);
    return 0;
}
