// Does an mbarrier.try_wait issued right after tcgen05.commit stall until the MMAs finish?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../composable_kernel_b200/csrc/ptx.cuh"
using namespace ckb;

template <int N, int NMMA>
__global__ void k(long long* out)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar = base + 65536, bar2 = bar + 8, slot = bar + 16;
    volatile int* flag = reinterpret_cast<volatile int*>(smem_raw + 70000);
    if(threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::mbar_init(bar2, 1); ptx::fence_mbar_init(); *flag = 1; }
    if(threadIdx.x < 32) { ptx::tmem_alloc<512>(slot); ptx::tmem_relinquish(); }
    ptx::tcgen05_fence_before_sync(); __syncthreads(); ptx::tcgen05_fence_after_sync();
    uint32_t tb; asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tb) : "r"(slot));
    if(threadIdx.x == 0)
    {
        ptx::mbar_arrive(bar2); // bar2 phase 0 complete
        const uint32_t idesc = ptx::make_idesc(1, 1, 1, 128, N);
        const uint64_t db = ptx::make_smem_desc_sw128(base + 32768, 1024);
        for(int rep = 0; rep < 3; ++rep)
        {
            long long t0 = clock64();
#pragma unroll
            for(int i = 0; i < NMMA; ++i) ptx::mma_f16_ts(tb + 256, tb + (i % 8) * 8, db + 2 * (i % 4), idesc, 1);
            long long t1 = clock64();
            ptx::mma_commit(bar);
            long long t2 = clock64();
            int f = *flag;                       // plain shared load after the commit
            long long t3 = clock64();
            bool ok = ptx::mbar_try_wait(bar2, 0); // unrelated, already-complete barrier
            long long t4 = clock64();
            ptx::mbar_wait(bar, rep & 1);
            long long t5 = clock64();
            out[0] = t1 - t0; out[1] = t2 - t0; out[2] = t3 - t0 + (f & 0); out[3] = t4 - t0 + (ok ? 0 : 1000000); out[4] = t5 - t0;
        }
    }
    __syncthreads();
    if(threadIdx.x < 32) ptx::tmem_dealloc<512>(tb);
}

template <int N, int NMMA>
void run(long long* out)
{
    cudaFuncSetAttribute(k<N, NMMA>, cudaFuncAttributeMaxDynamicSharedMemorySize, 80000);
    k<N, NMMA><<<1, 128, 80000>>>(out);
    cudaError_t e = cudaDeviceSynchronize();
    if(e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
    printf("N=%3d nmma=%2d: mmas issued %lld | commit issued %lld | ld.shared done %lld | try_wait(other) done %lld | own barrier done %lld\n",
           N, NMMA, out[0], out[1], out[2], out[3], out[4]);
}

int main()
{
    long long* out; cudaMallocManaged(&out, 64);
    run<16, 8>(out); run<16, 32>(out); run<256, 8>(out); run<256, 16>(out);
    return 0;
}
