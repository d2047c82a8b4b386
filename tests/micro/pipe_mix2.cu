// Microbenchmark (round 2): which SM pipes the int4 -> 16-bit unpack instructions use on sm_100a, whether the alu and
// fma pipes issue in parallel, and what the unpack + HMMA loop costs per warp-word for several instruction mixes and
// warp counts.  Results: profiles/r2_micro_pipe_mix2.txt.  clk = SM cycles per warp-instruction per SM sub-partition.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define NCH 8
enum Op { LOP3, SHF, PRMT, IMAD, MULHI, MULWIDE, HADD2BF, HFMA2BF, HADD2F16, FFMA, FMULI, LOP3_HADD, LOP3_IMAD, LOP3_MULHI, LOP3_FFMA,
          SHF_LOP3, LOP3_HFMA, IADD, NOPS };
static const char* op_name[] = {"lop3", "shf.r", "prmt", "mad.lo (IMAD)", "mul.hi.u32", "mul.wide.u32", "sub.bf16x2", "fma.bf16x2",
                                "sub.f16x2", "ffma", "fmul imm", "lop3 + sub.bf16x2 (pair)", "lop3 + mad.lo (pair)",
                                "lop3 + mul.hi (pair)", "lop3 + ffma (pair)", "shf + lop3 (pair)", "lop3 + fma.bf16x2 (pair)", "add.u32"};

template <int OP>
__global__ void k_rate(int iters, long long* out, uint32_t* sink, uint32_t seed)
{
    uint32_t r[NCH], q[NCH];
#pragma unroll
    for(int i = 0; i < NCH; ++i) { r[i] = seed * (threadIdx.x + 3 + i); q[i] = seed ^ (threadIdx.x * 7 + i); }
    const uint32_t c1 = seed | 0x000F000Fu, c2 = 0x43004300u ^ (seed & 1);
    __syncthreads();
    long long t0 = clock64();
    for(int it = 0; it < iters; ++it)
    {
#pragma unroll
        for(int rep = 0; rep < 4; ++rep)
#pragma unroll
            for(int i = 0; i < NCH; ++i)
            {
                if(OP == LOP3 || OP >= LOP3_HADD && OP != SHF_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(c1), "r"(c2));
                if(OP == SHF || OP == SHF_LOP3) asm volatile("shr.u32 %0, %0, 4;" : "+r"(r[i]));
                if(OP == SHF_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(q[i]) : "r"(c1), "r"(c2));
                if(OP == PRMT) asm volatile("prmt.b32 %0, %0, %1, 0x4140;" : "+r"(r[i]) : "r"(c2));
                if(OP == IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c1), "r"(c2));
                if(OP == LOP3_IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(q[i]) : "r"(c1), "r"(c2));
                if(OP == MULHI) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(c1));
                if(OP == LOP3_MULHI) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(q[i]) : "r"(c1));
                if(OP == MULWIDE) { uint64_t w; asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(r[i]), "r"(c1)); r[i] = uint32_t(w >> 32); }
                if(OP == HADD2BF) asm volatile("sub.rn.bf16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c2));
                if(OP == LOP3_HADD) asm volatile("sub.rn.bf16x2 %0, %0, %1;" : "+r"(q[i]) : "r"(c2));
                if(OP == HFMA2BF) asm volatile("fma.rn.bf16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c1), "r"(c2));
                if(OP == LOP3_HFMA) asm volatile("fma.rn.bf16x2 %0, %0, %1, %2;" : "+r"(q[i]) : "r"(c1), "r"(c2));
                if(OP == HADD2F16) asm volatile("sub.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c2));
                if(OP == FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(*(float*)&r[i]) : "f"(__uint_as_float(c1)), "f"(__uint_as_float(c2)));
                if(OP == LOP3_FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(*(float*)&q[i]) : "f"(__uint_as_float(c1)), "f"(__uint_as_float(c2)));
                if(OP == FMULI) asm volatile("mul.rn.f32 %0, %0, 0f3F800001;" : "+f"(*(float*)&r[i]));
                if(OP == IADD) asm volatile("add.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(c1));
            }
    }
    long long t1 = clock64();
    uint32_t x = 0;
#pragma unroll
    for(int i = 0; i < NCH; ++i) x ^= r[i] ^ q[i];
    if(x == 0x12345u) sink[0] = x;
    if(threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

template <int OP>
void rate(long long* out, uint32_t* sink)
{
    const int iters = 500;
    for(int warps : {4, 16})
    {
        k_rate<OP><<<148, warps * 32>>>(iters, out, sink, 0x9E3779B9u);
        cudaDeviceSynchronize();
        const int per_it = 4 * NCH * ((OP >= LOP3_HADD && OP != IADD) ? 2 : 1);
        printf("%-28s %2d warps/SM: %.2f clk per warp-instr per SMSP\n", op_name[OP], warps, double(out[0]) / (double(iters) * per_it * warps / 4.0));
    }
}

// ---------------------------------------------------------------------------------------------------------------------
template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
    return r;
}
__device__ __forceinline__ uint32_t hsub2(uint32_t a, uint32_t b) { uint32_t r; asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ uint32_t hsub2h(uint32_t a, uint32_t b) { uint32_t r; asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ uint32_t hfma2h(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
__device__ __forceinline__ uint32_t mulhi(uint32_t a, uint32_t b) { uint32_t r; asm("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ void mma_bf16(float* d, const uint32_t* a, uint32_t b0, uint32_t b1)
{
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_f16(float* d, const uint32_t* a, uint32_t b0, uint32_t b1)
{
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// unpack variants (one 32-bit word = 8 codes -> 4 registers of 2 x 16 bit)
//   0: bf16  3 shf + 4 lop3 + 4 sub            (round-1 kernel)
//   1: bf16  3 shf + 4 lop3                    (zero point folded into an activation sum)
//   2: bf16  3 mul.hi + 4 lop3 + 4 sub         (shifts on the fma pipe)
//   3: bf16  3 mul.hi + 4 lop3
//   4: fp16  1 shf + 4 lop3 + 2 sub + 2 fma    (x1 and x16 nibble positions)
//   5: fp16  1 shf + 4 lop3                    (activations pre-scaled for the x16 positions, zero point folded)
template <int V>
__device__ __forceinline__ void unpack(uint32_t w, uint32_t zm, uint32_t* a)
{
    if(V == 0 || V == 1)
    {
#pragma unroll
        for(int i = 0; i < 4; ++i)
        {
            uint32_t t = lop3<0xEA>(w >> (4 * i), 0x000F000Fu, 0x43004300u);
            a[i] = V == 0 ? hsub2(t, zm) : t;
        }
    }
    else if(V == 2 || V == 3)
    {
        const uint32_t w1 = mulhi(w, 1u << 28), w2 = mulhi(w, 1u << 24), w3 = mulhi(w, 1u << 20);
        const uint32_t ws[4] = {w, w1, w2, w3};
#pragma unroll
        for(int i = 0; i < 4; ++i)
        {
            uint32_t t = lop3<0xEA>(ws[i], 0x000F000Fu, 0x43004300u);
            a[i] = V == 2 ? hsub2(t, zm) : t;
        }
    }
    else
    {
        const uint32_t w8 = w >> 8;
        uint32_t t0 = lop3<0xEA>(w, 0x000F000Fu, 0x64006400u), t1 = lop3<0xEA>(w, 0x00F000F0u, 0x64006400u);
        uint32_t t2 = lop3<0xEA>(w8, 0x000F000Fu, 0x64006400u), t3 = lop3<0xEA>(w8, 0x00F000F0u, 0x64006400u);
        if(V == 4)
        {
            t0 = hsub2h(t0, zm); t2 = hsub2h(t2, zm);
            t1 = hfma2h(t1, 0x2C002C00u /* 1/16 */, zm ^ 0x80008000u); t3 = hfma2h(t3, 0x2C002C00u, zm ^ 0x80008000u);
        }
        a[0] = t0; a[1] = t1; a[2] = t2; a[3] = t3;
    }
}

// G groups per warp, KB 64-k blocks per iteration; every (group, block) = one LDS.128 + 4 x (unpack + HMMA); the partial sums of
// a block get the group scale (4 FFMA) -- the structure of the decode kernel's inner loop
template <int V, int G, int NT>
__global__ void __launch_bounds__(512, 1) k_loop(int iters, long long* out, float* sink, uint32_t zm, float scf)
{
    extern __shared__ uint4 sm[];
    for(int i = threadIdx.x; i < 8 * 16 * 32; i += blockDim.x) sm[i] = make_uint4(i * 2654435761u, i * 40503u, i ^ 0x5bd1e995u, i + 77u);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float acc[G][NT][4];
#pragma unroll
    for(int g = 0; g < G; ++g)
#pragma unroll
        for(int n = 0; n < NT; ++n) acc[g][n][0] = acc[g][n][1] = acc[g][n][2] = acc[g][n][3] = 0.f;
    uint32_t b[NT][8];
#pragma unroll
    for(int n = 0; n < NT; ++n)
#pragma unroll
        for(int i = 0; i < 8; ++i) b[n][i] = 0x3f803f80u + lane + i + n;
    long long t0 = clock64();
    for(int it = 0; it < iters; ++it)
    {
#pragma unroll
        for(int g = 0; g < G; ++g)
        {
            const uint4 wv = sm[(((g + it) & 7) * 16 + (warp & 15)) * 32 + lane];
            const uint32_t w4[4] = {wv.x, wv.y, wv.z, wv.w};
            float p[NT][4];
#pragma unroll
            for(int n = 0; n < NT; ++n) p[n][0] = p[n][1] = p[n][2] = p[n][3] = 0.f;
#pragma unroll
            for(int ks = 0; ks < 4; ++ks)
            {
                uint32_t a[4];
                unpack<V>(w4[ks], zm, a);
#pragma unroll
                for(int n = 0; n < NT; ++n)
                {
                    if(V >= 4) mma_f16(p[n], a, b[n][2 * ks], b[n][2 * ks + 1]);
                    else mma_bf16(p[n], a, b[n][2 * ks], b[n][2 * ks + 1]);
                }
            }
#pragma unroll
            for(int n = 0; n < NT; ++n)
            {
                acc[g][n][0] = fmaf(scf, p[n][0], acc[g][n][0]); acc[g][n][1] = fmaf(scf, p[n][1], acc[g][n][1]);
                acc[g][n][2] = fmaf(scf, p[n][2], acc[g][n][2]); acc[g][n][3] = fmaf(scf, p[n][3], acc[g][n][3]);
            }
        }
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for(int g = 0; g < G; ++g)
#pragma unroll
        for(int n = 0; n < NT; ++n) s += acc[g][n][0] + acc[g][n][1] + acc[g][n][2] + acc[g][n][3];
    if(s == 12345.f) sink[0] = s;
    if(threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

template <int V, int G, int NT>
void loop(long long* out, float* sink, const char* name)
{
    const int iters = 400;
    cudaFuncSetAttribute(k_loop<V, G, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 16 * 32 * 16);
    for(int warps : {8, 16})
    {
        k_loop<V, G, NT><<<148, warps * 32, 8 * 16 * 32 * 16>>>(iters, out, sink, 0x43084308u, 0.0078125f);
        cudaError_t e = cudaDeviceSynchronize();
        if(e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return; }
        const double clk_word = double(out[0]) / (iters * (warps / 4.0) * G * 4.0);
        printf("%-52s G=%d NT=%d %2d warps: %.1f clk per warp-word per SMSP (HBM budget 22.9)\n", name, G, NT, warps, clk_word);
    }
}

int main()
{
    long long* out; cudaMallocManaged(&out, 1024 * 8);
    uint32_t* sink; cudaMalloc(&sink, 64);
    rate<LOP3>(out, sink); rate<SHF>(out, sink); rate<PRMT>(out, sink); rate<IADD>(out, sink); rate<IMAD>(out, sink); rate<MULHI>(out, sink);
    rate<MULWIDE>(out, sink); rate<HADD2BF>(out, sink); rate<HFMA2BF>(out, sink); rate<HADD2F16>(out, sink); rate<FFMA>(out, sink);
    rate<FMULI>(out, sink); rate<LOP3_HADD>(out, sink); rate<LOP3_HFMA>(out, sink); rate<LOP3_IMAD>(out, sink); rate<LOP3_MULHI>(out, sink);
    rate<LOP3_FFMA>(out, sink); rate<SHF_LOP3>(out, sink);
    float* fs = (float*)sink;
    loop<0, 2, 1>(out, fs, "bf16 3shf+4lop3+4sub + hmma");
    loop<1, 2, 1>(out, fs, "bf16 3shf+4lop3 + hmma");
    loop<2, 2, 1>(out, fs, "bf16 3mulhi+4lop3+4sub + hmma");
    loop<3, 2, 1>(out, fs, "bf16 3mulhi+4lop3 + hmma");
    loop<4, 2, 1>(out, fs, "fp16 1shf+4lop3+2sub+2fma + hmma");
    loop<5, 2, 1>(out, fs, "fp16 1shf+4lop3 + hmma");
    loop<0, 4, 1>(out, fs, "bf16 3shf+4lop3+4sub + hmma");
    loop<1, 4, 1>(out, fs, "bf16 3shf+4lop3 + hmma");
    loop<3, 4, 1>(out, fs, "bf16 3mulhi+4lop3 + hmma");
    loop<0, 2, 2>(out, fs, "bf16 3shf+4lop3+4sub + 2 hmma");
    loop<1, 2, 2>(out, fs, "bf16 3shf+4lop3 + 2 hmma");
    loop<2, 2, 2>(out, fs, "bf16 3mulhi+4lop3+4sub + 2 hmma");
    loop<0, 4, 2>(out, fs, "bf16 3shf+4lop3+4sub + 2 hmma");
    loop<4, 2, 2>(out, fs, "fp16 1shf+4lop3+2sub+2fma + 2 hmma");
    cudaError_t e = cudaGetLastError();
    if(e != cudaSuccess) printf("error %s\n", cudaGetErrorString(e));
    return 0;
}
