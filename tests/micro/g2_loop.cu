// Microbenchmark (round 2): the W4 decode kernel's per-warp math loop on shared-memory-resident data, 8 warps per CTA,
// in the code shapes tried for wq_gemv2_kernel.cuh.  clk per warp-word (8 weights x 32 lanes) per SM sub-partition; the HBM
// budget is 22.9.  Results: profiles/r2_micro_g2_loop.txt.
//   VAR 0: groups sequential, one 8-HMMA chain per group, warp-uniform early exit on the last group (first gemv2 version)
//   VAR 1: no early exit (a slot with GW - 1 groups computes one dead group)
//   VAR 2: all groups' weight words loaded first, then ks-outer / group-inner (consecutive HMMAs hit different chains)
//   VAR 3: VAR 2 + two accumulation chains per group (even / odd ks)
//   VAR 4: VAR 2 + next scale group's weights prefetched into registers before this one's math (software pipeline)
//   VAR 5: VAR 1 with the zero point folded: unpack = shift + lop3 only (128 + q), activation sums of the scale group
//          from 8 extra HMMAs with an all-ones A fragment, correction s * (part - (128 + zp) * asum) on the fma pipe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint4 lds128(uint32_t addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr)
{
    uint16_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
    return v;
}
template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
    return r;
}
__device__ __forceinline__ uint32_t hsub2(uint32_t a, uint32_t b) { uint32_t r; asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ void mma(float* d, const uint32_t* a, uint32_t b0, uint32_t b1)
{
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void unpack(uint32_t w, uint32_t zm_lo, uint32_t zm_hi, uint32_t* a)
{
#pragma unroll
    for(int i = 0; i < 4; ++i) a[i] = hsub2(lop3<0xEA>(w >> (4 * i), 0x000F000Fu, 0x43004300u), (i & 1) ? zm_hi : zm_lo);
}

// one 128-k scale group: 2 64-k blocks x G groups; layout as in the kernel: weights [G][2 ksg][512 B], scales [ksg][16 G] bf16
template <int NT, int GW, int VAR>
__device__ __forceinline__ void scale_group(uint32_t a_base, uint32_t w_q, uint32_t s_q, uint32_t w_gstride, int G,
                                            const uint32_t (&a_addr)[2][NT][2], const bool (&a_ok)[NT], uint32_t zc,
                                            float (&acc)[GW][NT][4], uint4 (&wpre)[GW][2], uint32_t w_next)
{
    uint4 bf[2][NT][2];
#pragma unroll
    for(int b = 0; b < 2; ++b)
#pragma unroll
        for(int nt = 0; nt < NT; ++nt)
        {
            if(a_ok[nt]) { bf[b][nt][0] = lds128(a_base + a_addr[b][nt][0]); bf[b][nt][1] = lds128(a_base + a_addr[b][nt][1]); }
            else bf[b][nt][0] = bf[b][nt][1] = make_uint4(0u, 0u, 0u, 0u);
        }
    const uint32_t zm = (0x4300u + zc) * 0x10001u;
    if(VAR == 5)
    {
        float asum[NT][4];
#pragma unroll
        for(int nt = 0; nt < NT; ++nt) asum[nt][0] = asum[nt][1] = asum[nt][2] = asum[nt][3] = 0.f;
        const uint32_t ones[4] = {0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u};
#pragma unroll
        for(int b = 0; b < 2; ++b)
#pragma unroll
            for(int ks = 0; ks < 4; ++ks)
#pragma unroll
                for(int nt = 0; nt < NT; ++nt)
                {
                    const uint4 bb = bf[b][nt][ks >> 1];
                    if(ks & 1) mma(asum[nt], ones, bb.z, bb.w);
                    else mma(asum[nt], ones, bb.x, bb.y);
                }
        const float zf = 128.f + float(zc);
#pragma unroll
        for(int gi = 0; gi < GW; ++gi)
        {
            const uint32_t sl = lds_u16(s_q + gi * 32), sh = lds_u16(s_q + gi * 32 + 16);
            const float s_lo = __uint_as_float(sl << 16), s_hi = __uint_as_float(sh << 16);
            float part[NT][4];
#pragma unroll
            for(int nt = 0; nt < NT; ++nt) part[nt][0] = part[nt][1] = part[nt][2] = part[nt][3] = 0.f;
#pragma unroll
            for(int b = 0; b < 2; ++b)
            {
                const uint4 wv       = lds128(w_q + gi * w_gstride + b * 512);
                const uint32_t w4[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                for(int ks = 0; ks < 4; ++ks)
                {
                    uint32_t a[4];
#pragma unroll
                    for(int i = 0; i < 4; ++i) a[i] = lop3<0xEA>(w4[ks] >> (4 * i), 0x000F000Fu, 0x43004300u);
#pragma unroll
                    for(int nt = 0; nt < NT; ++nt)
                    {
                        const uint4 bb = bf[b][nt][ks >> 1];
                        if(ks & 1) mma(part[nt], a, bb.z, bb.w);
                        else mma(part[nt], a, bb.x, bb.y);
                    }
                }
            }
#pragma unroll
            for(int nt = 0; nt < NT; ++nt)
            {
                acc[gi][nt][0] = fmaf(s_lo, fmaf(-zf, asum[nt][0], part[nt][0]), acc[gi][nt][0]);
                acc[gi][nt][1] = fmaf(s_lo, fmaf(-zf, asum[nt][1], part[nt][1]), acc[gi][nt][1]);
                acc[gi][nt][2] = fmaf(s_hi, fmaf(-zf, asum[nt][2], part[nt][2]), acc[gi][nt][2]);
                acc[gi][nt][3] = fmaf(s_hi, fmaf(-zf, asum[nt][3], part[nt][3]), acc[gi][nt][3]);
            }
        }
    }
    else if(VAR <= 1)
    {
#pragma unroll
        for(int gi = 0; gi < GW; ++gi)
        {
            if(VAR == 0 && gi == GW - 1 && gi >= G) break;
            const uint32_t sl = lds_u16(s_q + gi * 32), sh = lds_u16(s_q + gi * 32 + 16);
            const float s_lo = __uint_as_float(sl << 16), s_hi = __uint_as_float(sh << 16);
            float part[NT][4];
#pragma unroll
            for(int nt = 0; nt < NT; ++nt) part[nt][0] = part[nt][1] = part[nt][2] = part[nt][3] = 0.f;
#pragma unroll
            for(int b = 0; b < 2; ++b)
            {
                const uint4 wv       = lds128(w_q + gi * w_gstride + b * 512);
                const uint32_t w4[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                for(int ks = 0; ks < 4; ++ks)
                {
                    uint32_t a[4];
                    unpack(w4[ks], zm, zm, a);
#pragma unroll
                    for(int nt = 0; nt < NT; ++nt)
                    {
                        const uint4 bb = bf[b][nt][ks >> 1];
                        if(ks & 1) mma(part[nt], a, bb.z, bb.w);
                        else mma(part[nt], a, bb.x, bb.y);
                    }
                }
            }
#pragma unroll
            for(int nt = 0; nt < NT; ++nt)
            {
                acc[gi][nt][0] = fmaf(s_lo, part[nt][0], acc[gi][nt][0]); acc[gi][nt][1] = fmaf(s_lo, part[nt][1], acc[gi][nt][1]);
                acc[gi][nt][2] = fmaf(s_hi, part[nt][2], acc[gi][nt][2]); acc[gi][nt][3] = fmaf(s_hi, part[nt][3], acc[gi][nt][3]);
            }
        }
    }
    else
    {
        constexpr int NCH = VAR == 3 ? 2 : 1;
        float part[GW][NCH][NT][4];
#pragma unroll
        for(int gi = 0; gi < GW; ++gi)
#pragma unroll
            for(int c = 0; c < NCH; ++c)
#pragma unroll
                for(int nt = 0; nt < NT; ++nt) part[gi][c][nt][0] = part[gi][c][nt][1] = part[gi][c][nt][2] = part[gi][c][nt][3] = 0.f;
        uint4 wv[GW][2];
#pragma unroll
        for(int gi = 0; gi < GW; ++gi)
#pragma unroll
            for(int b = 0; b < 2; ++b) wv[gi][b] = VAR == 4 ? wpre[gi][b] : lds128(w_q + gi * w_gstride + b * 512);
        if(VAR == 4)
        {
#pragma unroll
            for(int gi = 0; gi < GW; ++gi)
#pragma unroll
                for(int b = 0; b < 2; ++b) wpre[gi][b] = lds128(w_next + gi * w_gstride + b * 512);
        }
#pragma unroll
        for(int b = 0; b < 2; ++b)
#pragma unroll
            for(int ks = 0; ks < 4; ++ks)
#pragma unroll
                for(int gi = 0; gi < GW; ++gi)
                {
                    const uint32_t w = ks == 0 ? wv[gi][b].x : (ks == 1 ? wv[gi][b].y : (ks == 2 ? wv[gi][b].z : wv[gi][b].w));
                    uint32_t a[4];
                    unpack(w, zm, zm, a);
#pragma unroll
                    for(int nt = 0; nt < NT; ++nt)
                    {
                        const uint4 bb = bf[b][nt][ks >> 1];
                        float* pp      = part[gi][NCH == 2 ? (ks & 1) : 0][nt];
                        if(ks & 1) mma(pp, a, bb.z, bb.w);
                        else mma(pp, a, bb.x, bb.y);
                    }
                }
#pragma unroll
        for(int gi = 0; gi < GW; ++gi)
        {
            const uint32_t sl = lds_u16(s_q + gi * 32), sh = lds_u16(s_q + gi * 32 + 16);
            const float s_lo = __uint_as_float(sl << 16), s_hi = __uint_as_float(sh << 16);
#pragma unroll
            for(int nt = 0; nt < NT; ++nt)
            {
                float p0 = part[gi][0][nt][0], p1 = part[gi][0][nt][1], p2 = part[gi][0][nt][2], p3 = part[gi][0][nt][3];
                if(NCH == 2) { p0 += part[gi][NCH - 1][nt][0]; p1 += part[gi][NCH - 1][nt][1]; p2 += part[gi][NCH - 1][nt][2]; p3 += part[gi][NCH - 1][nt][3]; }
                acc[gi][nt][0] = fmaf(s_lo, p0, acc[gi][nt][0]); acc[gi][nt][1] = fmaf(s_lo, p1, acc[gi][nt][1]);
                acc[gi][nt][2] = fmaf(s_hi, p2, acc[gi][nt][2]); acc[gi][nt][3] = fmaf(s_hi, p3, acc[gi][nt][3]);
            }
        }
    }
}

// per warp: `ring` slots of ksg scale groups; the slot loop is a runtime loop as in the kernel (UQ = 1: the q loop is unrolled x2)
template <int NT, int GW, int VAR, int UQ, bool CLAMP>
__global__ void __launch_bounds__(256, NT == 1 ? 4 : 2) k_loop(int iters, int ksg, int G, int RB, int ring, long long* out, float* sink, uint32_t zc)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint32_t* sm32      = reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int a_bytes = 2 * ksg * RB * 128, w_bytes = GW * 2 * ksg * 512, s_bytes = (ksg * 32 * GW + 127) & ~127;
    const int total = 8 * ring * (a_bytes + w_bytes + s_bytes);
    for(int i = threadIdx.x; i < total / 4; i += blockDim.x) sm32[i] = (i * 2654435761u) & 0x3f7f3f7fu;
    __syncthreads();
    const uint32_t ring_a = base + warp * ring * a_bytes;
    const uint32_t ring_w = base + 8 * ring * a_bytes + warp * ring * w_bytes;
    const uint32_t ring_s = base + 8 * ring * (a_bytes + w_bytes) + warp * ring * s_bytes;
    const int g = lane >> 2, t = lane & 3;
    uint32_t a_addr[2][NT][2];
    bool a_ok[NT];
#pragma unroll
    for(int nt = 0; nt < NT; ++nt)
    {
        // CLAMP: lanes whose token row does not exist read the last row (a broadcast; their output columns are never stored)
        const int r = CLAMP ? min(8 * nt + g, RB - 1) : 8 * nt + g;
        a_ok[nt]    = CLAMP ? true : r < RB;
#pragma unroll
        for(int b = 0; b < 2; ++b)
        {
            const uint32_t rix = b * RB + r, sw = RB >= 4 ? (rix & 7) : 0;
            a_addr[b][nt][0] = rix * 128 + (((2 * t) ^ sw) << 4);
            a_addr[b][nt][1] = rix * 128 + (((2 * t + 1) ^ sw) << 4);
        }
    }
    float acc[GW][NT][4];
#pragma unroll
    for(int gi = 0; gi < GW; ++gi)
#pragma unroll
        for(int nt = 0; nt < NT; ++nt) acc[gi][nt][0] = acc[gi][nt][1] = acc[gi][nt][2] = acc[gi][nt][3] = 0.f;
    uint4 wpre[GW][2];
#pragma unroll
    for(int gi = 0; gi < GW; ++gi)
#pragma unroll
        for(int b = 0; b < 2; ++b) wpre[gi][b] = lds128(ring_w + lane * 16 + gi * (ksg * 1024) + b * 512);
    int rs = 0;
    long long t0 = clock64();
    for(int it = 0; it < iters; ++it)
    {
        const uint32_t a_s = ring_a + rs * a_bytes, w_s = ring_w + rs * w_bytes + lane * 16, s_s = ring_s + rs * s_bytes + g * 2;
        const int rs_n     = rs + 1 == ring ? 0 : rs + 1;
        const uint32_t w_n = ring_w + rs_n * w_bytes + lane * 16;
        if(UQ)
        {
#pragma unroll 2
            for(int q = 0; q < ksg; ++q)
                scale_group<NT, GW, VAR>(a_s + q * 2 * RB * 128, w_s + q * 1024, s_s + q * 32 * G, ksg * 1024, G, a_addr, a_ok, zc, acc, wpre,
                                         q + 1 < ksg ? w_s + (q + 1) * 1024 : w_n);
        }
        else
        {
#pragma unroll 1
            for(int q = 0; q < ksg; ++q)
                scale_group<NT, GW, VAR>(a_s + q * 2 * RB * 128, w_s + q * 1024, s_s + q * 32 * G, ksg * 1024, G, a_addr, a_ok, zc, acc, wpre,
                                         q + 1 < ksg ? w_s + (q + 1) * 1024 : w_n);
        }
        rs = rs_n;
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for(int gi = 0; gi < GW; ++gi)
#pragma unroll
        for(int nt = 0; nt < NT; ++nt) s += acc[gi][nt][0] + acc[gi][nt][1] + acc[gi][nt][2] + acc[gi][nt][3];
    if(s == 12345.f) sink[0] = s + wpre[0][0].x;
    if(threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

template <int NT, int GW, int VAR, int UQ, bool CLAMP = true>
void run(long long* out, float* sink, int ksg, int G, int RB, int ctas_per_sm)
{
    const int iters = 300, ring = 2;
    const int smem = 8 * ring * (2 * ksg * RB * 128 + GW * 2 * ksg * 512 + ((ksg * 32 * GW + 127) & ~127)) + 1024;
    cudaFuncSetAttribute(k_loop<NT, GW, VAR, UQ, CLAMP>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(k_loop<NT, GW, VAR, UQ, CLAMP>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    k_loop<NT, GW, VAR, UQ, CLAMP><<<148 * ctas_per_sm, 256, smem>>>(iters, ksg, G, RB, ring, out, sink, 8);
    cudaError_t e = cudaDeviceSynchronize();
    if(e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return; }
    // per SMSP: 2 warps per CTA x ksg x G(real) groups x 8 words per iteration
    const double words = double(iters) * 2.0 * ctas_per_sm * ksg * (VAR == 0 ? G : GW) * 8.0;
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, k_loop<NT, GW, VAR, UQ, CLAMP>);
    printf("NT=%d GW=%d G=%d var=%d clamp=%d ksg=%d RB=%d ctas/SM=%d regs=%d: %.1f clk per warp-word per SMSP\n", NT, GW, G, VAR, int(CLAMP), ksg, RB,
           ctas_per_sm, fa.numRegs, double(out[0]) / words);
}

int main()
{
    long long* out; cudaMallocManaged(&out, 4096 * 8);
    float* sink; cudaMalloc(&sink, 64);
    // 16 warps per SM (two co-resident 8-warp CTAs) unless noted
    run<1, 1, 1, 0, false>(out, sink, 4, 1, 1, 2);
    run<1, 1, 1, 0>(out, sink, 4, 1, 1, 2);
    run<1, 1, 3, 0>(out, sink, 4, 1, 1, 2);
    run<1, 1, 5, 0>(out, sink, 4, 1, 1, 2);
    run<1, 2, 1, 0>(out, sink, 2, 2, 1, 2);
    run<1, 2, 2, 0>(out, sink, 2, 2, 1, 2);
    run<1, 2, 5, 0>(out, sink, 2, 2, 1, 2);
    run<1, 3, 1, 0>(out, sink, 1, 3, 1, 2);
    run<1, 3, 5, 0>(out, sink, 1, 3, 1, 2);
    run<1, 4, 1, 0>(out, sink, 1, 4, 1, 2);
    run<1, 4, 5, 0>(out, sink, 1, 4, 1, 2);
    run<1, 2, 1, 0>(out, sink, 2, 2, 4, 2);
    run<1, 2, 5, 0>(out, sink, 2, 2, 4, 2);
    run<1, 2, 1, 0>(out, sink, 2, 2, 8, 2);
    run<1, 1, 1, 0>(out, sink, 4, 1, 1, 1);
    run<1, 2, 1, 0>(out, sink, 2, 2, 1, 1);
    run<1, 2, 5, 0>(out, sink, 2, 2, 1, 1);
    run<1, 4, 1, 0>(out, sink, 1, 4, 1, 1);
    run<2, 2, 1, 0>(out, sink, 1, 2, 16, 1);
    run<2, 4, 1, 0>(out, sink, 1, 4, 16, 1);
    run<2, 2, 1, 0>(out, sink, 1, 2, 16, 2);
    run<2, 2, 5, 0>(out, sink, 1, 2, 16, 2);
    return 0;
}
