// ptx.cuh -- thin inline-PTX wrappers for sm_100a: mbarrier, TMA (cp.async.bulk[.tensor]),
// tcgen05 (alloc / mma / commit / ld / st / fences), cache policies, PDL.
// Everything here is single-instruction glue; the kernels own all control flow.
#pragma once
#include <cstdint>
#include <cuda.h>

namespace ckb {
namespace ptx {

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint32_t lane_id()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%laneid;" : "=r"(r));
    return r;
}

__device__ __forceinline__ bool elect_one()
{
    uint32_t pred = 0;
    asm volatile(
        "{\n\t"
        ".reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ uint64_t global_timer_ns()
{
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}

// non-blocking probe (try_wait may suspend the thread for a hardware time limit when the phase is pending)
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}

// Bounded wait: a protocol bug becomes a trap (launch error) instead of a hung GPU.
#ifndef CKB_WAIT_TIMEOUT_NS
#define CKB_WAIT_TIMEOUT_NS 4000000000ull
#endif
// slow path kept out of line: the printf/trap sequence is large and would otherwise be replicated at
// every wait site (instruction-cache footprint is what limits the warp-specialised kernels)
static __device__ __noinline__ void mbar_wait_slow(uint32_t bar, uint32_t parity)
{
    const uint64_t t0 = global_timer_ns();
    uint32_t spins    = 0;
    while(!mbar_try_wait(bar, parity))
    {
        if((++spins & 0x3ff) == 0 && global_timer_ns() - t0 > CKB_WAIT_TIMEOUT_NS)
        {
            printf("ckb: mbarrier wait timeout (block %d thread %d bar 0x%x parity %u)\n",
                   int(blockIdx.x), int(threadIdx.x), bar, parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    if(mbar_try_wait(bar, parity)) return;
    mbar_wait_slow(bar, parity);
}

// Probe several barriers in ONE asm block so the (~90-cycle) try_wait latencies overlap instead of
// chaining (each selp would otherwise stall in-order issue behind its own try_wait).
__device__ __forceinline__ uint32_t mbar_try_wait2(uint32_t b0, uint32_t p0, uint32_t b1, uint32_t p1)
{
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P0, P1;\n\t"
        ".reg .b32 r0, r1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P0, [%1], %2;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%3], %4;\n\t"
        "selp.u32 r0, 1, 0, P0;\n\t"
        "selp.u32 r1, 2, 0, P1;\n\t"
        "or.b32 %0, r0, r1;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(b0), "r"(p0), "r"(b1), "r"(p1)
        : "memory");
    return ok;
}
__device__ __forceinline__ uint32_t
mbar_try_wait3(uint32_t b0, uint32_t p0, uint32_t b1, uint32_t p1, uint32_t b2, uint32_t p2)
{
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P0, P1, P2;\n\t"
        ".reg .b32 r0, r1, r2;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P0, [%1], %2;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%3], %4;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P2, [%5], %6;\n\t"
        "selp.u32 r0, 1, 0, P0;\n\t"
        "selp.u32 r1, 2, 0, P1;\n\t"
        "selp.u32 r2, 4, 0, P2;\n\t"
        "or.b32 r0, r0, r1;\n\t"
        "or.b32 %0, r0, r2;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(b0), "r"(p0), "r"(b1), "r"(p1), "r"(b2), "r"(p2)
        : "memory");
    return ok;
}
// six NON-BLOCKING probes in one block (test_wait: try_wait would suspend the thread on a pending
// barrier for a hardware time limit); bit k set = barrier k already complete
__device__ __forceinline__ uint32_t mbar_test_wait6(const uint32_t* b, const uint32_t* p)
{
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P0, P1, P2, P3, P4, P5;\n\t"
        ".reg .b32 r0, r1, r2, r3, r4, r5;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P0, [%1], %2;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P1, [%3], %4;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P2, [%5], %6;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P3, [%7], %8;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P4, [%9], %10;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P5, [%11], %12;\n\t"
        "selp.u32 r0, 1, 0, P0;\n\t"
        "selp.u32 r1, 2, 0, P1;\n\t"
        "selp.u32 r2, 4, 0, P2;\n\t"
        "selp.u32 r3, 8, 0, P3;\n\t"
        "selp.u32 r4, 16, 0, P4;\n\t"
        "selp.u32 r5, 32, 0, P5;\n\t"
        "or.b32 r0, r0, r1;\n\t"
        "or.b32 r2, r2, r3;\n\t"
        "or.b32 r4, r4, r5;\n\t"
        "or.b32 r0, r0, r2;\n\t"
        "or.b32 %0, r0, r4;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(b[0]), "r"(p[0]), "r"(b[1]), "r"(p[1]), "r"(b[2]), "r"(p[2]), "r"(b[3]), "r"(p[3]), "r"(b[4]),
          "r"(p[4]), "r"(b[5]), "r"(p[5])
        : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_wait2(uint32_t b0, uint32_t p0, uint32_t b1, uint32_t p1)
{
    const uint32_t ok = mbar_try_wait2(b0, p0, b1, p1);
    if(!(ok & 1)) mbar_wait(b0, p0);
    if(!(ok & 2)) mbar_wait(b1, p1);
}
__device__ __forceinline__ void
mbar_wait3(uint32_t b0, uint32_t p0, uint32_t b1, uint32_t p1, uint32_t b2, uint32_t p2)
{
    const uint32_t ok = mbar_try_wait3(b0, p0, b1, p1, b2, p2);
    if(!(ok & 1)) mbar_wait(b0, p0);
    if(!(ok & 2)) mbar_wait(b1, p1);
    if(!(ok & 4)) mbar_wait(b2, p2);
}

// ------------------------------------------------------------------ proxies / fences
__device__ __forceinline__ void fence_proxy_async_smem()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_before_sync()
{
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after_sync()
{
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// barrier id is an immediate: a register id makes ptxas reserve all 16 hardware barriers for the CTA,
// which silently limits residency to one CTA per SM
template <int ID>
__device__ __forceinline__ void named_bar_sync(uint32_t nthreads)
{
    asm volatile("bar.sync %0, %1;" ::"n"(ID), "r"(nthreads) : "memory");
}

// ------------------------------------------------------------------ system-scope signalling (peer GPUs over NVLink)
__device__ __forceinline__ void red_release_sys_add_u32(unsigned* p, unsigned v)
{
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned atom_acq_rel_gpu_add_u32(unsigned* p, unsigned v)
{
    unsigned old;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ unsigned ld_relaxed_sys_u32(const unsigned* p)
{
    unsigned v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p)
{
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// ------------------------------------------------------------------ L2 cache policies
__device__ __forceinline__ uint64_t policy_evict_first()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_last()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}

// ------------------------------------------------------------------ TMA
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m)
{
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tiled load global -> smem, completion on an mbarrier (complete_tx::bytes)
__device__ __forceinline__ void
tma_load_2d(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int32_t x, int32_t y)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4}], [%2];" ::"r"(smem_dst),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(x), "r"(y)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(
    uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int32_t x, int32_t y, uint64_t policy)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
                 ".L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_dst),
                 "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(x), "r"(y), "l"(policy)
                 : "memory");
}
// 3-D tiled load global -> smem with an L2 cache hint
__device__ __forceinline__ void tma_load_3d_hint(
    uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int32_t x, int32_t y, int32_t z, uint64_t policy)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
                 ".L2::cache_hint [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(smem_dst),
                 "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(x), "r"(y), "r"(z), "l"(policy)
                 : "memory");
}
__device__ __forceinline__ void
tma_load_3d(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int32_t x, int32_t y, int32_t z)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_dst),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(x), "r"(y), "r"(z)
        : "memory");
}
// 2-D tiled store smem -> global (bulk_group completion)
__device__ __forceinline__ void
tma_store_2d(const CUtensorMap* m, uint32_t smem_src, int32_t x, int32_t y)
{
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];" ::
                     "l"(reinterpret_cast<uint64_t>(m)),
                 "r"(smem_src), "r"(x), "r"(y)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit()
{
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait()
{
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// 1-D bulk copy global -> smem (bytes % 16 == 0, 16-byte aligned both sides)
__device__ __forceinline__ void
bulk_load_1d(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t bar, uint64_t policy)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
                 "[%0], [%1], %2, [%3], %4;" ::"r"(smem_dst),
                 "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(bar), "l"(policy)
                 : "memory");
}

// ------------------------------------------------------------------ warp-level MMA (HMMA) + shared loads
// D[16x8] += A[16x16] * B[16x8], 16-bit operands, fp32 accumulate
template <bool F16>
__device__ __forceinline__ void mma_m16n8k16(float* d, const uint32_t* a, uint32_t b0, uint32_t b1)
{
    if(F16)
        asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    else
        asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v)
{
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
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
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

// ------------------------------------------------------------------ TMEM
// whole-warp, .sync.aligned
template <int NCOLS>
__device__ __forceinline__ void tmem_alloc(uint32_t smem_result)
{
    static_assert(NCOLS >= 32 && NCOLS <= 512 && (NCOLS & (NCOLS - 1)) == 0, "power of two 32..512");
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_result),
                 "n"(NCOLS)
                 : "memory");
}
__device__ __forceinline__ void tmem_relinquish()
{
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS)
                 : "memory");
}

// ------------------------------------------------------------------ tcgen05.mma
// Shared-memory matrix descriptor, K-major operand, 128-byte swizzle (what TMA SWIZZLE_128B writes):
//   rows are 128 B apart, 8-row groups (one swizzle atom, 1024 B) are SBO apart.
//   bits [0,14) addr>>4 | [16,30) LBO>>4 (unused for swizzled K-major) | [32,46) SBO>>4 |
//   [46,48) version = 1 | [61,64) layout = 2 (SWIZZLE_128B)      (cute/arch/mma_sm100_desc.hpp)
__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t smem_addr, uint32_t sbo_bytes)
{
    uint64_t d = 0;
    d |= uint64_t((smem_addr & 0x3FFFF) >> 4);
    d |= uint64_t(1) << 16;
    d |= uint64_t((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= uint64_t(1) << 46;
    d |= uint64_t(2) << 61;
    return d;
}

// Instruction descriptor (32 bit) for kind::f16 / kind::i8, dense, K-major A and B.
//   [4,6) D format: 1 = F32, 2 = S32 | [7,10) A format | [10,13) B format
//   (kind::f16: 0 = F16, 1 = BF16; kind::i8: 0 = U8, 1 = S8) | [15] A major | [16] B major
//   [17,23) N >> 3 | [24,29) M >> 4
__host__ __device__ constexpr uint32_t
make_idesc(uint32_t d_fmt, uint32_t a_fmt, uint32_t b_fmt, uint32_t M, uint32_t N)
{
    return (d_fmt << 4) | (a_fmt << 7) | (b_fmt << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]
__device__ __forceinline__ void
mma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void
mma_i8_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]   (A: lane = row, 16-bit elements packed two per 32-bit column)
__device__ __forceinline__ void
mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// all previously issued tcgen05.mma of this thread arrive (count 1) on the mbarrier when complete;
// implies tcgen05.fence::before_thread_sync
__device__ __forceinline__ void mma_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
                 : "memory");
}


// ------------------------------------------------------------------ CTA pair (cta_group::2) variants
// Shared-window addresses of the two CTAs of a pair differ in bit 24; clearing it names the even (leader) CTA's
// copy of the same location (cute/arch/copy_sm100_tma.hpp: Sm100MmaPeerBitMask).
__device__ __forceinline__ uint32_t leader_addr(uint32_t smem_addr) { return smem_addr & 0xFEFFFFFFu; }
__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on a barrier of any CTA of the cluster (shared::cluster address)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster_addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
// Same arrive with the default semantics (.release at CTA scope), as CUTLASS's ClusterBarrier::arrive does for a remote
// barrier: .release.cluster compiles to MEMBAR.ALL.GPU + ERRBAR in front of the SYNCS.ARRIVE (ncu: the membar stall was the
// longest link of the transform chain of the fused prefill kernel).  Correct where the data the barrier guards was made
// visible by other means -- fence.proxy.async for operand stages read by the tensor core, tcgen05 fences for TMEM.
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar_cluster_addr)
{
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
// tile load into THIS CTA's shared memory, completion bytes credited to the LEADER CTA's barrier
__device__ __forceinline__ void tma_load_2d_pair(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int32_t x, int32_t y)
{
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
                 "[%0], [%1, {%3, %4}], [%2];" ::"r"(smem_dst),
                 "l"(reinterpret_cast<uint64_t>(m)), "r"(leader_addr(bar)), "r"(x), "r"(y)
                 : "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t smem_result)
{
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_result), "n"(NCOLS) : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair()
{
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr)
{
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS) : "memory");
}
__device__ __forceinline__ void
mma_f16_ss_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
                 "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
                 : "memory");
}
__device__ __forceinline__ void
mma_i8_ss_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
                 "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
                 : "memory");
}
// completion of all prior MMAs of the pair -> the barrier at the same offset in BOTH CTAs
__device__ __forceinline__ void mma_commit_pair(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((uint16_t)3)
                 : "memory");
}

// ------------------------------------------------------------------ tcgen05.ld / st (32 lanes x 32 bit, xN columns)
// warp w may only touch TMEM lanes [32*(w%4), 32*(w%4)+32); taddr = (lane << 16) | column
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void tmem_ld_x8(uint32_t taddr, uint32_t* r)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
                   "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t* r)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
                   "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]),
                   "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_x32(uint32_t taddr, uint32_t* r)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
                   "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]),
                   "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
                   "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                   "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]),
                   "=r"(r[31])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_st_x8(uint32_t taddr, const uint32_t* r)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr),
                 "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_st_x16(uint32_t taddr, const uint32_t* r)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
                 "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
                 "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                 "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]),
                 "r"(r[15])
                 : "memory");
}

// ------------------------------------------------------------------ programmatic dependent launch
__device__ __forceinline__ void pdl_wait_prior_grids() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ------------------------------------------------------------------ global memory helpers
__device__ __forceinline__ void red_add_f32(float* p, float v)
{
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add_v4_f32(float* p, float a, float b, float c, float d)
{
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

} // namespace ptx
} // namespace ckb
