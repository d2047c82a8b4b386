// api.cu -- the C ABI of libckb.so (include/ckb.h): argument validation (CK's IsSupportedArgument),
// kernel selection (CK's instance choice), workspace sizing, launch.  No CPU fallback exists here.
#include "common.cuh"
#include <memory>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "kernels.h"
#include "tile_map.cuh"

namespace ckb {
namespace {

bool is_f16ish(int dt) { return dt == CKB_F16 || dt == CKB_BF16; }

int unsupported(ckb_gemm_plan* plan, const char* fmt, ...)
{
    char buf[256];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error()    = buf;
    plan->supported = 0;
    plan->kernel_id = KID_NONE;
    return CKB_OK;
}

uint64_t d_bytes(const ckb_gemm_problem& p, int i)
{
    const uint64_t e = dtype_bytes(p.d_dtype[i]);
    const uint64_t m = p.d_sm[i] ? p.M : 1, n = p.d_sn[i] ? p.N : 1;
    return e * m * n;
}

uint64_t algo_bytes(const ckb_gemm_problem& p)
{
    uint64_t b = uint64_t(p.M) * p.K * dtype_bytes(p.a_dtype);
    if(p.b_dtype == CKB_I4) b += uint64_t(p.N) * p.K / 2;
    else b += uint64_t(p.N) * p.K * dtype_bytes(p.b_dtype);
    if((p.b_op == CKB_BOP_GROUP || p.b_op == CKB_BOP_GROUP_RTN) && p.group_size > 0)
    {
        const uint64_t g = (p.K + p.group_size - 1) / p.group_size;
        b += g * p.N * dtype_bytes(p.s_dtype);
        if(p.zero) b += g * p.N;
    }
    else if(p.b_op == CKB_BOP_MULTIPLY || p.b_op == CKB_BOP_FPAINTB)
        b += uint64_t(p.N) * dtype_bytes(p.s_dtype);
    for(int i = 0; i < p.num_d; ++i) b += d_bytes(p, i);
    b += uint64_t(p.M) * p.N * dtype_bytes(p.e_dtype);
    return b;
}

int needed_d(int cde)
{
    static const int need[] = {0, 1, 0, 1, 1, 2, 1, 2, 2, 0, 1, 1, 2};
    return (cde >= 0 && cde <= CKB_CDE_ADD_ACT_MUL2_CLAMP) ? need[cde] : 99;
}

} // namespace
} // namespace ckb

namespace ckb {
namespace {
// Small host->device upload done by SMs reading MAPPED pinned host memory: for the 8-350 KB activation matrices of a
// decode step a copy-engine transfer costs ~8 us of latency per direction inside the stream, a kernel ~3 us.
__global__ void __launch_bounds__(256) upload_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, long long n16,
                                                      const unsigned char* src_tail, unsigned char* dst_tail, int tail)
{
    // Programmatic dependent launch (used while a stream is being captured): let the GEMM that follows become resident right away
    // -- it fetches its weights while this copy and the previous GEMM of the stream are still running -- then wait for the
    // previous GEMM, which may still be reading the staging buffer this copy overwrites.  No-ops in an ordinary launch.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    for(long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (long long)gridDim.x * blockDim.x) dst[i] = src[i];
    if(blockIdx.x == 0 && threadIdx.x < tail) dst_tail[threadIdx.x] = src_tail[threadIdx.x];
}

// device-visible alias of a pinned (mapped) host pointer, or nullptr when `p` is pageable / not mapped
void* mapped_alias(const void* p)
{
    cudaPointerAttributes at{};
    if(cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if(at.type != cudaMemoryTypeHost || !at.devicePointer) return nullptr;
    return at.devicePointer;
}
} // namespace
} // namespace ckb

using namespace ckb;

extern "C" {

const char* ckb_last_error(void) { return last_error().c_str(); }
int ckb_version(void) { return CKB_VERSION; }
int ckb_device_sm_count(void) { return device_info().ok ? device_info().sm_count : -1; }

int ckb_tile_coord(int32_t t, int32_t num_m_blk, int32_t num_n_blk, int32_t group, int32_t* mb, int32_t* nb)
{
    if(!mb || !nb || num_m_blk <= 0 || num_n_blk <= 0 || group <= 0 || t < 0 || t >= num_m_blk * num_n_blk)
        return set_error(CKB_ERR_INVALID, "ckb_tile_coord: bad argument");
    int m = 0, n = 0;
    tile_coord(t, num_m_blk, num_n_blk, m, n, group);
    *mb = m; *nb = n;
    return CKB_OK;
}

int ckb_streamk_range(int64_t total_units, int32_t cta, int32_t grid, int64_t* begin, int64_t* end)
{
    if(!begin || !end || grid <= 0 || cta < 0 || cta >= grid || total_units < 0) return set_error(CKB_ERR_INVALID, "ckb_streamk_range: bad argument");
    long long b = 0, e = 0;
    streamk_range(total_units, cta, grid, b, e);
    *begin = b; *end = e;
    return CKB_OK;
}

int ckb_gemv_schedule(int32_t M, int32_t N, int32_t K, int32_t zero_points, int32_t sms, int32_t mode, int32_t* out)
{
    if(!out || N <= 0 || K <= 0 || sms <= 0) return set_error(CKB_ERR_INVALID, "ckb_gemv_schedule: bad argument");
    G2Sched s{};
    if(int rc = schedule_wq_gemv2(M, N, K, zero_points != 0, sms, s, mode)) return rc;
    const int32_t v[24] = {s.nt, s.rb, s.n_groups, s.kb_total, s.grid, s.passes, s.n_slots, s.slot_q, s.slot_r, s.g_hi, s.wn_log2, s.gw,
                           s.kbu, s.upw, s.n_chunks, s.ring, s.stage_bytes, s.off_w, s.off_s, s.off_z, s.red_off, s.bar_off, s.smem_bytes,
                           s.nt == 1 ? 16 : 8};
    for(int i = 0; i < 24; ++i) out[i] = v[i];
    return CKB_OK;
}

int ckb_peer_wait(void* flag, void* epoch_state, uint32_t arrivals_per_epoch, void* stream)
{
    if(!flag || !epoch_state || arrivals_per_epoch == 0) return set_error(CKB_ERR_INVALID, "ckb_peer_wait: null counter or zero arrivals");
    if(!device_info().ok) return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device");
    return launch_peer_wait(flag, epoch_state, arrivals_per_epoch, static_cast<cudaStream_t>(stream));
}

int ckb_gemm_query(const ckb_gemm_problem* pp, ckb_gemm_plan* plan)
{
    if(!pp || !plan) return set_error(CKB_ERR_INVALID, "null problem/plan");
    const ckb_gemm_problem& p = *pp;
    memset(plan, 0, sizeof *plan);
    if(p.M < 0 || p.N < 0 || p.K < 0) return set_error(CKB_ERR_INVALID, "negative GEMM size");
    plan->flops      = 2ull * p.M * p.N * p.K;
    plan->algo_bytes = algo_bytes(p);
    if(p.K <= 0) return unsupported(plan, "K must be positive");
    if(p.num_d < needed_d(p.cde_op) || p.num_d > CKB_MAX_D)
        return unsupported(plan, "cde_op %d needs %d D tensors, got %d", p.cde_op, needed_d(p.cde_op), p.num_d);
    if(p.a_layout != CKB_ROW) return unsupported(plan, "A must be row-major (mk_*)");
    for(int i = 0; i < needed_d(p.cde_op); ++i)
        if(!p.d[i]) return unsupported(plan, "D%d pointer is null", i);
    if(p.M == 0 || p.N == 0)
    {
        plan->supported = 1; // nothing to compute (empty problem): run() is a no-op
        plan->kernel_id = KID_NONE;
        snprintf(plan->kernel_name, sizeof plan->kernel_name, "empty");
        return CKB_OK;
    }
    if(!p.a || !p.b || !p.e) return unsupported(plan, "A/B/E pointer is null");

    if(p.a_dtype == CKB_I8 || (is_f16ish(p.a_dtype) && p.b_dtype == p.a_dtype))
    {
        // ---- K3: SS GEMM ----
        if(p.b_dtype != p.a_dtype) return unsupported(plan, "int8 A needs int8 B");
        if(p.b_layout != CKB_COL) return unsupported(plan, "SS kernel needs K-major (column-major) B");
        if(p.b_op != CKB_BOP_PASSTHROUGH) return unsupported(plan, "SS kernel supports PassThrough B only");
        const bool i8 = p.a_dtype == CKB_I8;
        if(i8)
        {
            switch(p.cde_op)
            {
            case CKB_CDE_PASSTHROUGH: case CKB_CDE_ACT_MUL_CLAMP: case CKB_CDE_ADD_ACT_MUL_CLAMP:
            case CKB_CDE_ACT_MUL2_CLAMP: case CKB_CDE_ADD_ACT_MUL2_CLAMP: case CKB_CDE_MULTIPLY_MULTIPLY: break;
            default: return unsupported(plan, "cde_op %d is not defined on int32 accumulators", p.cde_op);
            }
            if((p.cde_op == CKB_CDE_ADD_ACT_MUL_CLAMP || p.cde_op == CKB_CDE_ADD_ACT_MUL2_CLAMP) && p.d_dtype[0] != CKB_I32)
                return unsupported(plan, "bias must be int32");
        }
        else if(p.cde_op > CKB_CDE_MULTIPLY_MULTIPLY)
            return unsupported(plan, "quantization epilogues need int8 inputs");
        const int eb = dtype_bytes(p.a_dtype);
        if((p.lda * eb) % 16 || (p.ldb * eb) % 16) return unsupported(plan, "lda/ldb must be 16-byte multiples (K %% %d)", 16 / eb);
        if(p.lda < p.K || p.ldb < p.K || p.lde < p.N) return unsupported(plan, "leading dimension smaller than extent");
        if(!device_info().ok) { plan->supported = 0; return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device"); }
        if(int rc = plan_gemm_ss(p, *plan)) { if(rc == CKB_ERR_UNSUPPORTED) { plan->supported = 0; return CKB_OK; } return rc; }
        plan->kernel_id = KID_GEMM_SS;
        plan->supported = 1;
        return CKB_OK;
    }

    if(!is_f16ish(p.a_dtype)) return unsupported(plan, "A must be fp16, bf16 or int8");
    // ---- K1/K2: weight-quantized GEMM ----
    if(p.cde_op > CKB_CDE_MULTIPLY_MULTIPLY) return unsupported(plan, "quantization epilogues need int8 inputs");
    if(p.e_dtype != CKB_F16 && p.e_dtype != CKB_BF16 && p.e_dtype != CKB_F32) return unsupported(plan, "E must be f16/bf16/f32");
    switch(p.b_op)
    {
    case CKB_BOP_GROUP:
    case CKB_BOP_GROUP_RTN:
        if(p.b_dtype != CKB_I4 || p.b_layout != CKB_PREPACKED) return unsupported(plan, "GROUP dequant needs pre-packed int4 B (ckb_prepack_b_int4)");
        if(p.group_size != 128) return unsupported(plan, "group_size must be 128");
        if(!p.scale || p.s_dtype != p.a_dtype) return unsupported(plan, "group scales must have A's type");
        if(p.N % 8 || (p.zero && p.N % 16)) return unsupported(plan, "N must be a multiple of 8 (16 with zero points)");
        if(p.zero_const < 0 || p.zero_const > 15) return unsupported(plan, "zero_const out of range");
        if(p.K % 64) return unsupported(plan, "pre-packed int4 B needs K to be a multiple of 64");
        break;
    case CKB_BOP_MULTIPLY:
        if(p.a_dtype != CKB_BF16 || p.b_dtype != CKB_I8 || p.s_dtype != CKB_BF16 || !p.scale)
            return unsupported(plan, "Multiply B-op: bf16 A, int8 B0, bf16 B1");
        break;
    case CKB_BOP_FPAINTB:
        if(p.a_dtype != CKB_F16 || p.b_dtype != CKB_U8 || p.s_dtype != CKB_F16 || !p.scale)
            return unsupported(plan, "fpAintB: fp16 A, biased uint8 B (ckb_prepack_b_u8), fp16 scale");
        break;
    case CKB_BOP_PASSTHROUGH:
        if(p.b_dtype != CKB_I8) return unsupported(plan, "PassThrough B-op needs int8 B");
        break;
    default: return unsupported(plan, "unknown b_op %d", p.b_op);
    }
    const bool w4 = p.b_op == CKB_BOP_GROUP || p.b_op == CKB_BOP_GROUP_RTN;
    if(p.num_peers < 0 || p.num_peers > CKB_MAX_PEERS) return set_error(CKB_ERR_INVALID, "num_peers out of range");
    if(p.num_peers > 0)
    {
        if(p.b_op != CKB_BOP_GROUP || p.M > 32) return unsupported(plan, "the fused all-gather epilogue is served by the W4A16 decode kernel (CKB_BOP_GROUP, M <= 32)");
        if(p.peer_rank < 0 || p.peer_rank >= p.num_peers) return set_error(CKB_ERR_INVALID, "peer_rank out of range");
        for(int g = 0; g < p.num_peers; ++g)
            if(!p.e_peer[g] || !p.flag_peer[g]) return set_error(CKB_ERR_INVALID, "null peer pointer");
    }
    if(!w4)
    {
        if(p.b_layout != CKB_COL) return unsupported(plan, "int8 B must be K-major (column-major: mk_nk_mn)");
        if(p.ldb % 16 || p.ldb < p.K) return unsupported(plan, "ldb must be a multiple of 16 and >= K");
    }
    if(p.lda % 8 || p.lda < p.K) return unsupported(plan, "lda must be a multiple of 8 and >= K");
    if(p.lde < p.N) return unsupported(plan, "lde < N");
    if(!device_info().ok) { plan->supported = 0; return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device"); }
    // The two W4A16 numerics contracts keep their definition at EVERY M (no switch at the decode / prefill boundary):
    // GROUP (fp32 group scale on partial sums) is always served by the HMMA decode kernels, in row slabs above 64 rows;
    // GROUP_RTN (one rounding into A's type) by the fused tcgen05 kernel above 64 rows and by the RTN instantiation of the
    // HMMA kernel below.
    if((p.flags & CKB_FLAG_PREFILL_TWO_PASS) && p.M <= 64) return unsupported(plan, "the two-pass prefill instance serves M > 64");
    if((p.flags & CKB_FLAG_W8_STREAMK) && (p.M > 64 || w4)) return unsupported(plan, "the stream-K decode instance serves int8 weights at M <= 64");
    if(p.M > 64 && p.K % 8 == 0 && p.b_op != CKB_BOP_GROUP)
    {
        // prefill-shaped: the fused TMA -> dequant -> tcgen05 pair kernel (no workspace).  CKB_PREFILL_TWO_PASS=1 selects the
        // round-1 form (expand the weight into the workspace, then the dense GEMM) for comparison.
        const bool two_pass = (p.flags & CKB_FLAG_PREFILL_TWO_PASS) || (getenv("CKB_PREFILL_TWO_PASS") && atoi(getenv("CKB_PREFILL_TWO_PASS")) == 1);
        if(int rc = two_pass ? plan_wq_prefill(p, *plan) : plan_wq_fused(p, *plan)) return rc;
        plan->kernel_id = two_pass ? KID_WQ_PREFILL : KID_WQ_FUSED;
        plan->supported = 1;
        return CKB_OK;
    }
    if(w4)
    {
        // W4A16, M <= 64 (larger M with K % 8 != 0: slabs): N-split HMMA kernels, no workspace.  Two schedules of the same
        // math (profiles/r2_gemv_variants.txt): the rotating-producer kernel (wq_gemv2, <= 32 rows per launch) wins where one
        // pass covers the CTA's channels and either M <= 8 or the whole K extent is resident in the ring; the single-producer
        // kernel (wq_gemv, <= 64 rows) wins on streaming shapes and on M > 8.  CKB_GEMV=1 / 2 forces one of them.
        const int force = getenv("CKB_GEMV") ? atoi(getenv("CKB_GEMV")) : ((p.flags & CKB_FLAG_GEMV_SINGLE_PRODUCER) ? 1 : 0);
        // RTN numerics and the fixed-K-order contract exist in the rotating-producer kernel only
        const bool g2_only = p.b_op == CKB_BOP_GROUP_RTN || (p.flags & CKB_FLAG_FIXED_K_ORDER) || p.num_peers > 0;
        // rows beyond 32 go through 32-row slabs: the single-producer kernel's 64-row variant (8 token tiles) is retired -- its
        // activation stage overflows the ring on K >= 4096 (mbarrier time-out found by tests/slab_probe.py)
        bool old_kernel = !g2_only && force == 1;
        if(!old_kernel && force != 2 && !g2_only)
        {
            G2Sched s{};
            const int sms = device_info().sm_count > 0 ? device_info().sm_count : 148;
            if(schedule_wq_gemv2(std::min(p.M, 32), p.N, p.K, p.zero != nullptr, sms, s) != CKB_OK) old_kernel = true;
            else old_kernel = !(s.passes == 1 && (s.nt == 1 || s.ring >= s.n_chunks));
        }
        ckb_gemm_problem q = p;
        q.M = std::min(p.M, 32);
        if(int rc = old_kernel ? plan_wq_gemv(q, *plan) : plan_wq_gemv2(q, *plan)) { if(rc == CKB_ERR_UNSUPPORTED) { plan->supported = 0; return CKB_OK; } return rc; }
        plan->kernel_id = old_kernel ? KID_WQ_GEMV : KID_WQ_GEMV2;
        plan->supported = 1;
        return CKB_OK;
    }
    // int8-weight formats at decode sizes: the N-split HMMA kernel of the W4 path (K-major int8 streamed as [kb][row][64 B] boxes,
    // 32 rows per launch, no workspace) when K is a whole number of 64-k blocks; the stream-K tcgen05 kernel otherwise and behind
    // CKB_FLAG_W8_STREAMK / CKB_W8_DECODE=streamk (second factory instance)
    {
        const char* e8 = getenv("CKB_W8_DECODE");
        const bool streamk = (p.flags & CKB_FLAG_W8_STREAMK) || (e8 && !strcmp(e8, "streamk"));
        if(!streamk && p.K % 64 == 0 && p.M >= 1)
        {
            ckb_gemm_problem q = p;
            q.M = std::min(p.M, 32);
            const int rc = plan_wq_gemv2(q, *plan);
            if(rc == CKB_OK) { plan->kernel_id = KID_WQ_GEMV2; plan->supported = 1; return CKB_OK; }
            if(rc != CKB_ERR_UNSUPPORTED) return rc;
        }
    }
    {
        // M <= 64: one pass of the stream-K decode kernel (larger M with K % 8 != 0: 64-row slabs of it)
        ckb_gemm_problem q = p;
        q.M = std::min(p.M, 64);
        if(int rc = plan_wq_decode(q, *plan)) { if(rc == CKB_ERR_UNSUPPORTED) { plan->supported = 0; return CKB_OK; } return rc; }
        plan->kernel_id = KID_WQ_DECODE;
        plan->supported = 1;
    }
    return CKB_OK;
}

int ckb_gemm_run(const ckb_gemm_problem* pp, const ckb_gemm_plan* plan, void* workspace, void* stream)
{
    if(!pp || !plan) return set_error(CKB_ERR_INVALID, "null problem/plan");
    if(!plan->supported) return set_error(CKB_ERR_UNSUPPORTED, "plan is not supported: %s", last_error().c_str());
    if(pp->M == 0 || pp->N == 0) return CKB_OK;
    if(!device_info().ok) return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch(plan->kernel_id)
    {
    case KID_GEMM_SS: return run_gemm_ss(*pp, *plan, st);
    case KID_WQ_PREFILL: return run_wq_prefill(*pp, *plan, workspace, st);
    case KID_WQ_FUSED: return run_wq_fused(*pp, *plan, st);
    case KID_WQ_GEMV:
    case KID_WQ_GEMV2:
    case KID_WQ_DECODE: {
        const int kid  = plan->kernel_id;
        const int slab = kid == KID_WQ_DECODE ? 64 : 32; // rows one launch of the kernel handles
        auto plan_one = [&](const ckb_gemm_problem& q, ckb_gemm_plan& pl) {
            return kid == KID_WQ_GEMV2 ? plan_wq_gemv2(q, pl) : (kid == KID_WQ_GEMV ? plan_wq_gemv(q, pl) : plan_wq_decode(q, pl));
        };
        auto run_one = [&](const ckb_gemm_problem& q, const ckb_gemm_plan& pl) {
            return kid == KID_WQ_GEMV2 ? run_wq_gemv2(q, pl, st) : (kid == KID_WQ_GEMV ? run_wq_gemv(q, pl, st) : run_wq_decode(q, pl, workspace, st));
        };
        if(pp->M <= slab) return run_one(*pp, *plan);
        // slabs of rows; each slab needs its own activation tensor map
        for(int m0 = 0; m0 < pp->M; m0 += slab)
        {
            ckb_gemm_problem q = *pp;
            q.M = std::min(slab, pp->M - m0);
            q.a = static_cast<const uint8_t*>(pp->a) + size_t(m0) * pp->lda * 2;
            q.e = static_cast<uint8_t*>(pp->e) + size_t(m0) * pp->lde * dtype_bytes(pp->e_dtype);
            for(int i = 0; i < pp->num_d; ++i)
                q.d[i] = static_cast<const uint8_t*>(pp->d[i]) + size_t(m0) * pp->d_sm[i] * dtype_bytes(pp->d_dtype[i]);
            ckb_gemm_plan sub;
            memset(&sub, 0, sizeof sub);
            if(int rc = plan_one(q, sub)) return rc;
            sub.supported = 1;
            if(int rc = run_one(q, sub)) return rc;
        }
        return CKB_OK;
    }
    default: return set_error(CKB_ERR_INVALID, "bad kernel id %d", plan->kernel_id);
    }
}

uint64_t ckb_static_scratch_bytes(void) { return CKB_STATIC_SCRATCH_BYTES; }

int ckb_gemm_run_static_scratch(const ckb_gemm_problem* pp, const ckb_gemm_plan* plan, void* stream)
{
    if(!pp || !plan) return set_error(CKB_ERR_INVALID, "null problem/plan");
    if(!plan->supported) return set_error(CKB_ERR_UNSUPPORTED, "plan is not supported: %s", last_error().c_str());
    if(plan->kernel_id != KID_WQ_DECODE)
        return set_error(CKB_ERR_UNSUPPORTED, "static scratch: kernel %s uses no scratch memory; call ckb_gemm_run with workspace = NULL", plan->kernel_name);
    if(pp->M > 64) return set_error(CKB_ERR_UNSUPPORTED, "static scratch: M = %d; this entry point serves one launch of the decode kernel (M <= 64)", pp->M);
    if(pp->M == 0 || pp->N == 0) return CKB_OK;
    if(!device_info().ok) return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device");
    return run_wq_decode_static(*pp, *plan, static_cast<cudaStream_t>(stream));
}

uint64_t ckb_gemm_host_staging_bytes(const ckb_gemm_problem* p)
{
    if(!p) return 0;
    auto al = [](uint64_t v) { return (v + 255) & ~uint64_t(255); };
    uint64_t b = al(uint64_t(p->M) * p->lda * dtype_bytes(p->a_dtype));
    for(int i = 0; i < p->num_d; ++i)
        b += al((uint64_t(p->M - 1) * p->d_sm[i] + uint64_t(p->N - 1) * p->d_sn[i] + 1) * dtype_bytes(p->d_dtype[i]));
    b += al(uint64_t(p->M) * p->lde * dtype_bytes(p->e_dtype));
    return b;
}

// Plans of the host-buffer entry points, keyed by the bytes of the staged problem.  Entries are never mutated after insertion
// and live until the cache is reset (when it exceeds its bound), so the returned pointer stays valid for the call.
namespace {
struct PlanCache
{
    struct Entry { ckb_gemm_problem prob; ckb_gemm_plan plan; };
    std::mutex mu;
    std::unordered_multimap<uint64_t, std::unique_ptr<Entry>> map;
    std::vector<std::unique_ptr<Entry>> retired; // entries of a reset cache stay alive: another thread may still hold one
};
PlanCache& plan_cache() { static PlanCache c; return c; }
thread_local int g_last_rc = CKB_OK;
int last_error_code() { return g_last_rc; }

const ckb_gemm_plan* cached_plan(const ckb_gemm_problem& q)
{
    static const bool enabled = !(getenv("CKB_PLAN_CACHE") && atoi(getenv("CKB_PLAN_CACHE")) == 0);
    uint64_t h = 1469598103934665603ull; // FNV-1a over the struct bytes
    const unsigned char* b = reinterpret_cast<const unsigned char*>(&q);
    for(size_t i = 0; i < sizeof q; ++i) h = (h ^ b[i]) * 1099511628211ull;
    PlanCache& c = plan_cache();
    if(enabled)
    {
        std::lock_guard<std::mutex> lk(c.mu);
        auto range = c.map.equal_range(h);
        for(auto it = range.first; it != range.second; ++it)
            if(memcmp(&it->second->prob, &q, sizeof q) == 0) return &it->second->plan;
    }
    auto e = std::make_unique<PlanCache::Entry>();
    e->prob = q;
    if(int rc = ckb_gemm_query(&q, &e->plan)) { g_last_rc = rc; return nullptr; }
    std::lock_guard<std::mutex> lk(c.mu);
    if(c.map.size() >= 16384 || c.retired.size() >= 16384)
    {
        for(auto& kv : c.map) c.retired.push_back(std::move(kv.second));
        c.map.clear();
        if(c.retired.size() >= 65536) c.retired.erase(c.retired.begin(), c.retired.begin() + 32768);
    }
    const ckb_gemm_plan* out = &e->plan;
    if(enabled) c.map.emplace(h, std::move(e));
    else { c.retired.push_back(std::move(e)); if(c.retired.size() > 1024) c.retired.erase(c.retired.begin(), c.retired.begin() + 512); }
    return out;
}
} // namespace

static int run_host_impl(const ckb_gemm_problem* pp, void* staging, void* workspace, void* stream, bool sync)
{
    if(!pp || !staging) return set_error(CKB_ERR_INVALID, "null problem/staging");
    if(pp->M == 0 || pp->N == 0) return CKB_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    auto al = [](uint64_t v) { return (v + 255) & ~uint64_t(255); };
    ckb_gemm_problem q = *pp;
    uint8_t* cur = static_cast<uint8_t*>(staging);
    const uint64_t a_bytes = (uint64_t(pp->M - 1) * pp->lda + pp->K) * dtype_bytes(pp->a_dtype);
    if(void* src = ((reinterpret_cast<uintptr_t>(pp->a) & 15) == 0 && a_bytes <= (4u << 20)) ? mapped_alias(pp->a) : nullptr)
    {
        const long long n16 = (long long)(a_bytes / 16);
        const int blocks    = int(std::min<long long>(device_info().sm_count, (n16 + 255) / 256 + 1));
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(256); cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if(cudaStreamIsCapturing(st, &cs) != cudaSuccess) { cudaGetLastError(); cs = cudaStreamCaptureStatusNone; }
        cfg.attrs = attr; cfg.numAttrs = cs == cudaStreamCaptureStatusActive ? 1 : 0; // eager PDL launches cost ~20 us of host time
        CKB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, upload_kernel, static_cast<const uint4*>(src), reinterpret_cast<uint4*>(cur), n16,
                                          static_cast<const unsigned char*>(src) + n16 * 16, static_cast<unsigned char*>(cur + n16 * 16),
                                          int(a_bytes % 16)));
    }
    else
        CKB_CUDA_CHECK(cudaMemcpyAsync(cur, pp->a, a_bytes, cudaMemcpyHostToDevice, st));
    q.a = cur;
    cur += al(uint64_t(pp->M) * pp->lda * dtype_bytes(pp->a_dtype));
    for(int i = 0; i < pp->num_d; ++i)
    {
        const uint64_t bytes = (uint64_t(pp->M - 1) * pp->d_sm[i] + uint64_t(pp->N - 1) * pp->d_sn[i] + 1) * dtype_bytes(pp->d_dtype[i]);
        CKB_CUDA_CHECK(cudaMemcpyAsync(cur, pp->d[i], bytes, cudaMemcpyHostToDevice, st));
        q.d[i] = cur;
        cur += al(bytes);
    }
    // E: the epilogue stores straight into mapped pinned host memory (posted PCIe writes) when it can; otherwise staging + copy
    void* e_alias = mapped_alias(pp->e);
    q.e = e_alias ? e_alias : cur;
    // the plan (validation + TMA descriptors) depends only on the staged problem: cached per distinct problem, so a serving
    // loop that re-issues the same GEMMs pays for ckb_gemm_query once (round 1: ~15 us of host time per call went there)
    const ckb_gemm_plan* plan = cached_plan(q);
    if(!plan) return last_error_code();
    if(!plan->supported) return set_error(CKB_ERR_UNSUPPORTED, "%s", last_error().c_str());
    if(int rc = ckb_gemm_run(&q, plan, workspace, stream)) return rc;
    const uint64_t e_bytes = (uint64_t(pp->M - 1) * pp->lde + pp->N) * dtype_bytes(pp->e_dtype);
    if(!e_alias) CKB_CUDA_CHECK(cudaMemcpyAsync(pp->e, cur, e_bytes, cudaMemcpyDeviceToHost, st));
    if(sync) CKB_CUDA_CHECK(cudaStreamSynchronize(st));
    return CKB_OK;
}

int ckb_gemm_run_host(const ckb_gemm_problem* pp, void* staging, void* workspace, void* stream)
{
    return run_host_impl(pp, staging, workspace, stream, true);
}

int ckb_gemm_run_host_async(const ckb_gemm_problem* pp, void* staging, void* workspace, void* stream)
{
    return run_host_impl(pp, staging, workspace, stream, false);
}

int ckb_prepack_b_u8(const void* src, void* dst, int64_t n, void* stream)
{
    if(!src || !dst || n < 0) return set_error(CKB_ERR_INVALID, "bad prepack arguments");
    if(!device_info().ok) return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device");
    return launch_prepack_u8(src, dst, n, static_cast<cudaStream_t>(stream));
}

int ckb_prepack_b_rowmajor_i8(const void* src, int64_t ld_src, int32_t K, int32_t N, void* dst, void* stream)
{
    if(!src || !dst || K < 0 || N < 0 || ld_src < N) return set_error(CKB_ERR_INVALID, "bad prepack arguments");
    if(!device_info().ok) return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device");
    return launch_transpose_i8(src, ld_src, K, N, dst, K, static_cast<cudaStream_t>(stream));
}

int ckb_transpose_a_colmajor(const void* src, int64_t ld_src, int32_t M, int32_t K, int32_t elem_bytes, void* dst, void* stream)
{
    if(!src || !dst || M < 0 || K < 0 || ld_src < M || (elem_bytes != 1 && elem_bytes != 2)) return set_error(CKB_ERR_INVALID, "bad transpose arguments");
    if(!device_info().ok) return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device");
    // src is [K rows][M cols] with pitch ld_src; dst[m * K + k]
    return elem_bytes == 2 ? launch_transpose_16(src, ld_src, K, M, dst, K, static_cast<cudaStream_t>(stream))
                           : launch_transpose_i8(src, ld_src, K, M, dst, K, static_cast<cudaStream_t>(stream));
}

uint64_t ckb_prepack_b_int4_bytes(int32_t N, int32_t K)
{
    if(N <= 0 || K <= 0) return 0;
    return uint64_t((N + 15) / 16) * ((K + 63) / 64) * 512; // [n16 group][k64 block][32 lanes][16 B]
}

int ckb_prepack_b_int4(const void* src, int64_t sk, int64_t sn, int32_t N, int32_t K, void* dst, void* stream)
{
    if(!src || !dst || N < 0 || K < 0) return set_error(CKB_ERR_INVALID, "bad prepack arguments");
    if(!device_info().ok) return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device");
    return launch_prepack_int4(src, sk, sn, N, K, dst, static_cast<cudaStream_t>(stream));
}

uint64_t ckb_grouped_gemm_arg_bytes(int32_t num_b, int32_t num_d)
{
    if(num_b < 1 || num_b > 2 || num_d < 0 || num_d > 1) return 0;
    return uint64_t(grouped_arg_bytes(num_b, num_d));
}

int ckb_grouped_gemm_query(const ckb_gemm_problem* proto, int32_t group_count)
{
    if(!proto) return set_error(CKB_ERR_INVALID, "null problem");
    return check_grouped(*proto, group_count);
}

int ckb_grouped_gemm_run(const ckb_gemm_problem* proto, const void* dev_args, int32_t group_count, int32_t sum_m, void* stream)
{
    if(!proto) return set_error(CKB_ERR_INVALID, "null problem");
    if(group_count < 0 || sum_m < 0) return set_error(CKB_ERR_INVALID, "negative group count / row count");
    if(int rc = check_grouped(*proto, group_count)) return rc;
    if(!device_info().ok) return set_error(CKB_ERR_NO_DEVICE, "no sm_100 device");
    return run_grouped(*proto, dev_args, group_count, sum_m, static_cast<cudaStream_t>(stream));
}

} // extern "C"
