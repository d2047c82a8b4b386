// wq_gemv2.cu -- host side of K2 (W4A16 decode kernel, wq_gemv2_kernel.cuh): slot / ring planning, TMA descriptors, launch.
#include "wq_gemv2_kernel.cuh"

#include <string>
#include <vector>

namespace ckb {
namespace {

int env_int(const char* name, int dflt)
{
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

} // namespace

int launch_gemv2_f0_nt1(int, int, int, const G2Plan*, const G2Args&, int, cudaStream_t);
int launch_gemv2_f0_nt2(int, int, int, const G2Plan*, const G2Args&, int, cudaStream_t);
int launch_gemv2_f0_nt4(int, int, int, const G2Plan*, const G2Args&, int, cudaStream_t);
int launch_gemv2_f1_nt1(int, int, int, const G2Plan*, const G2Args&, int, cudaStream_t);
int launch_gemv2_f1_nt2(int, int, int, const G2Plan*, const G2Args&, int, cudaStream_t);
int launch_gemv2_f1_nt4(int, int, int, const G2Plan*, const G2Args&, int, cudaStream_t);

int launch_gemv2_variant(bool f16, int nt, int gw, int kbu, const G2Plan* gp, const G2Args& a, int grid, cudaStream_t st)
{
    if(nt == 1) return f16 ? launch_gemv2_f1_nt1(nt, gw, kbu, gp, a, grid, st) : launch_gemv2_f0_nt1(nt, gw, kbu, gp, a, grid, st);
    if(nt == 2) return f16 ? launch_gemv2_f1_nt2(nt, gw, kbu, gp, a, grid, st) : launch_gemv2_f0_nt2(nt, gw, kbu, gp, a, grid, st);
    return f16 ? launch_gemv2_f1_nt4(nt, gw, kbu, gp, a, grid, st) : launch_gemv2_f0_nt4(nt, gw, kbu, gp, a, grid, st);
}

// Host-side schedule of the decode kernel (also exercised without a GPU by tests/test_capi_cpu.py through
// ckb_gemv_schedule): how the n16 channel groups and the K range are cut into (CTA, pass, stage, warp) work.
static int schedule_wq_gemv2_cap(int M, int N, int K, bool zp, int sms, G2Sched& s, int mode, int gcap);

int schedule_wq_gemv2(int M, int N, int K, bool zp, int sms, G2Sched& s, int mode)
{
    // groups per slot are capped by what a warp can hold; when even one ring stage of such a slot does not fit in shared memory
    // (int8 weights: 1 KB per group and 64-k block, e.g. cfg3's N = 14336 -> 7 groups per CTA), the slots are made smaller: more
    // passes.  A schedule with empty slots (more slots than groups) is not accepted: the kernel assumes every slot holds a group.
    int gcap = 99;
    for(;;)
    {
        const int rc = schedule_wq_gemv2_cap(M, N, K, zp, sms, s, mode, gcap);
        if(rc == CKB_OK) return (gcap == 99 || s.slot_q >= 1) ? CKB_OK : set_error(CKB_ERR_UNSUPPORTED, "W8 decode kernel: no schedule without empty slots fits in shared memory");
        if(!(mode & 8) || (mode & 1) || s.grid < 1 || s.passes < 1 || s.g_hi < 2) return rc;
        // measured (profiles/r2_w8_multipass.txt): the multi-pass form beats the stream-K kernel by 10 - 20 % at 14336 x 4096; at
        // 28672 x 8192 it lost 2 - 8 % to stream-K with two CTAs per SM, and wins by 7 - 8 % now that stream-K runs one CTA per SM
        // (plan_wq_decode; 64.3 / 76.6 us against 69.2 / 83.3 us at M = 1 / 16, profiles/r2_streamk_two_cta_defect.txt): no size limit
        const int per_cta = (s.n_groups + s.grid - 1) / s.grid;
        gcap = std::min((per_cta + s.passes) / (s.passes + 1), s.g_hi - 1); // groups per slot with one more pass
        if(gcap < 1) return rc;
    }
}

static int schedule_wq_gemv2_cap(int M, int N, int K, bool zp, int sms, G2Sched& s, int mode, int gcap)
{
    const bool w8 = mode & 8; // int8-weight formats: 1024 B per (16-channel group, 64-k block), no scale / zero-point rows in the stage
    const bool fixed = mode & 1, rtn = (mode & 2) || (mode & 4) || w8; // bit 2 (peer-store epilogue): 128-k units like RTN, fp32-scale numerics
    if(M < 1 || M > 32) return set_error(CKB_ERR_UNSUPPORTED, "W4 decode kernel handles 1 <= M <= 32 per launch");
    s.nt = M <= 8 ? 1 : (M <= 16 ? 2 : 4);
    s.rb = s.nt == 1 ? (M <= 1 ? 1 : (M <= 2 ? 2 : (M <= 4 ? 4 : 8))) : 8 * s.nt;
    const int nw   = g2_warps(s.nt);
    const int gwmx = g2_gw_max(s.nt);
    s.n_groups = (N + 15) / 16;
    s.kb_total = (K + 63) / 64;
    s.grid     = std::max(1, std::min(sms, s.n_groups));
    s.grid     = std::max(1, std::min(env_int("CKB_GEMV_GRID", s.grid), s.n_groups));
    // groups per slot: up to 2 x GW with the warps split into two channel halves (M <= 8), GW otherwise
    const int gmax    = std::max(1, std::min({(s.nt == 1 ? 2 : 1) * gwmx, fixed ? 99 : env_int("CKB_G2_GMAX", 99), gcap}));
    const int per_cta = (s.n_groups + s.grid - 1) / s.grid;
    s.passes  = (per_cta + gmax - 1) / gmax;
    s.n_slots = s.grid * s.passes;
    s.slot_q  = s.n_groups / s.n_slots;
    s.slot_r  = s.n_groups % s.n_slots;
    s.g_hi    = s.slot_q + (s.slot_r ? 1 : 0);
    // 16 warps: all on K (1 x 16) while a warp can hold the slot's groups in <= 64 registers (<= 2), else 2 x 8
    s.wn_log2 = (s.nt == 1 && s.g_hi > 2) ? 1 : 0;
    if(s.nt == 1 && !fixed) s.wn_log2 = std::max(0, std::min(1, env_int("CKB_G2_WN", s.wn_log2 + 1) - 1));
    // fixed K order: the warp grid and the K units of a stage must not depend on N or on the grid (see CKB_FLAG_FIXED_K_ORDER)
    if(fixed) s.wn_log2 = s.nt == 1 ? 1 : 0;
    if((s.g_hi + (1 << s.wn_log2) - 1 >> s.wn_log2) > gwmx) s.wn_log2 = 1;
    const int wn = 1 << s.wn_log2, wk = nw / wn;
    s.gw      = (s.g_hi + wn - 1) / wn;
    // fixed K order: a warp with one group runs two interleaved accumulation chains (NCH = 2), which is a different fp32
    // order -- keep at least two group slots per warp (the second one may stay empty)
    if(fixed) s.gw = std::max(s.gw, 2);
    const int g_alloc = wn * s.gw; // groups a stage has room for (a warp with gw - 1 groups reads one dead group)
    auto al = [](int v, int a) { return (v + a - 1) / a * a; };
    const int red_bytes = nw * s.gw * s.nt * 512;
    const int bar_bytes = 2 * G2_MAX_RING * 8;
    // Stage shape: a warp takes `upw` K units of `kbu` 64-k blocks per stage.  Measured (profiles/r2_gemv_stage_shapes.txt):
    // the per-stage cost (barrier round trip, ~50 instructions per warp) is what hurts, so stages are made as long along K
    // as the ring allows -- a resident problem (every stage of the CTA fits the ring at once: no refills, no empty
    // barriers) first, then kbu = 2 (one scale application per 128 k) with >= 2 stages in the 112 KB that keep two CTAs per
    // SM (the next GEMM of the stream prefetches its weights while this one computes), then the whole SM.
    int best_kbu = 0, best_upw = 1, best_ring = 0, best_stage = 0, best_score = -1;
    for(int kbu : {2, 1})
        for(int upw : {4, 2, 1})
        {
            if(fixed) { if(kbu != 2 || upw != 1) continue; }
            else
            {
                if(int e = env_int("CKB_G2_KBU", 0)) { if(e != kbu) continue; }
                if(int e = env_int("CKB_G2_UPW", 0)) { if(e != upw) continue; }
            }
            if(rtn && kbu != 2) continue;
            if(kbu == 1 && upw > 1 && s.rb == 4) continue; // the swizzle phase of the activation rows would depend on the unit
            const int chk    = wk * upw * kbu, rows = chk / 2;
            if(chk > 256 || (upw > 1 && wk * (upw / 2) * kbu >= s.kb_total)) continue; // TMA box limit; no point in empty units
            const int a_b    = al(chk * s.rb * 128, 1024);
            const int w_b    = g_alloc * chk * (w8 ? 1024 : 512);
            const int s_b    = w8 ? 0 : al(rows * 32 * g_alloc + 64, 128);
            const int z_b    = (zp && !w8) ? al(rows * 16 * g_alloc + 64, 128) : 0;
            const int stage  = al(a_b + w_b + s_b + z_b, 1024);
            const int chunks = (s.kb_total + chk - 1) / chk;
            const int steps  = s.passes * chunks;
            const int own    = s.passes > 1 ? red_bytes : 0;
            for(int budget : {112 * 1024, 224 * 1024})
            {
                if(int e = env_int("CKB_GEMV_SMEM_KB", 0)) budget = e * 1024;
                const int ring = std::min({G2_MAX_RING, steps, (budget - bar_bytes - 1024 - own) / stage});
                if(ring < 1 || (ring < 2 && steps > 1)) continue;
                if(!own && ring * stage < red_bytes) continue;
                const bool resident = ring >= steps, two_ctas = budget <= 112 * 1024;
                const int score = (resident ? 1000 : 0) + (kbu == 2 ? 400 : 0) + (two_ctas ? (resident ? 200 : 20) : 0) + (resident ? 40 : std::min(ring, 4) * 10) +
                                  (upw == 2 ? 3 : (upw == 4 ? 2 : 0)) * (resident ? 1 : 30);
                if(score > best_score) { best_score = score; best_kbu = kbu; best_upw = upw; best_ring = ring; best_stage = stage; }
            }
        }
    if(best_kbu == 0) return set_error(CKB_ERR_UNSUPPORTED, "W4 decode kernel: a ring stage does not fit in shared memory");
    s.kbu = best_kbu; s.upw = best_upw; s.ring = best_ring; s.stage_bytes = best_stage;
    const int chk = wk * s.upw * s.kbu, rows = chk / 2;
    s.n_chunks = (s.kb_total + chk - 1) / chk;
    s.off_w    = al(chk * s.rb * 128, 1024);
    s.off_s    = s.off_w + g_alloc * chk * (w8 ? 1024 : 512);
    s.off_z    = s.off_s + (w8 ? 0 : al(rows * 32 * g_alloc + 64, 128));
    int end    = s.ring * s.stage_bytes;
    if(s.passes > 1) { s.red_off = end; end += red_bytes; }
    else s.red_off = 0;
    s.bar_off    = end;
    s.smem_bytes = s.bar_off + bar_bytes + 1024;
    return CKB_OK;
}

int plan_wq_gemv2(const ckb_gemm_problem& p, ckb_gemm_plan& plan)
{
    static_assert(sizeof(G2Plan) <= sizeof(plan.opaque), "opaque too small");
    G2Plan* gp     = reinterpret_cast<G2Plan*>(plan.opaque);
    const bool f16 = p.a_dtype == CKB_F16;
    const bool zp  = p.zero != nullptr;
    const int sms  = device_info().sm_count > 0 ? device_info().sm_count : 148;
    G2Sched s{};
    const bool w8  = p.b_op != CKB_BOP_GROUP && p.b_op != CKB_BOP_GROUP_RTN;
    // functor epilogue (D tensors, a CDE functor, fp32 E) on the fp32-scale W4 format: served by the instantiation with the epilogue
    // inlined, which exists for 128-k units only (mode bit 2 asks the schedule for them, as the peer-store epilogue does)
    const bool functor = !(p.cde_op == CKB_CDE_PASSTHROUGH && (p.e_dtype == CKB_BF16 || p.e_dtype == CKB_F16));
    const bool epi     = functor && !w8 && p.b_op == CKB_BOP_GROUP && p.num_peers == 0;
    const int mode = ((p.flags & CKB_FLAG_FIXED_K_ORDER) ? 1 : 0) | (p.b_op == CKB_BOP_GROUP_RTN ? 2 : 0) | ((p.num_peers > 0 || epi) ? 4 : 0) | (w8 ? 8 : 0);
    if(int rc = schedule_wq_gemv2(p.M, p.N, p.K, zp, sms, s, mode)) return rc;
    gp->nt = s.nt; gp->gw = s.gw; gp->passes = s.passes; gp->slot_q = s.slot_q; gp->slot_r = s.slot_r; gp->g_hi = s.g_hi;
    gp->wn_log2 = s.wn_log2; gp->kbu = s.kbu; gp->upw = s.upw; gp->n_chunks = s.n_chunks; gp->ring = s.ring; gp->stage_bytes = s.stage_bytes;
    gp->off_w = s.off_w; gp->off_s = s.off_s; gp->off_z = s.off_z;
    gp->red_off = s.red_off; gp->bar_off = s.bar_off; gp->smem_bytes = s.smem_bytes; gp->rb = s.rb;
    gp->rtn = (mode & 2) ? 1 : 0;
    gp->wf  = !w8 ? 0 : (p.b_op == CKB_BOP_MULTIPLY ? 1 : (p.b_op == CKB_BOP_FPAINTB ? 2 : 3));

    const int chk = (g2_warps(s.nt) >> s.wn_log2) * s.upw * s.kbu;
    const CUtensorMapDataType adt = f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    {   // activations as [K/64][M][64]: box = one stage of k-blocks x rb rows x 64 k (rows past M are zero-filled)
        const uint64_t dims[3] = {64, uint64_t(p.M), uint64_t(s.kb_total)};
        const uint64_t str[2]  = {uint64_t(p.lda) * 2, 128};
        const uint32_t box[3]  = {64, uint32_t(s.rb), uint32_t(chk)};
        if(int rc = encode_tmap_nd(&gp->tmA, adt, 3, p.a, dims, str, box,
                                   s.rb >= 4 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
    }
    const int kg_total = (p.K + 127) / 128;
    for(int i = 0; i < 2 && w8; ++i)
    {
        // int8 weights B[n][k] (K-major, pitch ldb) viewed as (64 B, N rows, K/64 blocks): a box lands as [k-block][row][64 B], so the
        // eight rows a quarter-warp reads are contiguous (conflict-free LDS.128)
        const int gb = i == 0 ? s.g_hi : std::max(1, s.g_hi - 1);
        const uint64_t dims[3] = {64, uint64_t(p.N), uint64_t(s.kb_total)};
        const uint64_t str[2]  = {uint64_t(p.ldb), 64};
        const uint32_t box[3]  = {64, uint32_t(16 * gb), uint32_t(chk)};
        if(int rc = encode_tmap_nd(&gp->tmW[i], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, p.b, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        gp->tmS[i] = gp->tmW[i];
        gp->tmZ[i] = gp->tmW[i];
    }
    for(int i = 0; i < 2 && !w8; ++i)
    {
        const int gb = i == 0 ? s.g_hi : std::max(1, s.g_hi - 1);
        {   // packed weights [n16 group][k64 block][128 words]
            const uint64_t dims[3] = {128, uint64_t(s.kb_total), uint64_t(s.n_groups)};
            const uint64_t str[2]  = {512, uint64_t(s.kb_total) * 512};
            const uint32_t box[3]  = {128, uint32_t(chk), uint32_t(gb)};
            if(int rc = encode_tmap_nd(&gp->tmW[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, p.b, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        {   // group scales [K/128][N]
            const uint64_t dims[2] = {uint64_t(p.N), uint64_t(kg_total)};
            const uint64_t str[1]  = {uint64_t(p.N) * 2};
            const uint32_t box[2]  = {uint32_t(16 * gb), uint32_t(chk / 2)};
            if(int rc = encode_tmap_nd(&gp->tmS[i], adt, 2, p.scale, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        if(zp)
        {
            const uint64_t dims[2] = {uint64_t(p.N), uint64_t(kg_total)};
            const uint64_t str[1]  = {uint64_t(p.N)};
            const uint32_t box[2]  = {uint32_t(16 * gb), uint32_t(chk / 2)};
            if(int rc = encode_tmap_nd(&gp->tmZ[i], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, p.zero, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        else
            gp->tmZ[i] = gp->tmS[i];
    }
    plan.grid            = s.grid;
    plan.block           = g2_threads(s.nt);
    plan.smem_bytes      = gp->smem_bytes;
    plan.split_units     = 1;
    plan.workspace_bytes = 0;
    static const char* wfn[] = {"w4g128", "w8mul", "w8fpaintb", "w8pass"};
    snprintf(plan.kernel_name, sizeof plan.kernel_name, "wq_gemv2_%s%s_%s_hmma_%dx%dwarps_g%d_p%d_m%d_u%dx%d_r%d%s%s", wfn[gp->wf], (mode & 2) ? "rtn" : "",
             f16 ? "f16" : "bf16", 1 << s.wn_log2, g2_warps(s.nt) >> s.wn_log2, s.gw, s.passes, 8 * s.nt, s.upw, s.kbu, s.ring,
             (mode & 1) ? "_korder" : "", (epi && s.kbu == 2) ? "_epi" : "");
    return CKB_OK;
}

// consumer side of the fused all-gather (ckb_peer_wait): one thread
static __global__ void peer_wait_kernel(const unsigned* flag, unsigned* epoch_state, unsigned per_epoch)
{
    ptx::pdl_launch_dependents();
    const unsigned epoch  = *epoch_state + 1;
    *epoch_state          = epoch;
    const unsigned target = epoch * per_epoch;
    // poll with relaxed loads (a system-scope acquire costs microseconds per iteration), then ONE acquire load pairs with the
    // producers' release adds
    while(int(ptx::ld_relaxed_sys_u32(flag) - target) < 0) { }
    (void)ptx::ld_acquire_sys_u32(flag);
}

int launch_peer_wait(void* flag, void* epoch_state, unsigned per_epoch, cudaStream_t stream)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(1); cfg.blockDim = dim3(1); cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if(cudaStreamIsCapturing(stream, &cs) != cudaSuccess) { cudaGetLastError(); cs = cudaStreamCaptureStatusNone; }
    cfg.attrs = attr; cfg.numAttrs = cs == cudaStreamCaptureStatusActive ? 1 : 0;
    CKB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, peer_wait_kernel, static_cast<const unsigned*>(flag), static_cast<unsigned*>(epoch_state), per_epoch));
    return CKB_OK;
}

int run_wq_gemv2(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream)
{
    const G2Plan* gp = reinterpret_cast<const G2Plan*>(plan.opaque);
    G2Args a{};
    a.M = p.M; a.N = p.N; a.K = p.K;
    a.kb_total = (p.K + 63) / 64;
    a.passes   = gp->passes;
    a.slot_q   = gp->slot_q;
    a.slot_r   = gp->slot_r;
    a.g_hi     = gp->g_hi;
    a.wn_log2  = gp->wn_log2;
    a.upw      = gp->upw;
    a.n_chunks = gp->n_chunks;
    a.ring     = gp->ring;
    a.stage_bytes = gp->stage_bytes;
    a.off_w = gp->off_w; a.off_s = gp->off_s; a.off_z = gp->off_z;
    a.red_off    = gp->red_off;
    a.bar_off    = gp->bar_off;
    a.rb         = gp->rb;
    a.zero_const = p.zero_const;
    a.has_zero   = p.zero != nullptr;
    a.rtn        = gp->rtn;
    a.wf         = gp->wf;
    a.ch_scale   = static_cast<const uint16_t*>(p.scale);
    // the optional paths live in the EXT instantiation (128-k units, fp32-scale numerics): peers force it through the schedule
    // (mode bit 2), the inputs-ready hint is honoured when the plan happens to use 128-k units and is ignored otherwise
    a.nodep      = ((p.flags & CKB_FLAG_INPUTS_READY) && gp->kbu == 2 && !gp->rtn) ? 1 : 0;
    a.ext        = (a.nodep || p.num_peers > 0) ? 1 : 0;
    {
        const bool functor = !(p.cde_op == CKB_CDE_PASSTHROUGH && (p.e_dtype == CKB_BF16 || p.e_dtype == CKB_F16));
        a.epi = (functor && gp->wf == 0 && !gp->rtn && gp->kbu == 2 && !a.ext) ? 1 : 0;
    }
    {
        // programmatic dependent launch pays inside a CUDA graph (PDL edges: the next kernel's prologue and weight prefill
        // overlap this one) but an EAGER launch with the attribute costs ~20 us of host time against ~6 us without
        // (tests/host_cost_probe.py), so it is only requested while the stream is being captured
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if(cudaStreamIsCapturing(stream, &cs) != cudaSuccess) { cudaGetLastError(); cs = cudaStreamCaptureStatusNone; }
        a.pdl = cs == cudaStreamCaptureStatusActive;
        a.pdl = env_int("CKB_GEMV_PDL", a.pdl);
    }
    a.num_peers = p.num_peers;
    a.peer_rank = p.peer_rank;
    for(int g = 0; g < p.num_peers && g < CKB_MAX_PEERS; ++g) { a.e_peer[g] = p.e_peer[g]; a.flag_peer[g] = static_cast<unsigned*>(p.flag_peer[g]); }
    a.dbg   = env_int("CKB_GEMV_DBG", 0);
    a.trace = nullptr;
    fill_epi_params(p, a.ep);
#ifdef CKB_ENABLE_TRACE
    if(const char* tf = getenv("CKB_G2_TRACE"))
    {
        // debug: per-warp timeline of up to 64 consecutive launches (each launch gets its own region, so a captured graph
        // records the steady state under programmatic dependent launch); dumped at process exit
        static constexpr int MAXL = 64, PER = 160 * 16 * 10;
        static long long* buf = nullptr;
        static int n_launch = 0;
        static int grids[MAXL];
        static std::string path;
        if(!buf)
        {
            cudaMalloc(&buf, size_t(MAXL) * PER * 8);
            cudaMemset(buf, 0, size_t(MAXL) * PER * 8);
            path = tf;
            atexit([] {
                cudaDeviceSynchronize();
                std::vector<long long> h(size_t(MAXL) * PER);
                cudaMemcpy(h.data(), buf, h.size() * 8, cudaMemcpyDeviceToHost);
                if(FILE* f = fopen(path.c_str(), "w"))
                {
                    for(int l = 0; l < std::min(n_launch, MAXL); ++l)
                        for(int c = 0; c < grids[l] * 16; ++c)
                        {
                            fprintf(f, "%d %d %d", l, c / 16, c % 16);
                            for(int k = 0; k < 10; ++k) fprintf(f, " %lld", h[size_t(l) * PER + c * 10 + k]);
                            fprintf(f, "\n");
                        }
                    fclose(f);
                }
            });
        }
        if(n_launch < MAXL && plan.grid <= 160) { a.trace = buf + size_t(n_launch) * PER; grids[n_launch] = plan.grid; ++n_launch; }
    }
#endif
    return launch_gemv2_variant(p.a_dtype == CKB_F16, gp->nt, gp->gw, gp->kbu, gp, a, plan.grid, stream);
}

} // namespace ckb
