// wq_gemv.cu -- host side of K2 (W4A16 decode kernel, wq_gemv_kernel.cuh): slot / ring planning, TMA descriptors, launch.
#include "wq_gemv_kernel.cuh"

#include <string>
#include <vector>

namespace ckb {
namespace {

struct GVLayout { int off_w, off_s, off_z, stage_bytes, scratch; };
GVLayout gv_layout(int M, int nt, int g, bool zp, int kbw)
{
    GVLayout l;
    l.off_w       = (kbw * GV_CHUNK_KB * M * 128 + 1023) & ~1023;
    l.off_s       = l.off_w + g * kbw * GV_CHUNK_KB * 512;
    l.off_z       = l.off_s + kbw * GV_PAIRS * g * 32;
    l.stage_bytes = (l.off_z + (zp ? kbw * GV_PAIRS * g * 16 : 0) + 1023) & ~1023;
    l.scratch     = GV_PAIRS * g * nt * 512;
    return l;
}

} // namespace

int launch_gemv_f0_nt1(int, int, const GVPlan*, const GVArgs&, int, cudaStream_t);
int launch_gemv_f0_nt2(int, int, const GVPlan*, const GVArgs&, int, cudaStream_t);
int launch_gemv_f0_nt4(int, int, const GVPlan*, const GVArgs&, int, cudaStream_t);
int launch_gemv_f1_nt1(int, int, const GVPlan*, const GVArgs&, int, cudaStream_t);
int launch_gemv_f1_nt2(int, int, const GVPlan*, const GVArgs&, int, cudaStream_t);
int launch_gemv_f1_nt4(int, int, const GVPlan*, const GVArgs&, int, cudaStream_t);

int launch_gemv_variant(bool f16, int nt, int gw, const GVPlan* gp, const GVArgs& a, int grid, cudaStream_t st)
{
    if(nt == 1) return f16 ? launch_gemv_f1_nt1(nt, gw, gp, a, grid, st) : launch_gemv_f0_nt1(nt, gw, gp, a, grid, st);
    if(nt == 2) return f16 ? launch_gemv_f1_nt2(nt, gw, gp, a, grid, st) : launch_gemv_f0_nt2(nt, gw, gp, a, grid, st);
    return f16 ? launch_gemv_f1_nt4(nt, gw, gp, a, grid, st) : launch_gemv_f0_nt4(nt, gw, gp, a, grid, st);
}

int plan_wq_gemv(const ckb_gemm_problem& p, ckb_gemm_plan& plan)
{
    static_assert(sizeof(GVPlan) <= sizeof(plan.opaque), "opaque too small");
    GVPlan* gp     = reinterpret_cast<GVPlan*>(plan.opaque);
    const bool f16 = p.a_dtype == CKB_F16;
    const bool zp  = p.zero != nullptr;
    const int nt   = p.M <= 8 ? 1 : (p.M <= 16 ? 2 : (p.M <= 32 ? 4 : 8));
    const int n_groups = (p.N + 15) / 16, kb_total = (p.K + 63) / 64, kg_total = (p.K + 127) / 128;
    int n_chunks       = (kb_total + GV_CHUNK_KB - 1) / GV_CHUNK_KB;
    const int sms      = device_info().sm_count > 0 ? device_info().sm_count : 148;
    int grid           = std::min(sms, n_groups);
    if(const char* e = getenv("CKB_GEMV_GRID")) grid = std::max(1, std::min(atoi(e), n_groups));
    int budget = 224 * 1024 - 1024 /*align*/ - 256 /*barriers*/;
    if(const char* e = getenv("CKB_GEMV_SMEM_KB")) budget = std::min(budget, atoi(e) * 1024 - 1280);
    // largest group count per slot whose ring (>= 2 stages, or the whole K range) + scratch fit
    int gmax = gv_gmax(nt);
    while(gmax > 1)
    {
        const GVLayout l = gv_layout(p.M, nt, gmax, zp, 1);
        if(std::min(2, n_chunks) * l.stage_bytes + l.scratch <= budget) break;
        --gmax;
    }
    const int per_cta = (n_groups + grid - 1) / grid;
    const int passes  = (per_cta + gmax - 1) / gmax;
    const int n_slots = grid * passes;
    const int g_hi    = (n_groups + n_slots - 1) / n_slots;
    const int g_lo    = std::max(1, n_groups / n_slots);
    const int kbw     = gv_kbw(nt, g_hi);   // 64-k blocks per warp and stage (compile-time property of the variant)
    const int chunk_kb = GV_CHUNK_KB * kbw;
    n_chunks          = (kb_total + chunk_kb - 1) / chunk_kb;
    const GVLayout l  = gv_layout(p.M, nt, g_hi, zp, kbw);
    gp->off_w = l.off_w; gp->off_s = l.off_s; gp->off_z = l.off_z; gp->stage_bytes = l.stage_bytes;
    {
        // small variants (NT = 1, GW <= 3) are compiled for two resident CTAs per SM so that, under programmatic dependent
        // launch, the NEXT kernel's prologue and weight prefill overlap this one: keep the footprint under half an SM
        // when the whole problem still fits in the ring (no refills); otherwise a deep ring matters more
        const int gw   = g_hi;
        const int half = 112 * 1024 - 1280;
        if(nt == 1 && gw <= 3 && !getenv("CKB_GEMV_SMEM_KB") && (half - l.scratch) / l.stage_bytes >= passes * n_chunks)
            budget = half;
    }
    gp->stages = std::min({GV_MAX_STAGES, (budget - l.scratch) / l.stage_bytes, passes * n_chunks});
    if(gp->stages < 1) return set_error(CKB_ERR_UNSUPPORTED, "W4 decode kernel: stage does not fit in shared memory");
    gp->scratch_off = gp->stages * gp->stage_bytes;
    gp->bar_off     = gp->scratch_off + l.scratch;
    gp->smem_bytes  = gp->bar_off + 256 + 1024;
    gp->nt = nt; gp->passes = passes; gp->g_hi = g_hi;

    const CUtensorMapDataType adt = f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    {   // activations as [K/64][M][64]: box = 16 k-blocks x M rows x 64 k, 128B swizzle
        const uint64_t dims[3] = {64, uint64_t(p.M), uint64_t(kb_total)};
        const uint64_t str[2]  = {uint64_t(p.lda) * 2, 128};
        const uint32_t box[3]  = {64, uint32_t(p.M), uint32_t(chunk_kb)};
        if(int rc = encode_tmap_nd(&gp->tmA, adt, 3, p.a, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    }
    for(int i = 0; i < 2; ++i)
    {
        const int gb = i == 0 ? g_hi : g_lo;
        {   // packed weights [n16 group][k64 block][128 words]
            const uint64_t dims[3] = {128, uint64_t(kb_total), uint64_t(n_groups)};
            const uint64_t str[2]  = {512, uint64_t(kb_total) * 512};
            const uint32_t box[3]  = {128, uint32_t(chunk_kb), uint32_t(gb)};
            if(int rc = encode_tmap_nd(&gp->tmW[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, p.b, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        {   // group scales [K/128][N]
            const uint64_t dims[2] = {uint64_t(p.N), uint64_t(kg_total)};
            const uint64_t str[1]  = {uint64_t(p.N) * 2};
            const uint32_t box[2]  = {uint32_t(16 * gb), uint32_t(GV_PAIRS * kbw)};
            if(int rc = encode_tmap_nd(&gp->tmS[i], adt, 2, p.scale, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        if(zp)
        {
            const uint64_t dims[2] = {uint64_t(p.N), uint64_t(kg_total)};
            const uint64_t str[1]  = {uint64_t(p.N)};
            const uint32_t box[2]  = {uint32_t(16 * gb), uint32_t(GV_PAIRS * kbw)};
            if(int rc = encode_tmap_nd(&gp->tmZ[i], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, p.zero, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
        }
        else
            gp->tmZ[i] = gp->tmS[i];
    }
    plan.grid            = grid;
    plan.block           = GV_THREADS;
    plan.smem_bytes      = gp->smem_bytes;
    plan.split_units     = 1;
    plan.workspace_bytes = 0;
    snprintf(plan.kernel_name, sizeof plan.kernel_name, "wq_gemv_w4g128_%s_hmma_nsplit_g%d_p%d_m%d_s%d", f16 ? "f16" : "bf16",
             g_hi, passes, 8 * nt, gp->stages);
    return CKB_OK;
}

int run_wq_gemv(const ckb_gemm_problem& p, const ckb_gemm_plan& plan, cudaStream_t stream)
{
    const GVPlan* gp = reinterpret_cast<const GVPlan*>(plan.opaque);
    GVArgs a{};
    a.M = p.M; a.N = p.N; a.K = p.K;
    a.n_groups = (p.N + 15) / 16;
    a.kb_total = (p.K + 63) / 64;
    a.passes   = gp->passes;
    a.n_slots  = plan.grid * gp->passes;
    a.g_hi     = gp->g_hi;
    a.stages   = gp->stages;
    a.stage_bytes = gp->stage_bytes;
    a.off_w = gp->off_w; a.off_s = gp->off_s; a.off_z = gp->off_z;
    a.scratch_off = gp->scratch_off;
    a.bar_off     = gp->bar_off;
    a.zero_const  = p.zero_const;
    a.has_zero    = p.zero != nullptr;
    a.act_rows    = p.M;
    {
        // programmatic dependent launch pays inside a CUDA graph (PDL edges: the next kernel's prologue and weight prefill
        // overlap this one) but an EAGER launch with the attribute costs ~20 us of host time against ~6 us without
        // (tests/host_cost_probe.py), so it is only requested while the stream is being captured
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if(cudaStreamIsCapturing(stream, &cs) != cudaSuccess) { cudaGetLastError(); cs = cudaStreamCaptureStatusNone; }
        a.pdl = cs == cudaStreamCaptureStatusActive;
        if(const char* e = getenv("CKB_GEMV_PDL")) a.pdl = atoi(e);
    }
    a.dbg         = getenv("CKB_GEMV_DBG") ? atoi(getenv("CKB_GEMV_DBG")) : 0;
    a.w           = static_cast<const uint8_t*>(p.b);
    a.gtrace      = nullptr;
    a.gspan       = nullptr;
    fill_epi_params(p, a.ep);
    const bool f16 = p.a_dtype == CKB_F16;
    auto go = [&]() -> int { return launch_gemv_variant(f16, gp->nt, gp->g_hi, gp, a, plan.grid, stream); };
#ifdef CKB_ENABLE_TRACE
    if(const char* sf = getenv("CKB_GSPAN"))
    {
        // debug: spans of consecutive launches (no synchronisation in between); dumped at process exit
        constexpr int MAXL = 4096;
        static unsigned long long* sbuf = nullptr;
        static long long* tbuf = nullptr; // per-CTA / per-chunk timeline of ONE launch in the middle of the sequence (CKB_GTRACE_AT)
        static int t_grid = 0;
        static int n_launch = 0;
        static std::string path;
        if(!sbuf)
        {
            cudaMalloc(&sbuf, MAXL * 4 * sizeof(unsigned long long));
            std::vector<unsigned long long> init(MAXL * 4);
            for(int i = 0; i < MAXL; ++i) { init[4 * i] = ~0ull; init[4 * i + 1] = 0; init[4 * i + 2] = ~0ull; init[4 * i + 3] = 0; }
            cudaMemcpy(sbuf, init.data(), init.size() * 8, cudaMemcpyHostToDevice);
            cudaMalloc(&tbuf, 8192 * sizeof(long long));
            cudaMemset(tbuf, 0, 8192 * sizeof(long long));
            path = sf;
            atexit([] {
                cudaDeviceSynchronize();
                std::vector<unsigned long long> h(MAXL * 4);
                cudaMemcpy(h.data(), sbuf, h.size() * 8, cudaMemcpyDeviceToHost);
                if(FILE* f = fopen(path.c_str(), "w"))
                {
                    for(int i = 0; i < n_launch && i < MAXL; ++i)
                        fprintf(f, "%d %llu %llu %llu %llu\n", i, h[4 * i], h[4 * i + 1], h[4 * i + 2], h[4 * i + 3]);
                    fclose(f);
                }
                std::vector<long long> hb(8192);
                cudaMemcpy(hb.data(), tbuf, hb.size() * 8, cudaMemcpyDeviceToHost);
                if(FILE* f = fopen((path + ".cta").c_str(), "w"))
                {
                    long long t0 = 0;
                    for(int i = 0; i < 8192; ++i) if(hb[i] && (!t0 || hb[i] < t0)) t0 = hb[i];
                    for(int b = 0; b < t_grid && b < 512; ++b)
                    {
                        fprintf(f, "%d", b);
                        for(int k = 0; k < 8; ++k) fprintf(f, " %lld", hb[b * 8 + k] ? hb[b * 8 + k] - t0 : -1);
                        fprintf(f, "\n");
                    }
                    for(int w = 0; w < 4; ++w)
                    {
                        fprintf(f, "# chunks cta%d warp%d:", (w >> 1) ? 74 : 0, (w & 1) ? 7 : 0);
                        for(int i = 0; i < 64; ++i)
                        {
                            const long long a0 = hb[4096 + (w * 64 + i) * 2], a1 = hb[4096 + (w * 64 + i) * 2 + 1];
                            if(a0) fprintf(f, " [%lld %lld]", a0 - t0, a1 - t0);
                        }
                        fprintf(f, "\n");
                    }
                    fclose(f);
                }
            });
        }
        if(n_launch < MAXL) a.gspan = sbuf + 4 * n_launch;
        if(getenv("CKB_GTRACE_AT") && n_launch == atoi(getenv("CKB_GTRACE_AT"))) { a.gtrace = tbuf; t_grid = plan.grid; }
        ++n_launch;
        return go();
    }
    if(const char* gf = getenv("CKB_GTRACE"))
    {
        static long long* gbuf = nullptr;
        const int NW = 8 * 1024;
        if(!gbuf) cudaMalloc(&gbuf, NW * sizeof(long long));
        cudaDeviceSynchronize();
        cudaMemset(gbuf, 0, NW * sizeof(long long));
        cudaDeviceSynchronize();
        a.gtrace = gbuf;
        const int rc = go();
        cudaDeviceSynchronize();
        std::vector<long long> hb(NW);
        cudaMemcpy(hb.data(), gbuf, NW * sizeof(long long), cudaMemcpyDeviceToHost);
        if(FILE* f = fopen(gf, "a"))
        {
            long long t0 = 0;
            for(int i = 0; i < NW; ++i) if(hb[i] && (!t0 || hb[i] < t0)) t0 = hb[i];
            fprintf(f, "# M%d N%d K%d grid %d\n", p.M, p.N, p.K, plan.grid);
            for(int b = 0; b < plan.grid && b < 512; ++b)
            {
                fprintf(f, "%d", b);
                for(int k = 0; k < 8; ++k) fprintf(f, " %lld", hb[b * 8 + k] ? hb[b * 8 + k] - t0 : -1);
                fprintf(f, "\n");
            }
            for(int w = 0; w < 4; ++w)
            {
                fprintf(f, "# chunks cta%d warp%d:", (w >> 1) ? 74 : 0, (w & 1) ? 7 : 0);
                for(int i = 0; i < 64; ++i)
                {
                    const long long a0 = hb[4096 + (w * 64 + i) * 2], a1 = hb[4096 + (w * 64 + i) * 2 + 1];
                    if(a0) fprintf(f, " [%lld %lld]", a0 - t0, a1 - t0);
                }
                fprintf(f, "\n");
            }
            fclose(f);
        }
        return rc;
    }
#endif
    return go();
}

} // namespace ckb
