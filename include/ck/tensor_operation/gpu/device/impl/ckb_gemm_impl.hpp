// ckb_gemm_impl.hpp -- the one place where the CK-style device operators cross into the C-ABI (include/ckb.h).
// In the reference each DeviceGemm* class builds grid descriptors in its Argument constructor, validates them
// in IsSupportedArgument and launches a HIP kernel through launch_and_time_kernel
// (include/ck/host_utility/kernel_launch.hpp:13-84).  Here the Argument constructor calls ckb_gemm_query
// (validation + kernel choice + TMA descriptors), IsSupportedArgument reads plan.supported, and Invoker::Run
// calls ckb_gemm_run, timing with CUDA events exactly as the reference does (cold_niters_ warm-ups, nrepeat_
// timed launches, mean).  Error conventions follow device_fpAintB_gemm_wmma.hpp:426-433: running an
// unsupported argument throws std::runtime_error.
#pragma once
#include <cuda_runtime_api.h>

#include <algorithm>
#include <cstdio>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>

#include "ck/host_utility/flush_cache.hpp"
#include "ck/host_utility/kernel_launch.hpp"
#include "ck/tensor_operation/gpu/device/device_base.hpp"
#include "ck/tensor_operation/gpu/device/tensor_layout.hpp"
#include "ck/tensor_operation/gpu/element/element_wise_operation.hpp"
#include "ck/utility/data_type.hpp"
#include "ckb.h"

namespace ck {
namespace tensor_operation {
namespace device {
namespace ckb_detail {

template <typename T>
constexpr int dtype_of()
{
    if constexpr(std::is_same_v<T, float>) return CKB_F32;
    else if constexpr(std::is_same_v<T, half_t>) return CKB_F16;
    else if constexpr(std::is_same_v<T, bhalf_t>) return CKB_BF16;
    else if constexpr(std::is_same_v<T, int8_t>) return CKB_I8;
    else if constexpr(std::is_same_v<T, uint8_t>) return CKB_U8;
    else if constexpr(std::is_same_v<T, int32_t>) return CKB_I32;
    else return -1;
}
template <typename L>
constexpr int layout_of()
{
    return std::is_same_v<L, tensor_layout::gemm::RowMajor> ? CKB_ROW : CKB_COL;
}
// D(m,n) strides from a CK layout + StrideD: Row -> {s,1}, Col -> {1,s}
template <typename L>
inline void d_strides(index_t stride, int64_t& sm, int64_t& sn)
{
    if constexpr(std::is_same_v<L, tensor_layout::gemm::RowMajor>) { sm = stride; sn = 1; }
    else { sm = 1; sn = stride; }
}

inline void cuda_check(cudaError_t e, const char* what)
{
    if(e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}

// name of the kernel chosen by the most recent ckb_gemm_query of this thread (for GetTypeString)
inline std::string& last_kernel_name()
{
    static thread_local std::string s;
    return s;
}

// Common Argument: owns the POD problem and the plan.
struct GemmArgument : public BaseArgument
{
    ckb_gemm_problem problem{};
    ckb_gemm_plan plan{};
    std::string why; // reason when unsupported
    bool queried = false;
    // The reference ops on this path need no caller workspace (GetWorkSpaceSize() == 0 there); the stream-K
    // kernels do, so when the caller has not supplied one the argument owns a small zeroed buffer.
    std::shared_ptr<void> own_workspace;
    // Row-major int8 B (CK's mk_kn_mn): the kernels stream K-major weights, so the op needs N*K bytes for a K-major copy.
    // MakeArgument has no side effects (as in the reference): it only records the request and validates the problem.  The
    // bytes are reported by GetWorkSpaceSize and taken through SetWorkSpacePointer (device_base.hpp:60-68), which enqueues
    // the transpose on the StreamConfig's stream -- a snapshot of the weight at that moment: call SetWorkSpacePointer again
    // after changing B.  A client that never supplies a workspace (the reference's client_example/30 does not) still runs:
    // the first Run allocates the copy itself and transposes on the run stream (not graph-capturable; documented fallback).
    // Column-major A (CK's km_*): activations change every call, so every Run transposes A on the launch stream into workspace
    // bytes reported by GetWorkSpaceSize (same fallback as above when the client supplies none); the GEMM then reads row-major A.
    const void* colmajor_a = nullptr;
    int64_t colmajor_lda   = 0;
    void* a_slot           = nullptr;
    std::shared_ptr<void> own_a;
    size_t ElemBytesA() const { return (problem.a_dtype == CKB_I8 || problem.a_dtype == CKB_U8) ? 1 : 2; }
    size_t TransposeABytes() const { return colmajor_a ? Align256(size_t(problem.M) * problem.K * ElemBytesA()) : 0; }
    void DeferColMajorA()
    {
        colmajor_a   = problem.a;
        colmajor_lda = problem.lda;
        problem.lda = problem.K; problem.a_layout = CKB_ROW;   // validated / planned as the row-major problem it becomes
    }
    void BindASlot(void* slot)
    {
        a_slot    = slot;
        problem.a = slot;
        if(b_ready) Query();
    }
    void TransposeA(cudaStream_t stream) const
    {
        if(ckb_transpose_a_colmajor(colmajor_a, colmajor_lda, problem.M, problem.K, int(ElemBytesA()), a_slot, stream) != CKB_OK)
            throw std::runtime_error(std::string("ckb_transpose_a_colmajor: ") + ckb_last_error());
    }
    const void* rowmajor_b = nullptr;
    int64_t rowmajor_ld    = 0;
    bool b_ready           = true;
    std::shared_ptr<void> own_b;
    static size_t Align256(size_t v) { return (v + 255) / 256 * 256; }
    size_t RepackBytes() const { return rowmajor_b ? Align256(size_t(problem.N) * problem.K) : 0; }
    void DeferRowMajorB()
    {
        rowmajor_b  = problem.b;
        rowmajor_ld = problem.ldb;
        b_ready     = false;
        // validate / plan as the K-major problem it will become (the pointer is replaced once the copy exists)
        problem.ldb = problem.K; problem.b_layout = CKB_COL;
        if(rowmajor_ld < problem.N) { why = "StrideB < N"; return; }
        Query();
    }
    void PrepareB(void* dst, cudaStream_t stream)
    {
        if(ckb_prepack_b_rowmajor_i8(rowmajor_b, rowmajor_ld, problem.K, problem.N, dst, stream) != CKB_OK)
            throw std::runtime_error(std::string("ckb_prepack_b_rowmajor_i8: ") + ckb_last_error());
        problem.b = dst;
        Query(); // tensor maps for the final pointer
        b_ready = true;
    }
    void* Workspace()
    {
        if(p_workspace_ || plan.workspace_bytes == 0) return p_workspace_;
        if(!own_workspace)
        {
            void* p = nullptr;
            cuda_check(cudaMalloc(&p, plan.workspace_bytes), "cudaMalloc(workspace)");
            cuda_check(cudaMemset(p, 0, plan.workspace_bytes), "cudaMemset(workspace)");
            own_workspace = std::shared_ptr<void>(p, [](void* q) { cudaFree(q); });
        }
        return own_workspace.get();
    }

    void Query()
    {
        const int rc = ckb_gemm_query(&problem, &plan);
        queried      = true;
        if(rc != CKB_OK || !plan.supported) why = ckb_last_error();
        if(rc != CKB_OK) plan.supported = 0;
        if(plan.supported) last_kernel_name() = plan.kernel_name;
    }
    // bytes of the A and B buffers (for the rotating copies of flush_cache.hpp)
    size_t SizeA() const
    {
        const size_t eb = (problem.a_dtype == CKB_I8 || problem.a_dtype == CKB_U8) ? 1 : 2;
        return size_t(problem.M) * size_t(problem.lda) * eb;
    }
    size_t SizeB() const
    {
        if(problem.b_layout == CKB_PREPACKED) return size_t(ckb_prepack_b_int4_bytes(problem.N, problem.K));
        return size_t(problem.N) * size_t(problem.ldb);
    }
    void Print() const
    {
        std::printf("ckb problem: M %d N %d K %d lda %lld ldb %lld lde %lld kernel %s\n", problem.M, problem.N, problem.K,
                    (long long)problem.lda, (long long)problem.ldb, (long long)problem.lde, plan.kernel_name);
    }
};

struct GemmInvoker : public BaseInvoker
{
    float Run(const GemmArgument& arg, const StreamConfig& sc = StreamConfig{})
    {
        if(!arg.plan.supported)
            throw std::runtime_error("wrong! the device op does not support this GEMM problem: " + arg.why);
        if(sc.log_level_ > 0 || EnvIsEnabled("CK_LOGGING")) arg.Print();
        if(!arg.b_ready)
        {
            auto& a = const_cast<GemmArgument&>(arg);
            void* t = nullptr;
            cuda_check(cudaMalloc(&t, a.RepackBytes()), "cudaMalloc(K-major copy of B; supply a workspace to avoid this)");
            a.own_b = std::shared_ptr<void>(t, [](void* q) { cudaFree(q); });
            a.PrepareB(t, sc.stream_id_);
        }
        if(arg.colmajor_a && !arg.a_slot)
        {
            auto& a = const_cast<GemmArgument&>(arg);
            void* t = nullptr;
            cuda_check(cudaMalloc(&t, a.TransposeABytes()), "cudaMalloc(row-major copy of A; supply a workspace to avoid this)");
            a.own_a = std::shared_ptr<void>(t, [](void* q) { cudaFree(q); });
            a.BindASlot(t);
        }
        auto launch = [&]() {
            if(arg.colmajor_a) arg.TransposeA(sc.stream_id_);
            const int rc = ckb_gemm_run(&arg.problem, &arg.plan, const_cast<GemmArgument&>(arg).Workspace(), sc.stream_id_);
            if(rc != CKB_OK) throw std::runtime_error(std::string("ckb_gemm_run: ") + ckb_last_error());
        };
        if(sc.flush_cache)
        {
            // cold-cache timing (flush_cache.hpp): rotating copies of A and B, L2 evicted before every timed launch, median
            auto& a = const_cast<GemmArgument&>(arg);
            utility::RotatingMem rot(a.problem.a, a.problem.b, a.SizeA(), a.SizeB(), std::size_t(std::max(1, sc.rotating_count)),
                                     [&](const void* pa, const void* pb) { a.problem.a = pa; a.problem.b = pb; a.Query(); });
            if(sc.log_level_ > 0 || EnvIsEnabled("CK_LOGGING")) rot.Print();
            auto pre = [&]() { rot.Next(); utility::flush_cache(sc.stream_id_); };
            return utility::launch_and_time_kernel_with_preprocess<false>(sc, pre, launch);
        }
        // launch_and_time_kernel (kernel_launch.hpp:37-69): warm-ups, then nrepeat launches between two events, mean
        return launch_and_time_kernel(sc, launch);
    }
    float Run(const BaseArgument* p_arg, const StreamConfig& sc = StreamConfig{}) override
    {
        const auto* a = dynamic_cast<const GemmArgument*>(p_arg);
        if(!a) throw std::runtime_error("wrong! the argument was not made by this device operation");
        return Run(*a, sc);
    }
};

// Workspace handling shared by every op: GetWorkSpaceSize = plan.workspace_bytes; SetWorkSpacePointer zeroes
// it once (the kernels keep the stream-K workspace zero between launches).
inline size_t workspace_size(const BaseArgument* p)
{
    const auto* a = dynamic_cast<const GemmArgument*>(p);
    if(!a) return 0;
    const size_t kbytes = size_t(a->plan.workspace_bytes); // stream-K partial sums (decode-shaped int8-weight problems)
    if(!a->rowmajor_b && !a->colmajor_a) return kbytes;
    return GemmArgument::Align256(kbytes) + a->RepackBytes() + a->TransposeABytes();
}
inline void set_workspace(BaseArgument* p_arg, void* p_workspace, const StreamConfig& sc)
{
    p_arg->p_workspace_ = p_workspace;
    auto* a = dynamic_cast<GemmArgument*>(p_arg);
    if(!a || !p_workspace) return;
    const size_t kbytes = size_t(a->plan.workspace_bytes);
    if(kbytes) cuda_check(cudaMemsetAsync(p_workspace, 0, kbytes, sc.stream_id_), "cudaMemsetAsync(workspace)");
    char* extra = static_cast<char*>(p_workspace) + GemmArgument::Align256(kbytes);
    if(a->rowmajor_b) a->PrepareB(extra, sc.stream_id_);
    if(a->colmajor_a) a->BindASlot(extra + a->RepackBytes());
}

template <typename Op, typename = void>
struct has_scale : std::false_type {};
template <typename Op>
struct has_scale<Op, std::void_t<decltype(std::declval<const Op&>().ckb_scale())>> : std::true_type {};
template <typename Op, typename = void>
struct has_act : std::false_type {};
template <typename Op>
struct has_act<Op, std::void_t<decltype(Op::kCkbAct)>> : std::true_type {};

template <typename Op>
inline float cde_scale_of(const Op& op)
{
    if constexpr(has_scale<Op>::value) return op.ckb_scale();
    else return 1.f;
}
template <typename Op>
constexpr int cde_act_of()
{
    if constexpr(has_act<Op>::value) return Op::kCkbAct;
    else return CKB_ACT_PASSTHROUGH;
}

} // namespace ckb_detail
} // namespace device
} // namespace tensor_operation
} // namespace ck
