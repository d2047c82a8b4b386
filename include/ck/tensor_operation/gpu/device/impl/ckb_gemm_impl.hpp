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

#include <cstdio>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>

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
    std::shared_ptr<void> own_b; // K-major copy of a row-major weight (see DeviceGemmMultipleABD_Xdl_CShuffle)
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
        auto launch = [&]() {
            const int rc = ckb_gemm_run(&arg.problem, &arg.plan, const_cast<GemmArgument&>(arg).Workspace(), sc.stream_id_);
            if(rc != CKB_OK) throw std::runtime_error(std::string("ckb_gemm_run: ") + ckb_last_error());
        };
        if(!sc.time_kernel_)
        {
            launch();
            return 0.f;
        }
        // launch_and_time_kernel (kernel_launch.hpp:37-69): warm-ups, then nrepeat launches between two events
        for(int i = 0; i < sc.cold_niters_; ++i) launch();
        cudaEvent_t start, stop;
        cuda_check(cudaEventCreate(&start), "cudaEventCreate");
        cuda_check(cudaEventCreate(&stop), "cudaEventCreate");
        cuda_check(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
        cuda_check(cudaEventRecord(start, sc.stream_id_), "cudaEventRecord");
        for(int i = 0; i < sc.nrepeat_; ++i) launch();
        cuda_check(cudaEventRecord(stop, sc.stream_id_), "cudaEventRecord");
        cuda_check(cudaEventSynchronize(stop), "cudaEventSynchronize");
        float ms = 0.f;
        cuda_check(cudaEventElapsedTime(&ms, start, stop), "cudaEventElapsedTime");
        cudaEventDestroy(start);
        cudaEventDestroy(stop);
        return ms / float(sc.nrepeat_);
    }
    float Run(const BaseArgument* p_arg, const StreamConfig& sc = StreamConfig{}) override
    {
        return Run(*dynamic_cast<const GemmArgument*>(p_arg), sc);
    }
};

// Workspace handling shared by every op: GetWorkSpaceSize = plan.workspace_bytes; SetWorkSpacePointer zeroes
// it once (the kernels keep the stream-K workspace zero between launches).
inline size_t workspace_size(const BaseArgument* p) { return size_t(dynamic_cast<const GemmArgument*>(p)->plan.workspace_bytes); }
inline void set_workspace(BaseArgument* p_arg, void* p_workspace, const StreamConfig& sc)
{
    p_arg->p_workspace_ = p_workspace;
    const size_t bytes  = workspace_size(p_arg);
    if(p_workspace && bytes) cuda_check(cudaMemsetAsync(p_workspace, 0, bytes, sc.stream_id_), "cudaMemsetAsync(workspace)");
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
