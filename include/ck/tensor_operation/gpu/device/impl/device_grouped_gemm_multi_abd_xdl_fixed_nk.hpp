// DeviceGroupedGemm_Xdl_Multi_ABD_Fixed_NK, B200 build -- same name and host behaviour as the reference class
// (device/impl/device_grouped_gemm_multi_abd_xdl_fixed_nk.hpp:148-850):
//   * the Argument constructor takes (ignored) pointer vectors and the gemm_descs; every desc must carry the same
//     M_ (= sum of all groups' M), N_ and K_, else it throws "wrong! M/N/K is not identical" (:460-463);
//   * the per-group pointers, M and strides come from a DEVICE array the caller fills and hands over with
//     SetDeviceKernelArgs (GetDeviceKernelArgSize bytes); Run without it throws (:597-600);
//   * SetKBatch is accepted and ignored (:429); SetElementwiseOps stores the functors.
// Below the class everything is ckb_grouped_gemm_query / ckb_grouped_gemm_run (include/ckb.h).
#pragma once
#include "ck/tensor_operation/gpu/device/device_grouped_gemm_multi_abd_fixed_nk.hpp"
#include "ck/tensor_operation/gpu/device/gemm_specialization.hpp"
#include "ck/tensor_operation/gpu/device/impl/ckb_gemm_impl.hpp"

namespace ck {
namespace tensor_operation {
namespace device {

template <typename AsLayout, typename BsLayout, typename DsLayout, typename ELayout, typename AsDataType, typename BsDataType,
          typename AccDataType, typename CShuffleDataType, typename DsDataType, typename EDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CDEElementwiseOperation,
          GemmSpecialization GemmSpec = GemmSpecialization::Default, typename... Unused>
struct DeviceGroupedGemm_Xdl_Multi_ABD_Fixed_NK
    : public DeviceGroupedGemmMultiABDFixedNK<AsLayout, BsLayout, DsLayout, ELayout, AsDataType, BsDataType, DsDataType, EDataType,
                                              AElementwiseOperation, BElementwiseOperation, CDEElementwiseOperation>
{
    using DeviceOp = DeviceGroupedGemm_Xdl_Multi_ABD_Fixed_NK;
    static constexpr index_t NumATensor = AsDataType::Size();
    static constexpr index_t NumBTensor = BsDataType::Size();
    static constexpr index_t NumDTensor = DsDataType::Size();
    using KernelArgument = GroupedGemmMultiABDKernelArgument<NumATensor, NumBTensor, NumDTensor>;

    struct Argument : public BaseArgument
    {
        Argument(std::vector<std::array<const void*, NumATensor>>&, std::vector<std::array<const void*, NumBTensor>>&,
                 std::vector<std::array<const void*, NumDTensor>>&, std::vector<void*>&, std::vector<GemmMultiABDDesc>& gemm_descs,
                 AElementwiseOperation a_op = AElementwiseOperation{}, BElementwiseOperation b_op = BElementwiseOperation{},
                 CDEElementwiseOperation c_op = CDEElementwiseOperation{})
            : a_element_op_{a_op}, b_element_op_{b_op}, c_element_op_{c_op}
        {
            static_assert(NumATensor == 1 && NumBTensor >= 1 && NumBTensor <= 2 && NumDTensor <= 1, "A0, B0[,B1], at most one D");
            group_count_ = ck::index_t(gemm_descs.size());
            if(group_count_ == 0) { why_ = "no groups"; return; }
            sum_of_m = gemm_descs[0].M_;
            for(const auto& d : gemm_descs)
                if(sum_of_m != d.M_ || gemm_descs[0].N_ != d.N_ || gemm_descs[0].K_ != d.K_)
                    throw std::runtime_error("wrong! M/N/K is not identical");
            const GemmMultiABDDesc& d0 = gemm_descs[0];
            ckb_gemm_problem& p        = proto_;
            p.M = sum_of_m; p.N = d0.N_; p.K = d0.K_;
            p.a_dtype  = ckb_detail::dtype_of<tuple_element_t<0, AsDataType>>();
            p.a_layout = ckb_detail::layout_of<tuple_element_t<0, AsLayout>>();
            p.b_dtype  = ckb_detail::dtype_of<tuple_element_t<0, BsDataType>>();
            p.b_layout = ckb_detail::layout_of<tuple_element_t<0, BsLayout>>();
            if constexpr(NumBTensor == 2)
            {
                p.b_op    = BElementwiseOperation::kCkbBOp;
                p.s_dtype = ckb_detail::dtype_of<tuple_element_t<1, BsDataType>>();
            }
            else
                p.b_op = CKB_BOP_PASSTHROUGH;
            p.num_d = NumDTensor;
            if constexpr(NumDTensor == 1) p.d_dtype[0] = ckb_detail::dtype_of<tuple_element_t<0, DsDataType>>();
            p.cde_op  = CDEElementwiseOperation::kCkbCde;
            p.e_dtype = ckb_detail::dtype_of<EDataType>();
            // every group must satisfy the same stride rules (the strides the kernel uses are the device-side ones;
            // the descs are the host's statement of them, exactly as in the reference)
            for(const auto& d : gemm_descs)
            {
                if(d.stride_As_.size() != size_t(NumATensor) || d.stride_Bs_.size() != size_t(NumBTensor) ||
                   d.stride_Ds_.size() != size_t(NumDTensor))
                { why_ = "gemm_desc stride vectors do not match the tensor counts"; return; }
                if(NumBTensor == 2 && d.stride_Bs_[1] != 0) { why_ = "B1 must be a per-channel vector (StrideB1 == 0)"; return; }
                p.lda = d.stride_As_[0]; p.ldb = d.stride_Bs_[0]; p.lde = d.stride_C_;
                if(ckb_grouped_gemm_query(&p, group_count_) != CKB_OK) { why_ = ckb_last_error(); return; }
            }
            supported_ = true;
        }
        void UpdateKBatch(index_t) {}

        ckb_gemm_problem proto_{};
        index_t group_count_ = 0, sum_of_m = 0;
        bool supported_ = false;
        std::string why_;
        const void* grouped_gemm_kernel_args_dev = nullptr;
        AElementwiseOperation a_element_op_;
        BElementwiseOperation b_element_op_;
        CDEElementwiseOperation c_element_op_;
    };

    struct Invoker : public BaseInvoker
    {
        float Run(const Argument& arg, const StreamConfig& sc = StreamConfig{})
        {
            if(!arg.supported_) throw std::runtime_error("wrong! the device op does not support this grouped GEMM: " + arg.why_);
            if(arg.grouped_gemm_kernel_args_dev == nullptr) throw std::runtime_error("wrong! grouped_gemm_kernel_args_dev is nullpr");
            auto launch = [&]() {
                if(ckb_grouped_gemm_run(&arg.proto_, arg.grouped_gemm_kernel_args_dev, arg.group_count_, arg.sum_of_m, sc.stream_id_) != CKB_OK)
                    throw std::runtime_error(std::string("ckb_grouped_gemm_run: ") + ckb_last_error());
            };
            if(!sc.time_kernel_) { launch(); return 0.f; }
            for(int i = 0; i < sc.cold_niters_; ++i) launch();
            cudaEvent_t start, stop;
            ckb_detail::cuda_check(cudaEventCreate(&start), "cudaEventCreate");
            ckb_detail::cuda_check(cudaEventCreate(&stop), "cudaEventCreate");
            ckb_detail::cuda_check(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
            ckb_detail::cuda_check(cudaEventRecord(start, sc.stream_id_), "cudaEventRecord");
            for(int i = 0; i < sc.nrepeat_; ++i) launch();
            ckb_detail::cuda_check(cudaEventRecord(stop, sc.stream_id_), "cudaEventRecord");
            ckb_detail::cuda_check(cudaEventSynchronize(stop), "cudaEventSynchronize");
            float ms = 0.f;
            ckb_detail::cuda_check(cudaEventElapsedTime(&ms, start, stop), "cudaEventElapsedTime");
            cudaEventDestroy(start);
            cudaEventDestroy(stop);
            return ms / float(sc.nrepeat_);
        }
        float Run(const BaseArgument* p_arg, const StreamConfig& sc = StreamConfig{}) override
        {
            return Run(*dynamic_cast<const Argument*>(p_arg), sc);
        }
    };

    static bool IsSupportedArgument(const Argument& arg) { return arg.supported_; }
    bool IsSupportedArgument(const BaseArgument* p_arg) override { return IsSupportedArgument(*dynamic_cast<const Argument*>(p_arg)); }

    static auto MakeArgument(std::vector<std::array<const void*, NumATensor>>& p_As, std::vector<std::array<const void*, NumBTensor>>& p_Bs,
                             std::vector<std::array<const void*, NumDTensor>>& p_Ds, std::vector<void*>& p_Es,
                             std::vector<GemmMultiABDDesc> gemm_descs, AElementwiseOperation a = AElementwiseOperation{},
                             BElementwiseOperation b = BElementwiseOperation{}, CDEElementwiseOperation c = CDEElementwiseOperation{})
    {
        return Argument{p_As, p_Bs, p_Ds, p_Es, gemm_descs, a, b, c};
    }
    static auto MakeInvoker() { return Invoker{}; }

    std::unique_ptr<BaseArgument>
    MakeArgumentPointer(std::vector<std::array<const void*, NumATensor>>& p_As, std::vector<std::array<const void*, NumBTensor>>& p_Bs,
                        std::vector<std::array<const void*, NumDTensor>>& p_Ds, std::vector<void*>& p_Es,
                        std::vector<GemmMultiABDDesc>& gemm_descs, AElementwiseOperation a = AElementwiseOperation{},
                        BElementwiseOperation b = BElementwiseOperation{}, CDEElementwiseOperation c = CDEElementwiseOperation{}) override
    {
        return std::make_unique<Argument>(p_As, p_Bs, p_Ds, p_Es, gemm_descs, a, b, c);
    }
    std::unique_ptr<BaseInvoker> MakeInvokerPointer() override { return std::make_unique<Invoker>(); }
    std::string GetTypeString() const override { return "DeviceGroupedGemm_Xdl_Multi_ABD_Fixed_NK<sm_100a, grouped_hmma_64x64x128_8w_persistent>"; }

    static void SetElementwiseOps(Argument& arg, AElementwiseOperation a, BElementwiseOperation b, CDEElementwiseOperation c)
    {
        arg.a_element_op_ = a; arg.b_element_op_ = b; arg.c_element_op_ = c;
    }
    static void SetDeviceKernelArgs(Argument& arg, const void* kernel_args) { arg.grouped_gemm_kernel_args_dev = kernel_args; }
    void SetDeviceKernelArgs(BaseArgument* p_arg, const void* kernel_args) const override
    {
        SetDeviceKernelArgs(*dynamic_cast<Argument*>(p_arg), kernel_args);
    }
    void SetElementwiseOps(BaseArgument* p_arg, AElementwiseOperation a, BElementwiseOperation b, CDEElementwiseOperation c) const override
    {
        SetElementwiseOps(*dynamic_cast<Argument*>(p_arg), a, b, c);
    }
    size_t GetDeviceKernelArgSize(const BaseArgument* p_arg) const override
    {
        static_assert(sizeof(KernelArgument) % 8 == 0, "argument struct must keep pointer alignment in an array");
        return size_t(dynamic_cast<const Argument*>(p_arg)->group_count_) * sizeof(KernelArgument);
    }
    static void SetKBatch(Argument& arg, index_t k_batch) { arg.UpdateKBatch(k_batch); }
    void SetKBatch(BaseArgument* p_arg, index_t k_batch) const override { SetKBatch(*dynamic_cast<Argument*>(p_arg), k_batch); }
};

} // namespace device
} // namespace tensor_operation
} // namespace ck
