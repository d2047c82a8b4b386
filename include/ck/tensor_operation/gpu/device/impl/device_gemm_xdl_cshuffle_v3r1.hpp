// DeviceGemm_Xdl_CShuffleV3R1, B200 build -- the split-K + reduce universal GEMM of
// device/impl/device_gemm_xdl_cshuffle_v3r1.hpp:73-703, served for its bf16 A x int8 B instances
// (library/.../gemm_universal_reduce.hpp:156-240, example/35_splitK_gemm/gemm_xdl_splitk_reduce_bf16A_i8B.cpp).
// The reference runs KSplit partial GEMMs into an fp32 workspace of M*N*KBatch*4 bytes (:694) and then a
// DeviceReduceThreadWiseMultiD pass that sums them and applies the Ds.  Here KSplit is a HINT (ckb_gemm_problem.kbatch):
// decode-sized problems run the stream-K kernel (fp32 red.add workspace of one tile set + last-CTA fix-up, no second kernel),
// larger ones the dense tcgen05 path with no split at all; GetWorkSpaceSize reports what the chosen kernel needs.
// Results are independent of KSplit up to fp32 summation order, as in the reference.
#pragma once
#include "ck/tensor_operation/gpu/device/device_gemm_v2.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_sm100.hpp"

namespace ck {
namespace tensor_operation {
namespace device {

template <typename ALayout, typename BLayout, typename DsLayout, typename CLayout, typename ADataType, typename BDataType,
          typename DsDataType, typename CDataType, typename GemmAccDataType, typename CShuffleDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation,
          GemmSpecialization GemmSpec = GemmSpecialization::Default, typename... Unused>
struct DeviceGemm_Xdl_CShuffleV3R1
    : public DeviceGemmV2R1<ALayout, BLayout, DsLayout, CLayout, ADataType, BDataType, DsDataType, CDataType, AElementwiseOperation,
                            BElementwiseOperation, CElementwiseOperation>
{
    using Argument = ckb_detail::GemmArgument;
    using Invoker  = ckb_detail::GemmInvoker;
    static constexpr index_t NumDTensor = DsDataType::Size();

    static Argument MakeArgument(const void* p_a, const void* p_b, std::array<const void*, NumDTensor> p_ds, void* p_c, index_t M,
                                 index_t N, index_t K, index_t StrideA, index_t StrideB, std::array<index_t, NumDTensor> StrideDs,
                                 index_t StrideC, index_t KSplit, AElementwiseOperation, BElementwiseOperation, CElementwiseOperation cde)
    {
        static_assert(NumDTensor <= CKB_MAX_D, "at most two D tensors");
        Argument arg;
        ckb_gemm_problem& p = arg.problem;
        p.M = M; p.N = N; p.K = K;
        p.a_dtype = ckb_detail::dtype_of<ADataType>(); p.a_layout = ckb_detail::layout_of<ALayout>(); p.a = p_a; p.lda = StrideA;
        p.b_dtype = ckb_detail::dtype_of<BDataType>(); p.b_layout = ckb_detail::layout_of<BLayout>(); p.b = p_b; p.ldb = StrideB;
        p.b_op    = CKB_BOP_PASSTHROUGH; // int8 -> bf16 exactly (PassThrough<bhalf_t,int8_t>)
        p.num_d   = NumDTensor;
        ckb_detail::fill_d<DsLayout, DsDataType, 0, size_t(NumDTensor)>(p, p_ds, StrideDs);
        p.cde_op = CElementwiseOperation::kCkbCde;
        p.cde_scale = ckb_detail::cde_scale_of(cde);
        p.e_dtype = ckb_detail::dtype_of<CDataType>(); p.e = p_c; p.lde = StrideC;
        p.kbatch  = KSplit;
        if(KSplit < 1) arg.why = "KSplit must be >= 1";
        if(p.b_layout == CKB_ROW && p.b_dtype == CKB_I8 && arg.why.empty())
            arg.DeferRowMajorB(); // mk_kn_mn: no allocation and no launch here (see GemmArgument::DeferRowMajorB)
        if(arg.why.empty()) arg.Query();
        return arg;
    }
    static Invoker MakeInvoker() { return Invoker{}; }
    static bool IsSupportedArgument(const Argument& arg) { return arg.plan.supported != 0; }
    bool IsSupportedArgument(const BaseArgument* p) override { return IsSupportedArgument(ckb_detail::checked(p)); }
    std::unique_ptr<BaseArgument>
    MakeArgumentPointer(const void* p_a, const void* p_b, std::array<const void*, NumDTensor> p_ds, void* p_c, index_t M, index_t N, index_t K,
                        index_t StrideA, index_t StrideB, std::array<index_t, NumDTensor> StrideDs, index_t StrideC, index_t KSplit,
                        AElementwiseOperation a, BElementwiseOperation b, CElementwiseOperation c) override
    {
        return std::make_unique<Argument>(MakeArgument(p_a, p_b, p_ds, p_c, M, N, K, StrideA, StrideB, StrideDs, StrideC, KSplit, a, b, c));
    }
    std::unique_ptr<BaseInvoker> MakeInvokerPointer() override { return std::make_unique<Invoker>(); }
    size_t GetWorkSpaceSize(const BaseArgument* p) const override { return ckb_detail::workspace_size(p); }
    void SetWorkSpacePointer(BaseArgument* p, void* ws, const StreamConfig& sc = StreamConfig{}) const override
    {
        ckb_detail::set_workspace(p, ws, sc);
    }
    std::string GetTypeString() const override { return ckb_detail::type_string("DeviceGemm_Xdl_CShuffleV3R1"); }
};

} // namespace device
} // namespace tensor_operation
} // namespace ck
