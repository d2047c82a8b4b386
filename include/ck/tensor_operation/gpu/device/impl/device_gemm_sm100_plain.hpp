// device_gemm_sm100_plain.hpp -- sm_100a implementations behind the reference's single-A / single-B / no-D interfaces:
//   DeviceGemm        (device/device_gemm.hpp:21-38)          -> DeviceGemm_Sm100
//   DeviceGemmSplitK  (device/device_gemm_splitk.hpp:25-46)   -> DeviceGemmSplitK_Sm100   (KBatch is a hint, like KSplit of V2R1)
//   DeviceGemmStreamK (device/device_gemm_streamk.hpp:24-44)  -> DeviceGemmStreamK_Sm100  (NumSKBlocks is a hint)
//   DeviceGemmMultipleD_ABScale (device/device_gemm_multiple_d_ab_scale.hpp:22-59) -> DeviceGemmMultipleD_ABScale_W4_Sm100,
//     the [EXT] W4A16 group-scale op under the block-scale interface (ScaleBlockK = 128, B = ckb_prepack_b_int4 buffer).
// Data types served: int8 x int8 (-> int32 / int8 / bf16) and bf16|fp16 x int8 (B element op PassThrough), i.e. the dense and
// weight-only kernels of this path; everything funnels into ckb_gemm_query / ckb_gemm_run like the other operators.
#pragma once
#include "ck/tensor_operation/gpu/device/device_gemm.hpp"
#include "ck/tensor_operation/gpu/device/device_gemm_multiple_d_ab_scale.hpp"
#include "ck/tensor_operation/gpu/device/device_gemm_splitk.hpp"
#include "ck/tensor_operation/gpu/device/device_gemm_streamk.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_sm100.hpp"

namespace ck {
namespace tensor_operation {
namespace device {
namespace ckb_detail {
template <typename ALayout, typename BLayout, typename ADataType, typename BDataType, typename CDataType, typename CElementwiseOperation>
inline GemmArgument make_plain(const void* p_a, const void* p_b, void* p_c, index_t M, index_t N, index_t K, index_t StrideA,
                               index_t StrideB, index_t StrideC, const CElementwiseOperation& c_op, index_t k_hint)
{
    GemmArgument arg;
    ckb_gemm_problem& p = arg.problem;
    p.M = M; p.N = N; p.K = K;
    p.a_dtype = dtype_of<ADataType>(); p.a_layout = layout_of<ALayout>(); p.a = p_a; p.lda = StrideA;
    p.b_dtype = dtype_of<BDataType>(); p.b_layout = layout_of<BLayout>(); p.b = p_b; p.ldb = StrideB;
    p.b_op    = CKB_BOP_PASSTHROUGH;
    p.cde_op  = CElementwiseOperation::kCkbCde;
    p.cde_scale = cde_scale_of(c_op);
    p.act     = cde_act_of<CElementwiseOperation>();
    p.e_dtype = dtype_of<CDataType>(); p.e = p_c; p.lde = StrideC;
    p.kbatch  = k_hint;
    if(k_hint < 0) { arg.why = "negative split hint"; return arg; }
    if(p.b_layout == CKB_ROW && p.b_dtype == CKB_I8) arg.DeferRowMajorB();
    else arg.Query();
    return arg;
}
} // namespace ckb_detail

#define CKB_PLAIN_OP_BODY(NAME)                                                                                              \
    using Argument = ckb_detail::GemmArgument;                                                                               \
    using Invoker  = ckb_detail::GemmInvoker;                                                                                \
    static Invoker MakeInvoker() { return Invoker{}; }                                                                       \
    static bool IsSupportedArgument(const Argument& arg) { return arg.plan.supported != 0; }                                 \
    bool IsSupportedArgument(const BaseArgument* p) override { return IsSupportedArgument(ckb_detail::checked(p)); }         \
    std::unique_ptr<BaseInvoker> MakeInvokerPointer() override { return std::make_unique<Invoker>(); }                       \
    size_t GetWorkSpaceSize(const BaseArgument* p) const override { return ckb_detail::workspace_size(p); }                  \
    void SetWorkSpacePointer(BaseArgument* p, void* ws, const StreamConfig& sc = StreamConfig{}) const override              \
    {                                                                                                                        \
        ckb_detail::set_workspace(p, ws, sc);                                                                                \
    }                                                                                                                        \
    std::string GetTypeString() const override { return ckb_detail::type_string(NAME); }

template <typename ALayout, typename BLayout, typename CLayout, typename ADataType, typename BDataType, typename CDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation>
struct DeviceGemm_Sm100 : public DeviceGemm<ALayout, BLayout, CLayout, ADataType, BDataType, CDataType, AElementwiseOperation,
                                            BElementwiseOperation, CElementwiseOperation>
{
    CKB_PLAIN_OP_BODY("DeviceGemm_Sm100")
    static Argument MakeArgument(const void* p_a, const void* p_b, void* p_c, index_t M, index_t N, index_t K, index_t StrideA,
                                 index_t StrideB, index_t StrideC, AElementwiseOperation, BElementwiseOperation, CElementwiseOperation c)
    {
        return ckb_detail::make_plain<ALayout, BLayout, ADataType, BDataType, CDataType>(p_a, p_b, p_c, M, N, K, StrideA, StrideB, StrideC, c, 0);
    }
    std::unique_ptr<BaseArgument> MakeArgumentPointer(const void* p_a, const void* p_b, void* p_c, index_t M, index_t N, index_t K,
                                                      index_t StrideA, index_t StrideB, index_t StrideC, AElementwiseOperation a,
                                                      BElementwiseOperation b, CElementwiseOperation c) override
    {
        return std::make_unique<Argument>(MakeArgument(p_a, p_b, p_c, M, N, K, StrideA, StrideB, StrideC, a, b, c));
    }
};

template <typename ALayout, typename BLayout, typename CLayout, typename ADataType, typename BDataType, typename CDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation, typename ComputeType = CDataType>
struct DeviceGemmSplitK_Sm100 : public DeviceGemmSplitK<ALayout, BLayout, CLayout, ADataType, BDataType, CDataType, AElementwiseOperation,
                                                        BElementwiseOperation, CElementwiseOperation, ComputeType>
{
    CKB_PLAIN_OP_BODY("DeviceGemmSplitK_Sm100")
    static Argument MakeArgument(const void* p_a, const void* p_b, void* p_c, index_t M, index_t N, index_t K, index_t StrideA,
                                 index_t StrideB, index_t StrideC, AElementwiseOperation, BElementwiseOperation, CElementwiseOperation c,
                                 index_t KBatch)
    {
        return ckb_detail::make_plain<ALayout, BLayout, ADataType, BDataType, CDataType>(p_a, p_b, p_c, M, N, K, StrideA, StrideB, StrideC, c, KBatch);
    }
    std::unique_ptr<BaseArgument> MakeArgumentPointer(const void* p_a, const void* p_b, void* p_c, index_t M, index_t N, index_t K,
                                                      index_t StrideA, index_t StrideB, index_t StrideC, AElementwiseOperation a,
                                                      BElementwiseOperation b, CElementwiseOperation c, index_t KBatch) override
    {
        return std::make_unique<Argument>(MakeArgument(p_a, p_b, p_c, M, N, K, StrideA, StrideB, StrideC, a, b, c, KBatch));
    }
};

template <typename ALayout, typename BLayout, typename CLayout, typename ADataType, typename BDataType, typename CDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation>
struct DeviceGemmStreamK_Sm100 : public DeviceGemmStreamK<ALayout, BLayout, CLayout, ADataType, BDataType, CDataType, AElementwiseOperation,
                                                          BElementwiseOperation, CElementwiseOperation>
{
    CKB_PLAIN_OP_BODY("DeviceGemmStreamK_Sm100")
    static Argument MakeArgument(const void* p_a, const void* p_b, void* p_c, index_t M, index_t N, index_t K, index_t StrideA,
                                 index_t StrideB, index_t StrideC, AElementwiseOperation, BElementwiseOperation, CElementwiseOperation c,
                                 index_t NumSKBlocks = 0)
    {
        return ckb_detail::make_plain<ALayout, BLayout, ADataType, BDataType, CDataType>(p_a, p_b, p_c, M, N, K, StrideA, StrideB, StrideC, c, NumSKBlocks);
    }
    std::unique_ptr<BaseArgument> MakeArgumentPointer(const void* p_a, const void* p_b, void* p_c, index_t M, index_t N, index_t K,
                                                      index_t StrideA, index_t StrideB, index_t StrideC, AElementwiseOperation a,
                                                      BElementwiseOperation b, CElementwiseOperation c, index_t NumSKBlocks = 0) override
    {
        return std::make_unique<Argument>(MakeArgument(p_a, p_b, p_c, M, N, K, StrideA, StrideB, StrideC, a, b, c, NumSKBlocks));
    }
};

// [EXT] W4A16 under the block-scale interface: A unscaled (p_a_scale must be null, ScaleBlockM = 1), B = pre-packed int4 with one
// scale per (ScaleBlockK = 128 k, ScaleBlockN = 1 channel), zero point 8.  BElementwiseOperation = DequantizeGroup | DequantizeGroupRound.
template <typename ALayout, typename DsLayout, typename ELayout, typename ADataType, typename DsDataType, typename EDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CDEElementwiseOperation>
struct DeviceGemmMultipleD_ABScale_W4_Sm100
    : public DeviceGemmMultipleD_ABScale<ALayout, tensor_layout::gemm::ColumnMajor, DsLayout, ELayout, ADataType, float, uint8_t, ADataType,
                                         DsDataType, EDataType, 1, 1, 128, AElementwiseOperation, BElementwiseOperation, CDEElementwiseOperation>
{
    CKB_PLAIN_OP_BODY("DeviceGemmMultipleD_ABScale_W4_Sm100")
    static constexpr index_t NumDTensor = DsDataType::Size();
    static Argument MakeArgument(const void* p_a, const void* p_b_prepacked, std::array<const void*, NumDTensor> p_ds, void* p_e, index_t M,
                                 index_t N, index_t K, index_t StrideA, index_t, std::array<index_t, NumDTensor> StrideDs, index_t StrideE,
                                 const void* p_a_scale, const void* p_b_scale, AElementwiseOperation, BElementwiseOperation,
                                 CDEElementwiseOperation cde)
    {
        static_assert(NumDTensor <= CKB_MAX_D, "at most two D tensors");
        Argument arg;
        ckb_gemm_problem& p = arg.problem;
        p.M = M; p.N = N; p.K = K;
        p.a_dtype = ckb_detail::dtype_of<ADataType>(); p.a_layout = ckb_detail::layout_of<ALayout>(); p.a = p_a; p.lda = StrideA;
        p.b_dtype = CKB_I4; p.b_layout = CKB_PREPACKED; p.b = p_b_prepacked;
        p.b_op = BElementwiseOperation::kCkbBOp; p.s_dtype = p.a_dtype; p.scale = p_b_scale; p.zero_const = 8; p.group_size = 128;
        p.num_d = NumDTensor;
        ckb_detail::fill_d<DsLayout, DsDataType, 0, size_t(NumDTensor)>(p, p_ds, StrideDs);
        p.cde_op = CDEElementwiseOperation::kCkbCde;
        p.cde_scale = ckb_detail::cde_scale_of(cde);
        p.e_dtype = ckb_detail::dtype_of<EDataType>(); p.e = p_e; p.lde = StrideE;
        if(p_a_scale) arg.why = "A is not block-scaled on this path (p_a_scale must be null)";
        else arg.Query();
        return arg;
    }
    std::unique_ptr<BaseArgument> MakeArgumentPointer(const void* p_a, const void* p_b, std::array<const void*, NumDTensor> p_ds, void* p_e,
                                                      const index_t M, const index_t N, const index_t K, const index_t StrideA,
                                                      const index_t StrideB, const std::array<index_t, NumDTensor> StrideDs,
                                                      const index_t StrideE, const void* p_a_scale, const void* p_b_scale,
                                                      AElementwiseOperation a, BElementwiseOperation b, CDEElementwiseOperation c) override
    {
        return std::make_unique<Argument>(MakeArgument(p_a, p_b, p_ds, p_e, M, N, K, StrideA, StrideB, StrideDs, StrideE, p_a_scale, p_b_scale, a, b, c));
    }
};
#undef CKB_PLAIN_OP_BODY

} // namespace device
} // namespace tensor_operation
} // namespace ck
