// device_gemm_sm100.hpp -- concrete device operators of the mixed-precision GEMM path, B200 build.
// Same class names, MakeArgument/MakeInvoker/IsSupportedArgument/Run surface and error behaviour as the
// reference implementations they replace:
//   DeviceFpAintBGemm_Wmma_CShuffle        device/impl/device_fpAintB_gemm_wmma.hpp:70-710
//   DeviceGemmMultipleABD_Xdl_CShuffle     device/impl/device_gemm_multiple_abd_xdl_cshuffle.hpp:71-747
//   DeviceGemmMultipleD_Xdl_CShuffle       device/impl/device_gemm_multiple_d_xdl_cshuffle.hpp
//   DeviceGemmMultiD_Xdl_CShuffle_V3       device/impl/device_gemm_multiple_d_xdl_cshuffle_v3.hpp:143-175
// This header holds the implementations (`*_Sm100`: type parameters + GemmSpec only -- one class serves every shape and picks
// the sm_100a kernel at run time through ckb_gemm_query).  The classes with the REFERENCE names and the reference's full
// template parameter lists (tile shapes, vector widths, pipeline enums: accepted, named, ignored) live in the headers of the
// same names as upstream -- device_fpAintB_gemm_wmma.hpp, device_gemm_multiple_abd_xdl_cshuffle.hpp,
// device_gemm_multiple_d_xdl_cshuffle.hpp, device_gemm_multiple_d_xdl_cshuffle_v3.hpp -- and derive from these, so that the
// reference's own instance declarations compile unchanged (tests/cpp/verbatim_instances.cpp).
#pragma once
#include <array>

#include "ck/tensor_operation/gpu/device/device_gemm_dequantB.hpp"
#include "ck/tensor_operation/gpu/device/device_gemm_multiple_abd.hpp"
#include "ck/tensor_operation/gpu/device/device_gemm_multiple_d.hpp"
#include "ck/tensor_operation/gpu/device/gemm_specialization.hpp"
#include "ck/tensor_operation/gpu/device/impl/ckb_gemm_impl.hpp"

namespace ck {
namespace tensor_operation {
namespace device {

namespace ckb_detail {
template <typename DsLayout, typename DsDataType, index_t I, size_t N>
inline void fill_d(ckb_gemm_problem& p, const std::array<const void*, N>& ptrs, const std::array<index_t, N>& strides)
{
    if constexpr(I < index_t(N))
    {
        p.d[I]       = ptrs[I];
        p.d_dtype[I] = dtype_of<tuple_element_t<I, DsDataType>>();
        d_strides<tuple_element_t<I, DsLayout>>(strides[I], p.d_sm[I], p.d_sn[I]);
        fill_d<DsLayout, DsDataType, I + 1, N>(p, ptrs, strides);
    }
}
// GetTypeString(): reference class name + the sm_100a kernel the most recent MakeArgument of this thread selected (the op
// object itself is stateless, as in the reference; the kernel is chosen per problem)
inline std::string type_string(const char* cls, const GemmArgument* a = nullptr)
{
    std::string s = std::string(cls) + "<sm_100a";
    const std::string& k = (a && a->queried) ? std::string(a->plan.kernel_name) : last_kernel_name();
    if(!k.empty()) s += ", " + k;
    return s + ">";
}
// null-safe downcast of the type-erased argument (the reference dereferences dynamic_cast unchecked)
inline const char* flags_tag(uint32_t f)
{
    return (f & CKB_FLAG_SS_1CTA) ? "[1cta]" : (f & CKB_FLAG_SS_2CTA) ? "[2cta]" : (f & CKB_FLAG_PREFILL_TWO_PASS) ? "[two_pass]" : (f & CKB_FLAG_GEMV_SINGLE_PRODUCER) ? "[single_producer]" : (f & CKB_FLAG_W8_STREAMK) ? "[stream_k]" : "";
}
inline const GemmArgument& checked(const BaseArgument* p)
{
    const auto* a = dynamic_cast<const GemmArgument*>(p);
    if(!a) throw std::runtime_error("wrong! the argument was not made by this device operation");
    return *a;
}
} // namespace ckb_detail

// ------------------------------------------------------------------------------------------------
// fp16 A x (biased uint8 B, fp16 per-channel scale): example/64_fpAintB_gemm
template <typename ALayout, typename BLayout, typename CLayout, typename ADataType, typename BDataType,
          typename ScaleDataType, typename CDataType, typename AccDataType, typename CShuffleDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation,
          GemmSpecialization GemmSpec = GemmSpecialization::Default, uint32_t KernelFlags = 0>
struct DeviceFpAintBGemm_Sm100
    : public DeviceGemm_dequantB<ALayout, BLayout, CLayout, ADataType, BDataType, CDataType, AElementwiseOperation,
                                 BElementwiseOperation, CElementwiseOperation>
{
    using Argument = ckb_detail::GemmArgument;
    using Invoker  = ckb_detail::GemmInvoker;

    static Argument MakeArgument(const ADataType* p_a, const BDataType* p_b, const ScaleDataType* p_scale, CDataType* p_c,
                                 index_t M, index_t N, index_t K, index_t StrideA, index_t StrideB, index_t StrideC,
                                 AElementwiseOperation, BElementwiseOperation, CElementwiseOperation)
    {
        Argument arg;
        ckb_gemm_problem& p = arg.problem;
        p.M = M; p.N = N; p.K = K;
        p.a_dtype = ckb_detail::dtype_of<ADataType>(); p.a_layout = ckb_detail::layout_of<ALayout>(); p.a = p_a; p.lda = StrideA;
        p.b_dtype = ckb_detail::dtype_of<BDataType>(); p.b_layout = ckb_detail::layout_of<BLayout>(); p.b = p_b; p.ldb = StrideB;
        // device_fpAintB_gemm_wmma.hpp:381: the scale is always read with stride 0 along K (one value per column)
        p.b_op = CKB_BOP_FPAINTB; p.s_dtype = ckb_detail::dtype_of<ScaleDataType>(); p.scale = p_scale;
        p.cde_op = CElementwiseOperation::kCkbCde;
        p.e_dtype = ckb_detail::dtype_of<CDataType>(); p.e = p_c; p.lde = StrideC;
        p.flags = KernelFlags;
        arg.Query();
        return arg;
    }
    static Invoker MakeInvoker() { return Invoker{}; }
    static bool IsSupportedArgument(const Argument& arg) { return arg.plan.supported != 0; }
    bool IsSupportedArgument(const BaseArgument* p) override { return IsSupportedArgument(ckb_detail::checked(p)); }

    std::unique_ptr<BaseArgument> MakeArgumentPointer(const void* p_a, const void* p_b, const void* p_scale, void* p_c,
                                                      index_t M, index_t N, index_t K, index_t StrideA, index_t StrideB,
                                                      index_t StrideC, AElementwiseOperation a, BElementwiseOperation b,
                                                      CElementwiseOperation c) override
    {
        return std::make_unique<Argument>(MakeArgument(static_cast<const ADataType*>(p_a), static_cast<const BDataType*>(p_b),
                                                       static_cast<const ScaleDataType*>(p_scale), static_cast<CDataType*>(p_c), M, N,
                                                       K, StrideA, StrideB, StrideC, a, b, c));
    }
    std::unique_ptr<BaseInvoker> MakeInvokerPointer() override { return std::make_unique<Invoker>(); }
    // honest workspace reporting (the reference ops on this path report 0 and need none): stream-K partial sums for decode-shaped
    // problems, plus the K-major copy of a row-major int8 B.  A client that ignores it still runs (the argument then owns the bytes).
    size_t GetWorkSpaceSize(const BaseArgument* p) const override { return ckb_detail::workspace_size(p); }
    void SetWorkSpacePointer(BaseArgument* p, void* ws, const StreamConfig& sc = StreamConfig{}) const override
    {
        ckb_detail::set_workspace(p, ws, sc);
    }
    virtual const char* ClassName() const { return "DeviceFpAintBGemm_Sm100"; }
    std::string GetTypeString() const override { return ckb_detail::type_string(ClassName()) + ckb_detail::flags_tag(KernelFlags); }
};

// ------------------------------------------------------------------------------------------------
// bf16 A x (int8 B0 [, bf16 B1 scale]) with D tensors: example/60, client_example/30
template <typename AsLayout, typename BsLayout, typename DsLayout, typename ELayout, typename AsDataType,
          typename BsDataType, typename AccDataType, typename CShuffleDataType, typename DsDataType, typename EDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CDEElementwiseOperation,
          GemmSpecialization GemmSpec = GemmSpecialization::Default, uint32_t KernelFlags = 0>
struct DeviceGemmMultipleABD_Sm100
    : public DeviceGemmMultipleABD<AsLayout, BsLayout, DsLayout, ELayout, AsDataType, BsDataType, DsDataType, EDataType,
                                   AElementwiseOperation, BElementwiseOperation, CDEElementwiseOperation>
{
    using Argument = ckb_detail::GemmArgument;
    using Invoker  = ckb_detail::GemmInvoker;
    static constexpr index_t NumATensor = AsDataType::Size();
    static constexpr index_t NumBTensor = BsDataType::Size();
    static constexpr index_t NumDTensor = DsDataType::Size();
    using A0DataType = tuple_element_t<0, AsDataType>;
    using B0DataType = tuple_element_t<0, BsDataType>;

    static Argument MakeArgument(std::array<const void*, NumATensor> p_as, std::array<const void*, NumBTensor> p_bs,
                                 std::array<const void*, NumDTensor> p_ds, void* p_e, index_t M, index_t N, index_t K,
                                 std::array<index_t, NumATensor> StrideAs, std::array<index_t, NumBTensor> StrideBs,
                                 std::array<index_t, NumDTensor> StrideDs, index_t StrideE, AElementwiseOperation,
                                 BElementwiseOperation, CDEElementwiseOperation cde)
    {
        static_assert(NumATensor == 1 && NumBTensor >= 1 && NumBTensor <= 2 && NumDTensor <= CKB_MAX_D, "A0, B0[,B1], <=2 Ds");
        Argument arg;
        ckb_gemm_problem& p = arg.problem;
        p.M = M; p.N = N; p.K = K;
        p.a_dtype = ckb_detail::dtype_of<A0DataType>(); p.a_layout = ckb_detail::layout_of<tuple_element_t<0, AsLayout>>();
        p.a = p_as[0]; p.lda = StrideAs[0];
        p.b_dtype = ckb_detail::dtype_of<B0DataType>(); p.b_layout = ckb_detail::layout_of<tuple_element_t<0, BsLayout>>();
        p.b = p_bs[0]; p.ldb = StrideBs[0];
        if constexpr(NumBTensor == 2)
        {
            // B1 = per-channel scale passed with StrideB1 = 0 (client_example/30...:180): one value per column
            p.b_op    = BElementwiseOperation::kCkbBOp;
            p.s_dtype = ckb_detail::dtype_of<tuple_element_t<1, BsDataType>>();
            p.scale   = p_bs[1];
            if(StrideBs[1] != 0) arg.why = "B1 must be a per-channel vector (StrideB1 == 0)";
        }
        else
            p.b_op = CKB_BOP_PASSTHROUGH;
        p.num_d = NumDTensor;
        ckb_detail::fill_d<DsLayout, DsDataType, 0, size_t(NumDTensor)>(p, p_ds, StrideDs);
        p.cde_op = CDEElementwiseOperation::kCkbCde;
        p.cde_scale = ckb_detail::cde_scale_of(cde);
        p.act = ckb_detail::cde_act_of<CDEElementwiseOperation>();
        p.e_dtype = ckb_detail::dtype_of<EDataType>(); p.e = p_e; p.lde = StrideE;
        p.flags = KernelFlags;
        if(p.a_layout == CKB_COL && arg.why.empty()) arg.DeferColMajorA(); // km_*: transposed on the launch stream in front of every Run
        if(p.b_layout == CKB_ROW && p.b_dtype == CKB_I8 && arg.why.empty())
            arg.DeferRowMajorB(); // mk_kn_mn: no allocation and no launch here (see GemmArgument::DeferRowMajorB)
        else if(arg.why.empty()) arg.Query();
        return arg;
    }
    static Invoker MakeInvoker() { return Invoker{}; }
    static bool IsSupportedArgument(const Argument& arg) { return arg.plan.supported != 0; }
    bool IsSupportedArgument(const BaseArgument* p) override { return IsSupportedArgument(ckb_detail::checked(p)); }
    std::unique_ptr<BaseArgument>
    MakeArgumentPointer(std::array<const void*, NumATensor> p_as, std::array<const void*, NumBTensor> p_bs,
                        std::array<const void*, NumDTensor> p_ds, void* p_e, index_t M, index_t N, index_t K,
                        std::array<index_t, NumATensor> StrideAs, std::array<index_t, NumBTensor> StrideBs,
                        std::array<index_t, NumDTensor> StrideDs, index_t StrideE, AElementwiseOperation a,
                        BElementwiseOperation b, CDEElementwiseOperation c) override
    {
        return std::make_unique<Argument>(MakeArgument(p_as, p_bs, p_ds, p_e, M, N, K, StrideAs, StrideBs, StrideDs, StrideE, a, b, c));
    }
    std::unique_ptr<BaseInvoker> MakeInvokerPointer() override { return std::make_unique<Invoker>(); }
    // honest workspace reporting (the reference ops on this path report 0 and need none): stream-K partial sums for decode-shaped
    // problems, plus the K-major copy of a row-major int8 B.  A client that ignores it still runs (the argument then owns the bytes).
    size_t GetWorkSpaceSize(const BaseArgument* p) const override { return ckb_detail::workspace_size(p); }
    void SetWorkSpacePointer(BaseArgument* p, void* ws, const StreamConfig& sc = StreamConfig{}) const override
    {
        ckb_detail::set_workspace(p, ws, sc);
    }
    virtual const char* ClassName() const { return "DeviceGemmMultipleABD_Sm100"; }
    std::string GetTypeString() const override { return ckb_detail::type_string(ClassName()) + ckb_detail::flags_tag(KernelFlags); }
};

// ------------------------------------------------------------------------------------------------
// single A, single B, D tensors: int8 x int8 quantized GEMM (example/14) and MultiplyMultiply (example/65)
template <typename ALayout, typename BLayout, typename DsLayout, typename ELayout, typename ADataType, typename BDataType,
          typename AccDataType, typename CShuffleDataType, typename DsDataType, typename EDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CDEElementwiseOperation,
          GemmSpecialization GemmSpec = GemmSpecialization::Default, uint32_t KernelFlags = 0>
struct DeviceGemmMultipleD_Sm100
    : public DeviceGemmMultipleD<ALayout, BLayout, DsLayout, ELayout, ADataType, BDataType, DsDataType, EDataType,
                                 AElementwiseOperation, BElementwiseOperation, CDEElementwiseOperation>
{
    using Argument = ckb_detail::GemmArgument;
    using Invoker  = ckb_detail::GemmInvoker;
    static constexpr index_t NumDTensor = DsDataType::Size();

    static Argument MakeArgument(const void* p_a, const void* p_b, std::array<const void*, NumDTensor> p_ds, void* p_e,
                                 index_t M, index_t N, index_t K, index_t StrideA, index_t StrideB,
                                 std::array<index_t, NumDTensor> StrideDs, index_t StrideE, AElementwiseOperation,
                                 BElementwiseOperation, CDEElementwiseOperation cde)
    {
        static_assert(NumDTensor <= CKB_MAX_D, "at most two D tensors");
        Argument arg;
        ckb_gemm_problem& p = arg.problem;
        p.M = M; p.N = N; p.K = K;
        p.a_dtype = ckb_detail::dtype_of<ADataType>(); p.a_layout = ckb_detail::layout_of<ALayout>(); p.a = p_a; p.lda = StrideA;
        p.b_dtype = ckb_detail::dtype_of<BDataType>(); p.b_layout = ckb_detail::layout_of<BLayout>(); p.b = p_b; p.ldb = StrideB;
        p.b_op = CKB_BOP_PASSTHROUGH;
        p.num_d = NumDTensor;
        ckb_detail::fill_d<DsLayout, DsDataType, 0, size_t(NumDTensor)>(p, p_ds, StrideDs);
        p.cde_op = CDEElementwiseOperation::kCkbCde;
        p.cde_scale = ckb_detail::cde_scale_of(cde);
        p.act = ckb_detail::cde_act_of<CDEElementwiseOperation>();
        p.e_dtype = ckb_detail::dtype_of<EDataType>(); p.e = p_e; p.lde = StrideE;
        p.flags = KernelFlags;
        arg.Query();
        return arg;
    }
    static Invoker MakeInvoker() { return Invoker{}; }
    static bool IsSupportedArgument(const Argument& arg) { return arg.plan.supported != 0; }
    bool IsSupportedArgument(const BaseArgument* p) override { return IsSupportedArgument(ckb_detail::checked(p)); }
    std::unique_ptr<BaseArgument> MakeArgumentPointer(const void* p_a, const void* p_b, std::array<const void*, NumDTensor> p_ds,
                                                      void* p_e, index_t M, index_t N, index_t K, index_t StrideA,
                                                      index_t StrideB, std::array<index_t, NumDTensor> StrideDs,
                                                      index_t StrideE, AElementwiseOperation a, BElementwiseOperation b,
                                                      CDEElementwiseOperation c) override
    {
        return std::make_unique<Argument>(MakeArgument(p_a, p_b, p_ds, p_e, M, N, K, StrideA, StrideB, StrideDs, StrideE, a, b, c));
    }
    std::unique_ptr<BaseInvoker> MakeInvokerPointer() override { return std::make_unique<Invoker>(); }
    // honest workspace reporting (the reference ops on this path report 0 and need none): stream-K partial sums for decode-shaped
    // problems, plus the K-major copy of a row-major int8 B.  A client that ignores it still runs (the argument then owns the bytes).
    size_t GetWorkSpaceSize(const BaseArgument* p) const override { return ckb_detail::workspace_size(p); }
    void SetWorkSpacePointer(BaseArgument* p, void* ws, const StreamConfig& sc = StreamConfig{}) const override
    {
        ckb_detail::set_workspace(p, ws, sc);
    }
    virtual const char* ClassName() const { return "DeviceGemmMultipleD_Sm100"; }
    std::string GetTypeString() const override { return ckb_detail::type_string(ClassName()) + ckb_detail::flags_tag(KernelFlags); }
};


// ------------------------------------------------------------------------------------------------
// [EXT] W4A16 with K-blocked scales / zero points on pre-packed int4 weights
template <typename ALayout, typename CLayout, typename ADataType, typename CDataType, index_t GroupSize,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation>
struct DeviceGemm_dequantB_GroupScale_Sm100
    : public DeviceGemm_dequantB_GroupScale<ALayout, CLayout, ADataType, CDataType, GroupSize, AElementwiseOperation,
                                            BElementwiseOperation, CElementwiseOperation>
{
    using Argument = ckb_detail::GemmArgument;
    using Invoker  = ckb_detail::GemmInvoker;
    static Argument MakeArgument(const void* p_a, const void* p_b_prepacked, const void* p_scale, const void* p_zero, void* p_c,
                                 index_t M, index_t N, index_t K, index_t StrideA, index_t StrideC, index_t ZeroPointDefault,
                                 AElementwiseOperation, BElementwiseOperation, CElementwiseOperation)
    {
        Argument arg;
        ckb_gemm_problem& p = arg.problem;
        p.M = M; p.N = N; p.K = K;
        p.a_dtype = ckb_detail::dtype_of<ADataType>(); p.a_layout = ckb_detail::layout_of<ALayout>(); p.a = p_a; p.lda = StrideA;
        p.b_dtype = CKB_I4; p.b_layout = CKB_PREPACKED; p.b = p_b_prepacked;
        p.b_op = BElementwiseOperation::kCkbBOp; p.s_dtype = p.a_dtype; p.scale = p_scale; p.zero = p_zero;
        p.zero_const = ZeroPointDefault; p.group_size = GroupSize;
        p.cde_op = CElementwiseOperation::kCkbCde;
        p.e_dtype = ckb_detail::dtype_of<CDataType>(); p.e = p_c; p.lde = StrideC;
        arg.Query();
        return arg;
    }
    static Invoker MakeInvoker() { return Invoker{}; }
    static bool IsSupportedArgument(const Argument& arg) { return arg.plan.supported != 0; }
    bool IsSupportedArgument(const BaseArgument* p) override { return IsSupportedArgument(ckb_detail::checked(p)); }
    std::unique_ptr<BaseArgument> MakeArgumentPointer(const void* p_a, const void* p_b, const void* p_scale, const void* p_zero,
                                                      void* p_c, index_t M, index_t N, index_t K, index_t StrideA, index_t StrideC,
                                                      index_t zp, AElementwiseOperation a, BElementwiseOperation b,
                                                      CElementwiseOperation c) override
    {
        return std::make_unique<Argument>(MakeArgument(p_a, p_b, p_scale, p_zero, p_c, M, N, K, StrideA, StrideC, zp, a, b, c));
    }
    std::unique_ptr<BaseInvoker> MakeInvokerPointer() override { return std::make_unique<Invoker>(); }
    size_t GetWorkSpaceSize(const BaseArgument* p) const override { return ckb_detail::workspace_size(p); }
    void SetWorkSpacePointer(BaseArgument* p, void* ws, const StreamConfig& sc = StreamConfig{}) const override
    {
        ckb_detail::set_workspace(p, ws, sc);
    }
    std::string GetTypeString() const override { return ckb_detail::type_string("DeviceGemm_dequantB_GroupScale_Sm100"); }
};

} // namespace device
} // namespace tensor_operation
} // namespace ck
