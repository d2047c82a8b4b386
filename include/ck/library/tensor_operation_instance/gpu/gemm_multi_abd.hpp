// Instance lookup for bf16 A x int8 B multi-ABD GEMM: the 16 add_device_gemm_xdl_multi_abd_* entry points and the
// factory specialisations of library/include/ck/library/tensor_operation_instance/gpu/gemm_multi_abd.hpp:26-750.
// In the reference each entry point registers 18 tile-shape instantiations (8 "comp" + 10 "mem",
// .../gemm_multi_abd/device_gemm_xdl_multi_abd_bf16_i8_bf16_mk_kn_mn_common.hpp:80-124); here it registers ONE
// sm_100a operator that chooses its kernel per problem at run time.
#pragma once
#include "ck/library/tensor_operation_instance/device_operation_instance_factory.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_multiple_abd_xdl_cshuffle.hpp"

namespace ck {
namespace tensor_operation {
namespace device {
namespace instance {

namespace multi_abd_detail {
template <typename ALayout, typename BsLayout, typename DsLayout, typename BsDataType, typename DsDataType, typename BOp, typename CDEOp>
using Iface = DeviceGemmMultipleABD<ck::Tuple<ALayout>, BsLayout, DsLayout, Row, ck::Tuple<BF16>, BsDataType, DsDataType, BF16,
                                    PassThrough, BOp, CDEOp>;
template <typename ALayout, typename BsLayout, typename DsLayout, typename BsDataType, typename DsDataType, typename BOp, typename CDEOp>
using Impl = DeviceGemmMultipleABD_Xdl_CShuffle<ck::Tuple<ALayout>, BsLayout, DsLayout, Row, ck::Tuple<BF16>, BsDataType, F32, F32,
                                                DsDataType, BF16, PassThrough, BOp, CDEOp>;
template <typename ALayout, typename BsLayout, typename DsLayout, typename BsDataType, typename DsDataType, typename BOp, typename CDEOp>
void add(std::vector<std::unique_ptr<Iface<ALayout, BsLayout, DsLayout, BsDataType, DsDataType, BOp, CDEOp>>>& v)
{
    v.push_back(std::make_unique<Impl<ALayout, BsLayout, DsLayout, BsDataType, DsDataType, BOp, CDEOp>>());
}
using B2 = ck::Tuple<I8, BF16>;
using B1 = ck::Tuple<I8>;
using D0 = ck::Tuple<>;
using D1 = ck::Tuple<BF16>;
using D2 = ck::Tuple<BF16, BF16>;
} // namespace multi_abd_detail

#define CKB_MULTI_ABD_ENTRY(name, AL, BL, DL, BT, DT, BOP, CDE)                                                              \
    inline void name(std::vector<std::unique_ptr<multi_abd_detail::Iface<AL, BL, DL, BT, DT, BOP, CDE>>>& instances)          \
    {                                                                                                                        \
        multi_abd_detail::add<AL, BL, DL, BT, DT, BOP, CDE>(instances);                                                      \
    }
#define CKB_T(...) ck::Tuple<__VA_ARGS__>
// B-op = Multiply(B0 int8, B1 bf16 scale): 3 layouts x {bias_gelu, bias, gelu, plain}   (gemm_multi_abd.hpp:26-182)
#define CKB_MULTI_ABD_LAYOUT(tag, AL, BL)                                                                                                         \
    CKB_MULTI_ABD_ENTRY(add_device_gemm_xdl_multi_abd_bf16_i8_bf16_##tag##_bias_gelu_v1_instances, AL, CKB_T(BL, BL), CKB_T(Row), multi_abd_detail::B2, multi_abd_detail::D1, Multiply, AddFastGelu) \
    CKB_MULTI_ABD_ENTRY(add_device_gemm_xdl_multi_abd_bf16_i8_bf16_##tag##_bias_v1_instances, AL, CKB_T(BL, BL), CKB_T(Row), multi_abd_detail::B2, multi_abd_detail::D1, Multiply, Add)              \
    CKB_MULTI_ABD_ENTRY(add_device_gemm_xdl_multi_abd_bf16_i8_bf16_##tag##_gelu_v1_instances, AL, CKB_T(BL, BL), CKB_T(), multi_abd_detail::B2, multi_abd_detail::D0, Multiply, FastGelu)           \
    CKB_MULTI_ABD_ENTRY(add_device_gemm_xdl_multi_abd_bf16_i8_bf16_##tag##_v1_instances, AL, CKB_T(BL, BL), CKB_T(), multi_abd_detail::B2, multi_abd_detail::D0, Multiply, PassThrough)
CKB_MULTI_ABD_LAYOUT(mk_kn_mn, Row, Row)
CKB_MULTI_ABD_LAYOUT(mk_nk_mn, Row, Col)
CKB_MULTI_ABD_LAYOUT(km_kn_mn, Col, Row)
// scale as D0 (B-op = PassThrough): mk_kn_mn only   (gemm_multi_abd.hpp:184-235)
CKB_MULTI_ABD_ENTRY(add_device_gemm_xdl_multi_abd_multiply_bf16_i8_bf16_mk_kn_mn_bias_gelu_v1_instances, Row, CKB_T(Row), CKB_T(Row, Row), multi_abd_detail::B1, multi_abd_detail::D2, PassThrough, MultiplyAddFastGelu)
CKB_MULTI_ABD_ENTRY(add_device_gemm_xdl_multi_abd_multiply_bf16_i8_bf16_mk_kn_mn_bias_v1_instances, Row, CKB_T(Row), CKB_T(Row, Row), multi_abd_detail::B1, multi_abd_detail::D2, PassThrough, MultiplyAdd)
CKB_MULTI_ABD_ENTRY(add_device_gemm_xdl_multi_abd_multiply_bf16_i8_bf16_mk_kn_mn_gelu_v1_instances, Row, CKB_T(Row), CKB_T(Row), multi_abd_detail::B1, multi_abd_detail::D1, PassThrough, MultiplyFastGelu)
CKB_MULTI_ABD_ENTRY(add_device_gemm_xdl_multi_abd_multiply_bf16_i8_bf16_mk_kn_mn_v1_instances, Row, CKB_T(Row), CKB_T(Row), multi_abd_detail::B1, multi_abd_detail::D1, PassThrough, Multiply)
#undef CKB_MULTI_ABD_LAYOUT
#undef CKB_MULTI_ABD_ENTRY
#undef CKB_T

// One factory specialisation covers the 8 of the reference (keyed there on the functor triple with
// `if constexpr` over layouts and data types, gemm_multi_abd.hpp:248-750).
template <typename AsLayout, typename BsLayout, typename DsLayout, typename ELayout, typename AsDataType, typename BsDataType,
          typename DsDataType, typename EDataType, typename AOp, typename BOp, typename CDEOp>
struct DeviceOperationInstanceFactory<
    DeviceGemmMultipleABD<AsLayout, BsLayout, DsLayout, ELayout, AsDataType, BsDataType, DsDataType, EDataType, AOp, BOp, CDEOp>>
{
    using DeviceOp = DeviceGemmMultipleABD<AsLayout, BsLayout, DsLayout, ELayout, AsDataType, BsDataType, DsDataType, EDataType, AOp,
                                           BOp, CDEOp>;
    static auto GetInstances()
    {
        std::vector<std::unique_ptr<DeviceOp>> op_ptrs;
        if constexpr(is_same_v<tuple_element_t<0, AsDataType>, BF16> && is_same_v<tuple_element_t<0, BsDataType>, I8> &&
                     is_same_v<EDataType, BF16> && is_same_v<ELayout, Row>)
        {
            op_ptrs.push_back(std::make_unique<DeviceGemmMultipleABD_Xdl_CShuffle<AsLayout, BsLayout, DsLayout, ELayout, AsDataType,
                                                                                  BsDataType, F32, F32, DsDataType, EDataType, AOp, BOp,
                                                                                  CDEOp>>());
            // second instance: the two-pass prefill form (dequant pass into workspace + dense tcgen05 GEMM) next to the fused kernel;
            // it serves M > 64 only and needs GetWorkSpaceSize() bytes
            op_ptrs.push_back(std::make_unique<DeviceGemmMultipleABD_Sm100<AsLayout, BsLayout, DsLayout, ELayout, AsDataType, BsDataType, F32, F32,
                                                                           DsDataType, EDataType, AOp, BOp, CDEOp, GemmSpecialization::Default,
                                                                           CKB_FLAG_PREFILL_TWO_PASS>>());
            // third instance: decode rows (M <= 64) on the stream-K tcgen05 kernel (weights as the TMEM A operand) instead of the
            // N-split HMMA kernel
            op_ptrs.push_back(std::make_unique<DeviceGemmMultipleABD_Sm100<AsLayout, BsLayout, DsLayout, ELayout, AsDataType, BsDataType, F32, F32,
                                                                           DsDataType, EDataType, AOp, BOp, CDEOp, GemmSpecialization::Default,
                                                                           CKB_FLAG_W8_STREAMK>>());
        }
        return op_ptrs;
    }
};

} // namespace instance
} // namespace device
} // namespace tensor_operation
} // namespace ck
