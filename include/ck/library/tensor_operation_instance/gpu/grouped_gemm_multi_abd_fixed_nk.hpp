// Instance lookup for grouped fixed-NK bf16 A x int8 B GEMM: the 12 add_device_grouped_gemm_xdl_fixed_nk_multi_abd_*
// entry points and the factory of library/include/ck/library/tensor_operation_instance/gpu/
// grouped_gemm_multi_abd_fixed_nk.hpp:25-470.  The reference registers 10 tile-shape instantiations per entry point
// (.../grouped_gemm_fixed_nk_multi_abd/..._mk_kn_mn_common.hpp); here each registers ONE sm_100a operator.
// All three layouts of the reference are served: mk_kn_mn, mk_nk_mn and km_kn_mn (column-major A: transposed on the way into shared memory).
#pragma once
#include "ck/library/tensor_operation_instance/device_operation_instance_factory.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_grouped_gemm_multi_abd_xdl_fixed_nk.hpp"

namespace ck {
namespace tensor_operation {
namespace device {
namespace instance {

namespace grouped_fixed_nk_detail {
template <typename AL, typename BL, typename DsLayout, typename DsDataType, typename CDEOp>
using Iface = DeviceGroupedGemmMultiABDFixedNK<ck::Tuple<AL>, ck::Tuple<BL, BL>, DsLayout, Row, ck::Tuple<BF16>, ck::Tuple<I8, BF16>,
                                               DsDataType, BF16, PassThrough, Multiply, CDEOp>;
template <typename AL, typename BL, typename DsLayout, typename DsDataType, typename CDEOp>
using Impl = DeviceGroupedGemm_Xdl_Multi_ABD_Fixed_NK<ck::Tuple<AL>, ck::Tuple<BL, BL>, DsLayout, Row, ck::Tuple<BF16>, ck::Tuple<I8, BF16>,
                                                      F32, F32, DsDataType, BF16, PassThrough, Multiply, CDEOp>;
} // namespace grouped_fixed_nk_detail

#define CKB_GROUPED_ENTRY(name, AL, BL, DL, DT, CDE)                                                                          \
    inline void name(std::vector<std::unique_ptr<grouped_fixed_nk_detail::Iface<AL, BL, DL, DT, CDE>>>& instances)             \
    {                                                                                                                         \
        instances.push_back(std::make_unique<grouped_fixed_nk_detail::Impl<AL, BL, DL, DT, CDE>>());                           \
    }
#define CKB_GROUPED_LAYOUT(tag, AL, BL)                                                                                                                   \
    CKB_GROUPED_ENTRY(add_device_grouped_gemm_xdl_fixed_nk_multi_abd_bf16_i8_bf16_##tag##_bias_gelu_instances, AL, BL, ck::Tuple<Row>, ck::Tuple<BF16>, AddFastGelu) \
    CKB_GROUPED_ENTRY(add_device_grouped_gemm_xdl_fixed_nk_multi_abd_bf16_i8_bf16_##tag##_bias_instances, AL, BL, ck::Tuple<Row>, ck::Tuple<BF16>, Add)              \
    CKB_GROUPED_ENTRY(add_device_grouped_gemm_xdl_fixed_nk_multi_abd_bf16_i8_bf16_##tag##_gelu_instances, AL, BL, ck::Tuple<>, ck::Tuple<>, FastGelu)               \
    CKB_GROUPED_ENTRY(add_device_grouped_gemm_xdl_fixed_nk_multi_abd_bf16_i8_bf16_##tag##_instances, AL, BL, ck::Tuple<>, ck::Tuple<>, PassThrough)
CKB_GROUPED_LAYOUT(mk_kn_mn, Row, Row)
CKB_GROUPED_LAYOUT(mk_nk_mn, Row, Col)
CKB_GROUPED_LAYOUT(km_kn_mn, Col, Row)
#undef CKB_GROUPED_LAYOUT
#undef CKB_GROUPED_ENTRY

template <typename AsLayout, typename BsLayout, typename DsLayout, typename ELayout, typename AsDataType, typename BsDataType,
          typename DsDataType, typename EDataType, typename AOp, typename BOp, typename CDEOp>
struct DeviceOperationInstanceFactory<
    DeviceGroupedGemmMultiABDFixedNK<AsLayout, BsLayout, DsLayout, ELayout, AsDataType, BsDataType, DsDataType, EDataType, AOp, BOp, CDEOp>>
{
    using DeviceOp = DeviceGroupedGemmMultiABDFixedNK<AsLayout, BsLayout, DsLayout, ELayout, AsDataType, BsDataType, DsDataType, EDataType,
                                                      AOp, BOp, CDEOp>;
    static auto GetInstances()
    {
        std::vector<std::unique_ptr<DeviceOp>> op_ptrs;
        if constexpr(is_same_v<tuple_element_t<0, AsDataType>, BF16> && is_same_v<tuple_element_t<0, BsDataType>, I8> &&
                     is_same_v<EDataType, BF16> && is_same_v<ELayout, Row>)
        {
            op_ptrs.push_back(std::make_unique<DeviceGroupedGemm_Xdl_Multi_ABD_Fixed_NK<AsLayout, BsLayout, DsLayout, ELayout, AsDataType,
                                                                                        BsDataType, F32, F32, DsDataType, EDataType, AOp,
                                                                                        BOp, CDEOp>>());
        }
        return op_ptrs;
    }
};

} // namespace instance
} // namespace device
} // namespace tensor_operation
} // namespace ck
