// Instance lookup for the split-K + reduce universal GEMM, bf16 A x int8 B -> bf16 (mk_kn_mn): the seven
// add_device_gemm_xdl_universal_reduce_bf16_i8_bf16_mk_kn_mn_* entry points and the factory of
// library/include/ck/library/tensor_operation_instance/gpu/gemm_universal_reduce.hpp:156-330.  The reference's comp / mem_v2
// x default / kpadding / mnpadding / mnkpadding variants are tile-shape and padding choices; the sm_100a operator handles
// every shape, so only the first entry point registers it and the others add nothing (no duplicate instances).
#pragma once
#include "ck/library/tensor_operation_instance/device_operation_instance_factory.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_xdl_cshuffle_v3r1.hpp"

namespace ck {
namespace tensor_operation {
namespace device {
namespace instance {

namespace universal_reduce_detail {
using Iface = DeviceGemmV2R1<Row, Row, ck::Tuple<>, Row, BF16, I8, ck::Tuple<>, BF16, PassThrough, PassThrough, PassThrough>;
using Impl  = DeviceGemm_Xdl_CShuffleV3R1<Row, Row, ck::Tuple<>, Row, BF16, I8, ck::Tuple<>, BF16, F32, BF16, PassThrough, PassThrough, PassThrough>;
} // namespace universal_reduce_detail

inline void add_device_gemm_xdl_universal_reduce_bf16_i8_bf16_mk_kn_mn_comp_default_instances(
    std::vector<std::unique_ptr<universal_reduce_detail::Iface>>& instances)
{
    instances.push_back(std::make_unique<universal_reduce_detail::Impl>());
}
#define CKB_UR_EMPTY(name) inline void name(std::vector<std::unique_ptr<universal_reduce_detail::Iface>>&) {}
CKB_UR_EMPTY(add_device_gemm_xdl_universal_reduce_bf16_i8_bf16_mk_kn_mn_comp_kpadding_instances)
CKB_UR_EMPTY(add_device_gemm_xdl_universal_reduce_bf16_i8_bf16_mk_kn_mn_comp_mnpadding_instances)
CKB_UR_EMPTY(add_device_gemm_xdl_universal_reduce_bf16_i8_bf16_mk_kn_mn_comp_mnkpadding_instances)
CKB_UR_EMPTY(add_device_gemm_xdl_universal_reduce_bf16_i8_bf16_mk_kn_mn_mem_v2_default_instances)
CKB_UR_EMPTY(add_device_gemm_xdl_universal_reduce_bf16_i8_bf16_mk_kn_mn_mem_v2_kpadding_instances)
CKB_UR_EMPTY(add_device_gemm_xdl_universal_reduce_bf16_i8_bf16_mk_kn_mn_mem_v2_mnkpadding_instances)
#undef CKB_UR_EMPTY

template <typename ALayout, typename BLayout, typename DsLayout, typename CLayout, typename DsDataType>
struct DeviceOperationInstanceFactory<
    DeviceGemmV2R1<ALayout, BLayout, DsLayout, CLayout, BF16, I8, DsDataType, BF16, PassThrough, PassThrough, PassThrough>>
{
    using DeviceOp = DeviceGemmV2R1<ALayout, BLayout, DsLayout, CLayout, BF16, I8, DsDataType, BF16, PassThrough, PassThrough, PassThrough>;
    static auto GetInstances()
    {
        std::vector<std::unique_ptr<DeviceOp>> op_ptrs;
        if constexpr(is_same_v<CLayout, Row> && DsDataType::Size() == 0)
            op_ptrs.push_back(std::make_unique<DeviceGemm_Xdl_CShuffleV3R1<ALayout, BLayout, DsLayout, CLayout, BF16, I8, DsDataType, BF16, F32,
                                                                           BF16, PassThrough, PassThrough, PassThrough>>());
        return op_ptrs;
    }
};

} // namespace instance
} // namespace device
} // namespace tensor_operation
} // namespace ck
