// int8 x int8 quantized GEMM and MultiplyMultiply instance lookup
// (library/include/ck/library/tensor_operation_instance/gpu/quantization/gemm_quantization.hpp,
//  .../gpu/gemm_multiply_multiply.hpp:20-215 -- the reference wires fp8 only; int8 is the [EXT] dtype of cfg4).
#pragma once
#include "ck/library/tensor_operation_instance/device_operation_instance_factory.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_multiple_d_xdl_cshuffle.hpp"

namespace ck {
namespace tensor_operation {
namespace device {
namespace instance {

template <typename ALayout, typename BLayout, typename DsLayout, typename ELayout, typename DsDataType, typename EDataType,
          typename CDEOp>
struct DeviceOperationInstanceFactory<
    DeviceGemmMultipleD<ALayout, BLayout, DsLayout, ELayout, I8, I8, DsDataType, EDataType, PassThrough, PassThrough, CDEOp>>
{
    using DeviceOp = DeviceGemmMultipleD<ALayout, BLayout, DsLayout, ELayout, I8, I8, DsDataType, EDataType, PassThrough, PassThrough, CDEOp>;
    static auto GetInstances()
    {
        std::vector<std::unique_ptr<DeviceOp>> op_ptrs;
        op_ptrs.push_back(std::make_unique<DeviceGemmMultipleD_Xdl_CShuffle<ALayout, BLayout, DsLayout, ELayout, I8, I8, I32, I32,
                                                                            DsDataType, EDataType, PassThrough, PassThrough, CDEOp>>());
        // the two tcgen05 tilings as explicit instances (the default instance above picks per shape): 128x256 tiles, one CTA each,
        // and 256x256 tiles per CTA pair (cta_group::2) -- a real trade-off on this part that the caller can time
        op_ptrs.push_back(std::make_unique<DeviceGemmMultipleD_Sm100<ALayout, BLayout, DsLayout, ELayout, I8, I8, I32, I32, DsDataType, EDataType,
                                                                     PassThrough, PassThrough, CDEOp, GemmSpecialization::Default, CKB_FLAG_SS_1CTA>>());
        op_ptrs.push_back(std::make_unique<DeviceGemmMultipleD_Sm100<ALayout, BLayout, DsLayout, ELayout, I8, I8, I32, I32, DsDataType, EDataType,
                                                                     PassThrough, PassThrough, CDEOp, GemmSpecialization::Default, CKB_FLAG_SS_2CTA>>());
        return op_ptrs;
    }
};

} // namespace instance
} // namespace device
} // namespace tensor_operation
} // namespace ck
