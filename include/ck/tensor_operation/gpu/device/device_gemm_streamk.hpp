// SPDX-License-Identifier: MIT
// Interface mirrored from Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence): names, template
// parameters and MakeArgumentPointer signature must match for client code to compile unchanged.
// DeviceGemmStreamK interface (include/ck/tensor_operation/gpu/device/device_gemm_streamk.hpp:15-60)
#pragma once
#include <memory>
#include "ck/tensor_operation/gpu/device/device_base.hpp"
namespace ck {
namespace tensor_operation {
namespace device {
template <typename ALayout, typename BLayout, typename CLayout, typename ADataType, typename BDataType, typename CDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation>
struct DeviceGemmStreamK : public BaseOperator
{
    virtual std::unique_ptr<BaseArgument>
    MakeArgumentPointer(const void* p_a, const void* p_b, void* p_c, ck::index_t M, ck::index_t N, ck::index_t K, ck::index_t StrideA,
                        ck::index_t StrideB, ck::index_t StrideC, AElementwiseOperation a_element_op,
                        BElementwiseOperation b_element_op, CElementwiseOperation c_element_op, ck::index_t NumSKBlocks = 0) = 0;
    virtual std::unique_ptr<BaseInvoker> MakeInvokerPointer() = 0;
};
template <typename ALayout, typename BLayout, typename CLayout, typename ADataType, typename BDataType, typename CDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation>
using DeviceGemmStreamKPtr = std::unique_ptr<DeviceGemmStreamK<ALayout, BLayout, CLayout, ADataType, BDataType, CDataType,
                                                               AElementwiseOperation, BElementwiseOperation, CElementwiseOperation>>;
} // namespace device
} // namespace tensor_operation
} // namespace ck
