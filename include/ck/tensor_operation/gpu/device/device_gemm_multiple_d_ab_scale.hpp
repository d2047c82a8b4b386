// SPDX-License-Identifier: MIT
// Interface mirrored from Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence): names, template
// parameters and MakeArgumentPointer signature must match for client code to compile unchanged.
// DeviceGemmMultipleD_ABScale interface (include/ck/tensor_operation/gpu/device/device_gemm_multiple_d_ab_scale.hpp:15-59):
// block-scaled operands, the reference pattern of the [EXT] W4A16 group-scale op (ScaleBlockK = the quantization group).
#pragma once
#include <array>
#include <memory>
#include "ck/tensor_operation/gpu/device/device_base.hpp"
#include "ck/utility/data_type.hpp"
namespace ck {
namespace tensor_operation {
namespace device {
template <typename ALayout, typename BLayout, typename DsLayout, typename ELayout, typename ADataType, typename AScaleType,
          typename BDataType, typename BScaleType, typename DsDataType, typename EDataType, index_t ScaleBlockM,
          index_t ScaleBlockN, index_t ScaleBlockK, typename AElementwiseOperation, typename BElementwiseOperation,
          typename CDEElementwiseOperation>
struct DeviceGemmMultipleD_ABScale : public BaseOperator
{
    static constexpr index_t NumDTensor = DsDataType::Size();
    virtual std::unique_ptr<BaseArgument>
    MakeArgumentPointer(const void* p_a, const void* p_b, std::array<const void*, NumDTensor> p_ds, void* p_e, const ck::index_t M,
                        const ck::index_t N, const ck::index_t K, const ck::index_t StrideA, const ck::index_t StrideB,
                        const std::array<ck::index_t, NumDTensor> StrideDs, const ck::index_t StrideE, const void* p_a_scale,
                        const void* p_b_scale, AElementwiseOperation a_element_op, BElementwiseOperation b_element_op,
                        CDEElementwiseOperation cde_element_op) = 0;
    virtual std::unique_ptr<BaseInvoker> MakeInvokerPointer() = 0;
};
} // namespace device
} // namespace tensor_operation
} // namespace ck
