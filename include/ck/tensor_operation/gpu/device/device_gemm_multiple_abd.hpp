// SPDX-License-Identifier: MIT
// Interface mirrored from Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence): declarations must match the
// reference for client code to compile unchanged; see the reference file cited below.
// DeviceGemmMultipleABD interface (include/ck/tensor_operation/gpu/device/device_gemm_multiple_abd.hpp:33-56)
#pragma once
#include <array>
#include <memory>
#include "ck/tensor_operation/gpu/device/device_base.hpp"
#include "ck/utility/data_type.hpp"
namespace ck {
namespace tensor_operation {
namespace device {
template <typename AsLayout, typename BsLayout, typename DsLayout, typename ELayout, typename AsDataType,
          typename BsDataType, typename DsDataType, typename EDataType, typename AElementwiseOperation,
          typename BElementwiseOperation, typename CDEElementwiseOperation>
struct DeviceGemmMultipleABD : public BaseOperator
{
    static constexpr index_t NumATensor = AsDataType::Size();
    static constexpr index_t NumBTensor = BsDataType::Size();
    static constexpr index_t NumDTensor = DsDataType::Size();
    virtual std::unique_ptr<BaseArgument>
    MakeArgumentPointer(std::array<const void*, NumATensor> p_as, std::array<const void*, NumBTensor> p_bs,
                        std::array<const void*, NumDTensor> p_ds, void* p_e, ck::index_t M, ck::index_t N, ck::index_t K,
                        std::array<ck::index_t, NumATensor> StrideAs, std::array<ck::index_t, NumBTensor> StrideBs,
                        std::array<ck::index_t, NumDTensor> StrideDs, ck::index_t StrideE,
                        AElementwiseOperation a_element_op, BElementwiseOperation b_element_op,
                        CDEElementwiseOperation cde_element_op) = 0;
    virtual std::unique_ptr<BaseInvoker> MakeInvokerPointer() = 0;
};
} // namespace device
} // namespace tensor_operation
} // namespace ck
