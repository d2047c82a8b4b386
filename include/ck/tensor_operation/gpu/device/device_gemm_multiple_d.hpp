// SPDX-License-Identifier: MIT
// Interface mirrored from Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence): declarations must match the
// reference for client code to compile unchanged; see the reference file cited below.
// DeviceGemmMultipleD interface (include/ck/tensor_operation/gpu/device/device_gemm_multiple_d.hpp:33-54)
#pragma once
#include <array>
#include <memory>
#include "ck/tensor_operation/gpu/device/device_base.hpp"
#include "ck/utility/data_type.hpp"
namespace ck {
namespace tensor_operation {
namespace device {
template <typename ALayout, typename BLayout, typename DsLayout, typename ELayout, typename ADataType,
          typename BDataType, typename DsDataType, typename EDataType, typename AElementwiseOperation,
          typename BElementwiseOperation, typename CDEElementwiseOperation>
struct DeviceGemmMultipleD : public BaseOperator
{
    static constexpr index_t NumDTensor = DsDataType::Size();
    virtual std::unique_ptr<BaseArgument>
    MakeArgumentPointer(const void* p_a, const void* p_b, std::array<const void*, NumDTensor> p_ds, void* p_e,
                        ck::index_t M, ck::index_t N, ck::index_t K, ck::index_t StrideA, ck::index_t StrideB,
                        std::array<ck::index_t, NumDTensor> StrideDs, ck::index_t StrideE,
                        AElementwiseOperation a_element_op, BElementwiseOperation b_element_op,
                        CDEElementwiseOperation cde_element_op) = 0;
    virtual std::unique_ptr<BaseInvoker> MakeInvokerPointer() = 0;
};
} // namespace device
} // namespace tensor_operation
} // namespace ck
