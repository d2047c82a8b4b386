// SPDX-License-Identifier: MIT
// Interface mirrored from Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence): declarations must match the
// reference for client code to compile unchanged; see the reference file cited below.
// DeviceGemmV2 / DeviceGemmV2R1 interfaces (include/ck/tensor_operation/gpu/device/device_gemm_v2.hpp:14-74): plain GEMM with a
// split-K argument (KBatch / KSplit); V2R1 adds D tensors applied by the reduction step.
#pragma once
#include <array>
#include <memory>

#include "ck/tensor_operation/gpu/device/device_base.hpp"
#include "ck/utility/data_type.hpp"

namespace ck {
namespace tensor_operation {
namespace device {

template <typename ALayout, typename BLayout, typename CLayout, typename ADataType, typename BDataType, typename CDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation>
struct DeviceGemmV2 : public BaseOperator
{
    virtual std::unique_ptr<BaseArgument> MakeArgumentPointer(const void* p_a, const void* p_b, void* p_c, ck::index_t M, ck::index_t N,
                                                              ck::index_t K, ck::index_t StrideA, ck::index_t StrideB, ck::index_t StrideC,
                                                              ck::index_t KSplit, AElementwiseOperation a_element_op,
                                                              BElementwiseOperation b_element_op, CElementwiseOperation c_element_op) = 0;
    virtual std::unique_ptr<BaseInvoker> MakeInvokerPointer() = 0;
};

template <typename ALayout, typename BLayout, typename DsLayout, typename CLayout, typename ADataType, typename BDataType,
          typename DsDataType, typename CDataType, typename AElementwiseOperation, typename BElementwiseOperation,
          typename CElementwiseOperation>
struct DeviceGemmV2R1 : public BaseOperator
{
    static constexpr index_t NumDTensor = DsDataType::Size();
    virtual std::unique_ptr<BaseArgument>
    MakeArgumentPointer(const void* p_a, const void* p_b, std::array<const void*, NumDTensor> p_ds, void* p_c, ck::index_t M, ck::index_t N,
                        ck::index_t K, ck::index_t StrideA, ck::index_t StrideB, std::array<ck::index_t, NumDTensor> DsStrides,
                        ck::index_t StrideC, ck::index_t KSplit, AElementwiseOperation a_element_op, BElementwiseOperation b_element_op,
                        CElementwiseOperation c_element_op) = 0;
    virtual std::unique_ptr<BaseInvoker> MakeInvokerPointer() = 0;
};

} // namespace device
} // namespace tensor_operation
} // namespace ck
