// SPDX-License-Identifier: MIT
// Interface mirrored from Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence): declarations must match the
// reference for client code to compile unchanged; see the reference file cited below.
// DeviceGemm_dequantB interface (include/ck/tensor_operation/gpu/device/device_gemm_dequantB.hpp:24-42)
#pragma once
#include <memory>
#include "ck/tensor_operation/gpu/device/device_base.hpp"
namespace ck {
namespace tensor_operation {
namespace device {
template <typename ALayout, typename BLayout, typename CLayout, typename ADataType, typename BDataType,
          typename CDataType, typename AElementwiseOperation, typename BElementwiseOperation,
          typename CElementwiseOperation>
struct DeviceGemm_dequantB : public BaseOperator
{
    virtual std::unique_ptr<BaseArgument>
    MakeArgumentPointer(const void* p_a, const void* p_b, const void* p_scale, void* p_c, ck::index_t M, ck::index_t N,
                        ck::index_t K, ck::index_t StrideA, ck::index_t StrideB, ck::index_t StrideC,
                        AElementwiseOperation a_element_op, BElementwiseOperation b_element_op,
                        CElementwiseOperation c_element_op) = 0;
    virtual std::unique_ptr<BaseInvoker> MakeInvokerPointer() = 0;
};

// [EXT] K-blocked scales + zero points + packed int4; shaped after DeviceGemmMultipleD_ABScale
// (device/device_gemm_multiple_d_ab_scale.hpp:22-59): extra p_zero argument and a GroupSize template integer.
template <typename ALayout, typename CLayout, typename ADataType, typename CDataType, ck::index_t GroupSize,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation>
struct DeviceGemm_dequantB_GroupScale : public BaseOperator
{
    // p_b: buffer produced by ckb_prepack_b_int4; p_scale [K/GroupSize, N] of ADataType; p_zero uint8 [K/GroupSize, N] or null
    virtual std::unique_ptr<BaseArgument>
    MakeArgumentPointer(const void* p_a, const void* p_b_prepacked, const void* p_scale, const void* p_zero, void* p_c,
                        ck::index_t M, ck::index_t N, ck::index_t K, ck::index_t StrideA, ck::index_t StrideC,
                        ck::index_t ZeroPointDefault, AElementwiseOperation a_element_op,
                        BElementwiseOperation b_element_op, CElementwiseOperation c_element_op) = 0;
    virtual std::unique_ptr<BaseInvoker> MakeInvokerPointer() = 0;
};
} // namespace device
} // namespace tensor_operation
} // namespace ck
