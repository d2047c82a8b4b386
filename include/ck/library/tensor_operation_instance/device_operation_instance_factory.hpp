// DeviceOperationInstanceFactory and the alias vocabulary of CK's instance library
// (library/include/ck/library/tensor_operation_instance/device_operation_instance_factory.hpp:16-141).
#pragma once
#include <memory>
#include <vector>

#include "ck/tensor_operation/gpu/device/tensor_layout.hpp"
#include "ck/tensor_operation/gpu/element/element_wise_operation.hpp"
#include "ck/utility/data_type.hpp"

namespace ck {
namespace tensor_operation {
namespace device {
namespace instance {

using F32  = float;
using F16  = ck::half_t;
using BF16 = ck::bhalf_t;
using I8   = int8_t;
using I32  = int32_t;

using Empty_Tuple = ck::Tuple<>;
using F16_Tuple   = ck::Tuple<F16>;
using BF16_Tuple  = ck::Tuple<BF16>;
using F32_Tuple   = ck::Tuple<F32>;
using I32_Tuple   = ck::Tuple<I32>;

using Row = ck::tensor_layout::gemm::RowMajor;
using Col = ck::tensor_layout::gemm::ColumnMajor;
using Row_Tuple = ck::Tuple<Row>;

using PassThrough         = ck::tensor_operation::element_wise::PassThrough;
using Relu                = ck::tensor_operation::element_wise::Relu;
using Scale               = ck::tensor_operation::element_wise::Scale;
using Add                 = ck::tensor_operation::element_wise::Add;
using Multiply            = ck::tensor_operation::element_wise::Multiply;
using FastGelu            = ck::tensor_operation::element_wise::FastGelu;
using AddFastGelu         = ck::tensor_operation::element_wise::AddFastGelu;
using MultiplyAdd         = ck::tensor_operation::element_wise::MultiplyAdd;
using MultiplyFastGelu    = ck::tensor_operation::element_wise::MultiplyFastGelu;
using MultiplyAddFastGelu = ck::tensor_operation::element_wise::MultiplyAddFastGelu;
using MultiplyMultiply    = ck::tensor_operation::element_wise::MultiplyMultiply;
template <typename Activation>
using Activation_Mul_Clamp = ck::tensor_operation::element_wise::Activation_Mul_Clamp<Activation>;
template <typename Activation>
using Add_Activation_Mul_Clamp = ck::tensor_operation::element_wise::Add_Activation_Mul_Clamp<Activation>;

template <typename DeviceOp, typename Tag = void>
struct DeviceOperationInstanceFactory;

// add_device_operation_instances (add_device_operation_instance.hpp:16-30)
template <typename BaseOp, typename NewOp>
void add_device_operation_instance(std::vector<std::unique_ptr<BaseOp>>& op_instances)
{
    op_instances.push_back(std::make_unique<NewOp>());
}

} // namespace instance
} // namespace device
} // namespace tensor_operation
} // namespace ck
