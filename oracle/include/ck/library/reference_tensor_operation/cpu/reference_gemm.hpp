// ORACLE (test infrastructure; never included by product code): CK-style host reference operators over
// Tensor<T>, restating
//   ReferenceGemm            library/include/ck/library/reference_tensor_operation/cpu/reference_gemm.hpp:26-153
//   ReferencefpAintBGemm     .../cpu/reference_fpAintB_gemm.hpp:25-173
//   ReferenceGemmMultipleD   .../cpu/reference_gemm_multiple_d.hpp:28-175
// Same template parameters, MakeArgument / MakeInvoker / Run surface, same naive k-ascending loop,
// ComputeType defaults (ComputeTypeA = CDataType, :24-25), ConvertBF16RTN bypass (:70-91) and
// std::thread::hardware_concurrency() parallelisation over the flattened (m,n) space.
#pragma once
#include <thread>

#include "ck/library/utility/host_tensor.hpp"
#include "ck/tensor_operation/gpu/device/device_base.hpp"
#include "ck/tensor_operation/gpu/element/element_wise_operation.hpp"

namespace ck {
namespace tensor_operation {
namespace host {

template <typename ADataType, typename BDataType, typename CDataType, typename AccDataType, typename AElementwiseOperation,
          typename BElementwiseOperation, typename CElementwiseOperation, typename ComputeTypeA = CDataType,
          typename ComputeTypeB = ComputeTypeA>
struct ReferenceGemm : public device::BaseOperator
{
    struct Argument : public device::BaseArgument
    {
        Argument(const Tensor<ADataType>& a_m_k, const Tensor<BDataType>& b_k_n, Tensor<CDataType>& c_m_n, AElementwiseOperation a_op,
                 BElementwiseOperation b_op, CElementwiseOperation c_op)
            : a_m_k_{a_m_k}, b_k_n_{b_k_n}, c_m_n_{c_m_n}, a_element_op_{a_op}, b_element_op_{b_op}, c_element_op_{c_op} {}
        const Tensor<ADataType>& a_m_k_;
        const Tensor<BDataType>& b_k_n_;
        Tensor<CDataType>& c_m_n_;
        AElementwiseOperation a_element_op_;
        BElementwiseOperation b_element_op_;
        CElementwiseOperation c_element_op_;
    };
    struct Invoker : public device::BaseInvoker
    {
        float Run(const Argument& arg)
        {
            auto f_mk_kn_mn = [&](auto m, auto n) {
                const int K = arg.a_m_k_.mDesc.GetLengths()[1];
                AccDataType v_acc = 0;
                ComputeTypeA v_a  = 0;
                ComputeTypeB v_b  = 0;
                for(int k = 0; k < K; ++k)
                {
                    if constexpr(is_same_v<AElementwiseOperation, element_wise::ConvertBF16RTN>)
                        element_wise::PassThrough{}(v_a, arg.a_m_k_(m, k));
                    else
                        arg.a_element_op_(v_a, arg.a_m_k_(m, k));
                    if constexpr(is_same_v<BElementwiseOperation, element_wise::ConvertBF16RTN>)
                        element_wise::PassThrough{}(v_b, arg.b_k_n_(k, n));
                    else
                        arg.b_element_op_(v_b, arg.b_k_n_(k, n));
                    v_acc += ck::type_convert<AccDataType>(v_a) * ck::type_convert<AccDataType>(v_b);
                }
                CDataType v_c = 0;
                arg.c_element_op_(v_c, v_acc);
                arg.c_m_n_(m, n) = v_c;
            };
            make_ParallelTensorFunctor(f_mk_kn_mn, arg.c_m_n_.mDesc.GetLengths()[0], arg.c_m_n_.mDesc.GetLengths()[1])(
                std::thread::hardware_concurrency());
            return 0;
        }
        float Run(const device::BaseArgument* p_arg, const StreamConfig& = StreamConfig{}) override
        {
            return Run(*dynamic_cast<const Argument*>(p_arg));
        }
    };
    bool IsSupportedArgument(const device::BaseArgument*) override { return true; }
    static auto MakeArgument(const Tensor<ADataType>& a, const Tensor<BDataType>& b, Tensor<CDataType>& c, AElementwiseOperation ao,
                             BElementwiseOperation bo, CElementwiseOperation co)
    {
        return Argument{a, b, c, ao, bo, co};
    }
    static auto MakeInvoker() { return Invoker{}; }
    std::string GetTypeString() const override { return "ReferenceGemm"; }
};

template <typename ADataType, typename BDataType, typename ScaleDataType, typename CDataType, typename AccDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation>
struct ReferencefpAintBGemm : public device::BaseOperator
{
    struct Argument : public device::BaseArgument
    {
        Argument(const Tensor<ADataType>& a_m_k, const Tensor<BDataType>& b_k_n, const Tensor<ScaleDataType>& scale_k_n,
                 Tensor<CDataType>& c_m_n, AElementwiseOperation a_op, BElementwiseOperation b_op, CElementwiseOperation c_op)
            : a_m_k_{a_m_k}, b_k_n_{b_k_n}, scale_k_n_{scale_k_n}, c_m_n_{c_m_n}, a_element_op_{a_op}, b_element_op_{b_op},
              c_element_op_{c_op} {}
        const Tensor<ADataType>& a_m_k_;
        const Tensor<BDataType>& b_k_n_;
        const Tensor<ScaleDataType>& scale_k_n_;
        Tensor<CDataType>& c_m_n_;
        AElementwiseOperation a_element_op_;
        BElementwiseOperation b_element_op_;
        CElementwiseOperation c_element_op_;
    };
    struct Invoker : public device::BaseInvoker
    {
        float Run(const Argument& arg)
        {
            auto f_mk_kn_mn = [&](auto m, auto n) {
                const int K = arg.a_m_k_.mDesc.GetLengths()[1];
                AccDataType v_acc = 0;
                for(int k = 0; k < K; ++k)
                {
                    ADataType v_a;
                    BDataType v_b;
                    ScaleDataType v_scale;
                    ADataType v_converted_b;
                    arg.a_element_op_(v_a, arg.a_m_k_(m, k));
                    arg.b_element_op_(v_b, arg.b_k_n_(k, n));
                    arg.b_element_op_(v_scale, arg.scale_k_n_(k, n));
                    v_converted_b = type_convert<ADataType>(v_b) * v_scale; // reference_fpAintB_gemm.hpp:110
                    v_acc += ck::type_convert<AccDataType>(v_a) * ck::type_convert<AccDataType>(v_converted_b);
                }
                AccDataType v_c;
                arg.c_element_op_(v_c, v_acc);
                arg.c_m_n_(m, n) = ck::type_convert<CDataType>(v_c);
            };
            make_ParallelTensorFunctor(f_mk_kn_mn, arg.c_m_n_.mDesc.GetLengths()[0], arg.c_m_n_.mDesc.GetLengths()[1])(
                std::thread::hardware_concurrency());
            return 0;
        }
        float Run(const device::BaseArgument* p_arg, const StreamConfig& = StreamConfig{}) override
        {
            return Run(*dynamic_cast<const Argument*>(p_arg));
        }
    };
    bool IsSupportedArgument(const device::BaseArgument*) override { return true; }
    static auto MakeArgument(const Tensor<ADataType>& a, const Tensor<BDataType>& b, const Tensor<ScaleDataType>& s,
                             Tensor<CDataType>& c, AElementwiseOperation ao, BElementwiseOperation bo, CElementwiseOperation co)
    {
        return Argument{a, b, s, c, ao, bo, co};
    }
    static auto MakeInvoker() { return Invoker{}; }
    std::string GetTypeString() const override { return "ReferencefpAintBGemm"; }
};

} // namespace host
} // namespace tensor_operation
} // namespace ck
