// ORACLE (test infrastructure; never included by product code): CK-style host ReferenceGemmMultipleD over Tensor<T>, restating
// library/include/ck/library/reference_tensor_operation/cpu/reference_gemm_multiple_d.hpp:28-175 -- same template parameters and
// MakeArgument / MakeInvoker / Run surface; k-ascending accumulation in AccDataType of the a_op / b_op images (ConvertBF16RTN
// bypassed exactly as upstream, :82-101), then ONE call of the CDE functor with the D values of (m, n), any number of Ds
// (upstream spells out 0, 1 and 2).  All D tensors share the first D's element type, as upstream (:30).
#pragma once
#include <array>
#include <thread>
#include <utility>

#include "ck/library/utility/host_tensor.hpp"
#include "ck/tensor_operation/gpu/device/device_base.hpp"
#include "ck/tensor_operation/gpu/element/element_wise_operation.hpp"

namespace ck {
namespace tensor_operation {
namespace host {

namespace detail {
template <typename Ds, bool Empty = (Ds::Size() == 0)>
struct first_d { using type = float; };
template <typename Ds>
struct first_d<Ds, false> { using type = tuple_element_t<0, Ds>; };
} // namespace detail

template <typename ADataType, typename BDataType, typename DsDataType, typename CDataType, typename AccDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CDEElementwiseOperation,
          typename ComputeTypeA = CDataType, typename ComputeTypeB = ComputeTypeA>
struct ReferenceGemmMultipleD : public device::BaseOperator
{
    using DDataType              = typename detail::first_d<DsDataType>::type;
    static constexpr size_t NumD = size_t(DsDataType::Size());
    using DsArray                = std::array<Tensor<DDataType>, NumD>;

    struct Argument : public device::BaseArgument
    {
        Argument(const Tensor<ADataType>& a_m_k, const Tensor<BDataType>& b_k_n, const DsArray& ds_m_n, Tensor<CDataType>& c_m_n,
                 AElementwiseOperation a_op, BElementwiseOperation b_op, CDEElementwiseOperation cde_op)
            : a_m_k_{a_m_k}, b_k_n_{b_k_n}, ds_m_n_{ds_m_n}, c_m_n_{c_m_n}, a_element_op_{a_op}, b_element_op_{b_op}, cde_element_op_{cde_op} {}
        const Tensor<ADataType>& a_m_k_;
        const Tensor<BDataType>& b_k_n_;
        const DsArray& ds_m_n_;
        Tensor<CDataType>& c_m_n_;
        AElementwiseOperation a_element_op_;
        BElementwiseOperation b_element_op_;
        CDEElementwiseOperation cde_element_op_;
    };

    struct Invoker : public device::BaseInvoker
    {
        template <size_t... I>
        static void apply_cde(const Argument& arg, CDataType& c, const AccDataType& acc, size_t m, size_t n, std::index_sequence<I...>)
        {
            arg.cde_element_op_(c, acc, arg.ds_m_n_[I](m, n)...);
        }
        float Run(const Argument& arg)
        {
            const int K     = arg.a_m_k_.mDesc.GetLengths()[1];
            auto one_output = [&](auto m, auto n) {
                AccDataType acc = 0;
                ComputeTypeA va = 0;
                ComputeTypeB vb = 0;
                for(int k = 0; k < K; ++k)
                {
                    if constexpr(is_same_v<AElementwiseOperation, element_wise::ConvertBF16RTN>) element_wise::PassThrough{}(va, arg.a_m_k_(m, k));
                    else arg.a_element_op_(va, arg.a_m_k_(m, k));
                    if constexpr(is_same_v<BElementwiseOperation, element_wise::ConvertBF16RTN>) element_wise::PassThrough{}(vb, arg.b_k_n_(k, n));
                    else arg.b_element_op_(vb, arg.b_k_n_(k, n));
                    acc += ck::type_convert<AccDataType>(va) * ck::type_convert<AccDataType>(vb);
                }
                CDataType c = 0;
                apply_cde(arg, c, acc, m, n, std::make_index_sequence<NumD>{});
                arg.c_m_n_(m, n) = c;
            };
            make_ParallelTensorFunctor(one_output, arg.c_m_n_.mDesc.GetLengths()[0], arg.c_m_n_.mDesc.GetLengths()[1])(
                std::thread::hardware_concurrency());
            return 0;
        }
        float Run(const device::BaseArgument* p_arg, const StreamConfig& = StreamConfig{}) override
        {
            return Run(*dynamic_cast<const Argument*>(p_arg));
        }
    };

    static constexpr bool IsValidCompilationParameter() { return true; }
    bool IsSupportedArgument(const device::BaseArgument*) override { return true; }
    static auto MakeArgument(const Tensor<ADataType>& a, const Tensor<BDataType>& b, const DsArray& ds, Tensor<CDataType>& c,
                             AElementwiseOperation ao, BElementwiseOperation bo, CDEElementwiseOperation co)
    {
        return Argument{a, b, ds, c, ao, bo, co};
    }
    static auto MakeInvoker() { return Invoker{}; }
    virtual std::unique_ptr<device::BaseInvoker> MakeInvokerPointer() { return std::make_unique<Invoker>(Invoker{}); }
    std::string GetTypeString() const override { return "ReferenceGemmMultipleD"; }
};

} // namespace host
} // namespace tensor_operation
} // namespace ck
