// Port of example/14_gemm_quantization (gemm_xdl_quantization_int8.cpp and gemm_xdl_bias_relu_quantization_int8.cpp) plus the
// [EXT] int8 MultiplyMultiply composition of cfg4 (example/65 epilogue on int32 accumulators).  Verification follows
// gemm_xdl_bias_relu_quantization_int8.cpp:97-103,213-229: ReferenceGemm<int8,int8,int32,int32> then the CDE functor on the
// host; integers are compared exactly (check_err atol = 0), so the int32 accumulators are proven bit-exact.
// usage: example_gemm_quantization_int8 [mode: 0 mul_clamp | 1 bias_relu_mul_clamp | 2 multiply_multiply | 3 raw int32] [M N K]
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/tensor_operation_instance/gpu/gemm_quantization.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "ck/library/utility/host_tensor_generator.hpp"

using I8 = int8_t; using I32 = int32_t; using F32 = float; using BF16 = ck::bhalf_t;
using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
using namespace ck::tensor_operation::element_wise;
using namespace ck::literals;

struct Problem
{
    ck::index_t M, N, K;
    Tensor<I8> a, b; Tensor<I32> c;
    Problem(ck::index_t m, ck::index_t n, ck::index_t k)
        : M(m), N(n), K(k), a(HostTensorDescriptor({std::size_t(m), std::size_t(k)}, {std::size_t(k), 1_uz})),
          b(HostTensorDescriptor({std::size_t(k), std::size_t(n)}, {1_uz, std::size_t(k)})), c({std::size_t(m), std::size_t(n)})
    {
        a.GenerateTensorValue(GeneratorTensor_2<I8>{-128, 127});
        b.GenerateTensorValue(GeneratorTensor_2<I8>{-128, 127});
        using Ref = ck::tensor_operation::host::ReferenceGemm<I8, I8, I32, I32, PassThrough, PassThrough, PassThrough>;
        auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a, b, c, PassThrough{}, PassThrough{}, PassThrough{});
        inv.Run(arg);
    }
};

template <typename DsLayout, typename DsDataType, typename EDataType, typename CDEOp, std::size_t ND, typename HostF>
int run(Problem& p, std::array<const void*, ND> ds, std::array<ck::index_t, ND> strides, CDEOp cde, HostF host_f, const char* name)
{
    Tensor<EDataType> e_host({std::size_t(p.M), std::size_t(p.N)}), e_dev({std::size_t(p.M), std::size_t(p.N)});
    DeviceMem a_buf(p.a.GetElementSpaceSize()), b_buf(p.b.GetElementSpaceSize()), e_buf(sizeof(EDataType) * e_dev.GetElementSpaceSize());
    a_buf.ToDevice(p.a.mData.data()); b_buf.ToDevice(p.b.mData.data());
    using DeviceOp = ck::tensor_operation::device::DeviceGemmMultipleD<Row, Col, DsLayout, Row, I8, I8, DsDataType, EDataType, PassThrough, PassThrough, CDEOp>;
    auto ops = ck::tensor_operation::device::instance::DeviceOperationInstanceFactory<DeviceOp>::GetInstances();
    auto& op = ops.at(0);
    auto arg = op->MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), ds, e_buf.GetDeviceBuffer(), p.M, p.N, p.K, p.K, p.K,
                                       strides, p.N, PassThrough{}, PassThrough{}, cde);
    if(!op->IsSupportedArgument(arg.get())) { std::cerr << name << ": unsupported" << std::endl; return 2; }
    const float ms = op->MakeInvokerPointer()->Run(arg.get(), StreamConfig{nullptr, true, 0, 2, 5});
    std::cout << "Perf: " << ms << " ms, " << 2.0 * p.M * p.N * p.K / 1.E9 / ms << " TOps, " << name << ", " << op->GetTypeString() << std::endl;
    e_buf.FromDevice(e_dev.mData.data());
    for(int m = 0; m < p.M; ++m) for(int n = 0; n < p.N; ++n) host_f(e_host(m, n), p.c(m, n), m, n);
    const bool ok = ck::utils::check_err(e_dev, e_host, "Error: Incorrect results!", 0, 0); // exact, also for the bf16 bit patterns
    std::cout << name << (ok ? ": verification PASSED (bit-exact)" : ": verification FAILED") << std::endl;
    return ok ? 0 : 1;
}

int main(int argc, char* argv[])
{
    int mode = 0; ck::index_t M = 1024, N = 1024, K = 1024;
    if(argc >= 2) mode = std::atoi(argv[1]);
    if(argc >= 5) { M = std::atoi(argv[2]); N = std::atoi(argv[3]); K = std::atoi(argv[4]); }
    Problem p(M, N, K);
    const float requant_scale = 0.03f;
    if(mode == 0)
    {
        auto cde = Activation_Mul_Clamp<PassThrough>{requant_scale, PassThrough{}};
        return run<ck::Tuple<>, ck::Tuple<>, I8>(p, std::array<const void*, 0>{}, std::array<ck::index_t, 0>{}, cde,
                                                 [&](I8& e, const I32& c, int, int) { cde(e, c); }, "int8 gemm + mul + clamp");
    }
    if(mode == 1)
    {
        Tensor<I32> bias(HostTensorDescriptor({std::size_t(M), std::size_t(N)}, {0_uz, 1_uz}));
        bias.GenerateTensorValue(GeneratorTensor_2<I32>{-5000, 5000});
        DeviceMem bias_buf(4 * bias.GetElementSpaceSize()); bias_buf.ToDevice(bias.mData.data());
        auto cde = Add_Activation_Mul_Clamp<Relu>{requant_scale, Relu{}};
        return run<ck::Tuple<Row>, ck::Tuple<I32>, I8>(p, std::array<const void*, 1>{bias_buf.GetDeviceBuffer()}, std::array<ck::index_t, 1>{0}, cde,
                                                       [&](I8& e, const I32& c, int m, int n) { cde(e, c, bias(m, n)); }, "int8 gemm + bias + relu + mul + clamp");
    }
    if(mode == 2)
    {
        Tensor<F32> d0(HostTensorDescriptor({std::size_t(M), std::size_t(N)}, {0_uz, 1_uz})), d1(HostTensorDescriptor({std::size_t(M), std::size_t(N)}, {1_uz, 0_uz}));
        d0.GenerateTensorValue([](auto, auto) { return 5e-4f + 1e-3f * float(std::rand()) / float(RAND_MAX); });
        d1.GenerateTensorValue([](auto, auto) { return 5e-4f + 1e-3f * float(std::rand()) / float(RAND_MAX); });
        DeviceMem d0_buf(4 * d0.GetElementSpaceSize()), d1_buf(4 * d1.GetElementSpaceSize());
        d0_buf.ToDevice(d0.mData.data()); d1_buf.ToDevice(d1.mData.data());
        auto cde = MultiplyMultiply{};
        return run<ck::Tuple<Row, Col>, ck::Tuple<F32, F32>, BF16>(p, std::array<const void*, 2>{d0_buf.GetDeviceBuffer(), d1_buf.GetDeviceBuffer()},
                                                                   std::array<ck::index_t, 2>{0, 0}, cde,
                                                                   [&](BF16& e, const I32& c, int m, int n) { cde(e, c, d0(m, n), d1(m, n)); },
                                                                   "int8 gemm multiply_multiply (row/col scales) -> bf16");
    }
    auto cde = PassThrough{};
    return run<ck::Tuple<>, ck::Tuple<>, I32>(p, std::array<const void*, 0>{}, std::array<ck::index_t, 0>{}, cde,
                                              [&](I32& e, const I32& c, int, int) { e = c; }, "int8 gemm raw int32 accumulators");
}
