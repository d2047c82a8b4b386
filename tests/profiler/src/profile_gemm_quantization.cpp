// ckProfiler gemm_quantization -- int8 x int8 GEMM with the requantisation epilogues of example/14_gemm_quantization and
// client_example/09_quantization/gemm_quantization.cpp: 0 = Activation_Mul_Clamp<PassThrough>, 1 = Add_Activation_Mul_Clamp<Relu>
// (int32 bias).  Verification is exact (atol = 0): ReferenceGemm<int8,int8,int32,int32> then the functor on the host.
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/tensor_operation_instance/gpu/quantization/gemm_quantization.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "ck/library/utility/host_tensor_generator.hpp"
#include "profiler/profile_util.hpp"
#include "profiler_operation_registry.hpp"

int profile_gemm_quantization(int argc, char* argv[])
{
    if(argc != 11 && argc != 13)
    {
        printf("arg1: tensor operation (gemm_quantization)\narg2: epilogue (0: mul + clamp; 1: bias + relu + mul + clamp)\n"
               "arg3: verification (0: no; 1: yes)\narg4: initialization (1: small integers; 2: full int8 range)\n"
               "arg5: print tensor value (0: no; 1: yes)\narg6: time kernel (0=no, 1=yes)\narg7 to 9: M, N, K\narg10: requant scale (e.g. 0.03)\n"
               "optional: arg11: warm-up cycles (default 1), arg12: iterations (default 10)\n");
        return 1;
    }
    using I8 = int8_t; using I32 = int32_t;
    using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
    using namespace ck::tensor_operation::element_wise;
    using namespace ck::literals;
    const int epi = std::stoi(argv[2]);
    const bool do_verification = std::stoi(argv[3]); const int init_method = std::stoi(argv[4]);
    const bool do_log = std::stoi(argv[5]), time_kernel = std::stoi(argv[6]);
    const int M = std::stoi(argv[7]), N = std::stoi(argv[8]), K = std::stoi(argv[9]);
    const float requant = std::stof(argv[10]);
    const int n_warmup = argc == 13 ? std::stoi(argv[11]) : 1, n_iter = argc == 13 ? std::stoi(argv[12]) : 10;
    const std::size_t m = M, n = N, k = K;
    Tensor<I8> a(HostTensorDescriptor({m, k}, {k, 1_uz})), b(HostTensorDescriptor({k, n}, {1_uz, k}));
    Tensor<I32> bias(HostTensorDescriptor({m, n}, {0_uz, 1_uz})), c({m, n});
    Tensor<I8> e_host({m, n}), e_dev({m, n});
    const int r = init_method == 1 ? 5 : 127;
    a.GenerateTensorValue(GeneratorTensor_2<I8>{-r, r}); b.GenerateTensorValue(GeneratorTensor_2<I8>{-r, r});
    bias.GenerateTensorValue(GeneratorTensor_2<I32>{-5000, 5000});
    DeviceMem a_buf(a.GetElementSpaceSize()), b_buf(b.GetElementSpaceSize()), bias_buf(4 * bias.GetElementSpaceSize()), e_buf(e_dev.GetElementSpaceSize());
    a_buf.ToDevice(a.mData.data()); b_buf.ToDevice(b.mData.data()); bias_buf.ToDevice(bias.mData.data());
    if(do_verification)
    {
        using Ref = ck::tensor_operation::host::ReferenceGemm<I8, I8, I32, I32, PassThrough, PassThrough, PassThrough>;
        auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a, b, c, PassThrough{}, PassThrough{}, PassThrough{});
        inv.Run(arg);
    }
    const double flop = 2.0 * M * N * K, bytes = 1.0 * M * K + 1.0 * K * N + 1.0 * M * N;
    auto verify = [&]() {
        e_buf.FromDevice(e_dev.mData.data());
        if(do_log) LogRangeAsType<int>(std::cout << "e_device: ", e_dev.mData, ",") << std::endl;
        return ck::utils::check_err(e_dev, e_host, "Error: Incorrect results!", 0, 0);
    };
    bool pass = false;
    if(epi == 0)
    {
        auto cde = Activation_Mul_Clamp<PassThrough>{requant, PassThrough{}};
        if(do_verification) for(int i = 0; i < M; ++i) for(int j = 0; j < N; ++j) cde(e_host(i, j), c(i, j));
        using DeviceOp = ck::tensor_operation::device::DeviceGemmMultipleD<Row, Col, ck::Tuple<>, Row, I8, I8, ck::Tuple<>, I8, PassThrough, PassThrough, decltype(cde)>;
        auto op_ptrs = ck::tensor_operation::device::instance::DeviceOperationInstanceFactory<DeviceOp>::GetInstances();
        pass = ck::profiler::run_instances(
            op_ptrs,
            [&](auto& op) {
                return op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), std::array<const void*, 0>{}, e_buf.GetDeviceBuffer(), M, N, K, K, K,
                                              std::array<ck::index_t, 0>{}, N, PassThrough{}, PassThrough{}, cde);
            },
            [&](auto&, auto*) { return 0; }, verify, do_verification, time_kernel, n_warmup, n_iter, flop, bytes, 2.0);
    }
    else
    {
        auto cde = Add_Activation_Mul_Clamp<Relu>{requant, Relu{}};
        if(do_verification) for(int i = 0; i < M; ++i) for(int j = 0; j < N; ++j) cde(e_host(i, j), c(i, j), bias(i, j));
        using DeviceOp = ck::tensor_operation::device::DeviceGemmMultipleD<Row, Col, ck::Tuple<Row>, Row, I8, I8, ck::Tuple<I32>, I8, PassThrough, PassThrough, decltype(cde)>;
        auto op_ptrs = ck::tensor_operation::device::instance::DeviceOperationInstanceFactory<DeviceOp>::GetInstances();
        pass = ck::profiler::run_instances(
            op_ptrs,
            [&](auto& op) {
                return op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), std::array<const void*, 1>{bias_buf.GetDeviceBuffer()},
                                              e_buf.GetDeviceBuffer(), M, N, K, K, K, std::array<ck::index_t, 1>{0}, N, PassThrough{}, PassThrough{}, cde);
            },
            [&](auto&, auto*) { return 0; }, verify, do_verification, time_kernel, n_warmup, n_iter, flop, bytes + 4.0 * N, 2.0);
    }
    return pass ? 0 : 1;
}

REGISTER_PROFILER_OPERATION("gemm_quantization", "GEMM int8 x int8 with requantisation epilogue (example 14)", profile_gemm_quantization);
