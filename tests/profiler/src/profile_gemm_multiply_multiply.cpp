// ckProfiler gemm_multiply_multiply -- argument list of profiler/src/profile_gemm_multiply_multiply.cpp:35-88, instance
// loop of profile_gemm_multiply_multiply_impl.hpp:39-326.  The reference wires fp8 -> bf16 only; this build serves the
// [EXT] int8 x int8 -> bf16 instances (data type 3), the cfg4 workload: E = bf16( int32 acc * D0[n] * D1[m] ).
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/tensor_operation_instance/gpu/gemm_multiply_multiply.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "ck/library/utility/host_tensor_generator.hpp"
#include "profiler/profile_util.hpp"
#include "profiler_operation_registry.hpp"

#define OP_NAME "gemm_multiply_multiply"
#define OP_DESC "GEMM_Multiply_Multiply"

int profile_gemm_multiply_multiply(int argc, char* argv[])
{
    if(argc != 16 && argc != 19)
    {
        printf("arg1: tensor operation (" OP_NAME ": " OP_DESC ")\n");
        printf("arg2: data type (3: int8 x int8 -> bf16, int32 accumulate, fp32 row/col scales; others: not built for sm_100a)\n");
        printf("arg3: matrix layout (1: A[m, k] * B[n, k] = C[m, n]; others: not built)\n");
        printf("arg4: verification (0: no; 1: yes)\n");
        printf("arg5: initialization (0: no init; 1: integer value; 2: decimal value)\n");
        printf("arg6: print tensor value (0: no; 1: yes)\n");
        printf("arg7: time kernel (0=no, 1=yes)\n");
        printf("arg8 to 15: M, N, K, StrideA, StrideB, StrideD0, StrideD1, StrideE\n");
        printf("optional:\narg16: number of warm-up cycles (default 1)\narg17: number of iterations (default 10)\n");
        printf("arg18: memory for rotating buffer (MB; > 0: cold-cache timing with rotating A/B copies and an L2 flush before each launch)\n");
        return 1;
    }
    const int data_type = std::stoi(argv[2]), layout = std::stoi(argv[3]);
    const bool do_verification = std::stoi(argv[4]);
    const int init_method      = std::stoi(argv[5]);
    const bool do_log = std::stoi(argv[6]), time_kernel = std::stoi(argv[7]);
    const int M = std::stoi(argv[8]), N = std::stoi(argv[9]), K = std::stoi(argv[10]);
    auto dflt = [](int v, int d) { return v < 0 ? d : v; }; // -1 = default stride, as the reference's f_get_default_stride
    const int StrideA = dflt(std::stoi(argv[11]), K), StrideB = dflt(std::stoi(argv[12]), K);
    const int StrideD0 = std::max(0, std::stoi(argv[13])), StrideD1 = std::max(0, std::stoi(argv[14]));
    const int StrideE = dflt(std::stoi(argv[15]), N);
    int n_warmup = 1, n_iter = 10;
    if(argc == 19) { n_warmup = std::stoi(argv[16]); n_iter = std::stoi(argv[17]); ck::profiler::options().rotating_mb = std::stoi(argv[18]); }
    if(data_type != 3 || layout != 1)
    {
        std::cout << "this data_type & layout is not implemented" << std::endl;
        return 1;
    }
    using I8 = int8_t; using I32 = int32_t; using F32 = float; using BF16 = ck::bhalf_t;
    using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
    using namespace ck::tensor_operation::element_wise;
    using namespace ck::literals;
    const std::size_t m = M, n = N, k = K;
    Tensor<I8> a(HostTensorDescriptor({m, k}, {std::size_t(StrideA), 1_uz})), b(HostTensorDescriptor({k, n}, {1_uz, std::size_t(StrideB)}));
    Tensor<F32> d0(HostTensorDescriptor({m, n}, {std::size_t(StrideD0), 1_uz})); // Row, stride 0: one scale per column
    Tensor<F32> d1(HostTensorDescriptor({m, n}, {1_uz, std::size_t(StrideD1)})); // Col, stride 0: one scale per row
    Tensor<BF16> e_host(HostTensorDescriptor({m, n}, {std::size_t(StrideE), 1_uz})), e_dev(e_host.mDesc);
    std::cout << "a_m_k: " << a.mDesc << "\nb_k_n: " << b.mDesc << "\ne_m_n: " << e_host.mDesc << std::endl;
    if(init_method == 1)
    {
        a.GenerateTensorValue(GeneratorTensor_2<I8>{-5, 5}); b.GenerateTensorValue(GeneratorTensor_2<I8>{-5, 5});
        d0.GenerateTensorValue([](auto, auto) { return float(1 + std::rand() % 3); });
        d1.GenerateTensorValue([](auto, auto) { return float(1 + std::rand() % 3); });
    }
    else if(init_method == 2)
    {
        a.GenerateTensorValue(GeneratorTensor_2<I8>{-128, 127}); b.GenerateTensorValue(GeneratorTensor_2<I8>{-128, 127});
        d0.GenerateTensorValue([](auto, auto) { return 5e-4f + 1e-3f * float(std::rand()) / float(RAND_MAX); });
        d1.GenerateTensorValue([](auto, auto) { return 5e-4f + 1e-3f * float(std::rand()) / float(RAND_MAX); });
    }
    DeviceMem a_buf(a.GetElementSpaceSize()), b_buf(b.GetElementSpaceSize()), d0_buf(4 * d0.GetElementSpaceSize()),
        d1_buf(4 * d1.GetElementSpaceSize()), e_buf(2 * e_dev.GetElementSpaceSize());
    a_buf.ToDevice(a.mData.data()); b_buf.ToDevice(b.mData.data()); d0_buf.ToDevice(d0.mData.data()); d1_buf.ToDevice(d1.mData.data());

    if(do_verification)
    {
        Tensor<I32> c({m, n});
        using Ref = ck::tensor_operation::host::ReferenceGemm<I8, I8, I32, I32, PassThrough, PassThrough, PassThrough>;
        auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a, b, c, PassThrough{}, PassThrough{}, PassThrough{});
        inv.Run(arg);
        for(int i = 0; i < M; ++i) for(int j = 0; j < N; ++j) MultiplyMultiply{}(e_host(i, j), c(i, j), d0(i, j), d1(i, j));
    }
    using DeviceOp = ck::tensor_operation::device::DeviceGemmMultipleD<Row, Col, ck::Tuple<Row, Col>, Row, I8, I8, ck::Tuple<F32, F32>, BF16,
                                                                       PassThrough, PassThrough, MultiplyMultiply>;
    auto op_ptrs = ck::tensor_operation::device::instance::DeviceOperationInstanceFactory<DeviceOp>::GetInstances();
    const double flop = 2.0 * M * N * K, bytes = 1.0 * M * K + 1.0 * K * N + 2.0 * M * N;
    const bool pass = ck::profiler::run_instances(
        op_ptrs,
        [&](auto& op) {
            return op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(),
                                          std::array<const void*, 2>{d0_buf.GetDeviceBuffer(), d1_buf.GetDeviceBuffer()}, e_buf.GetDeviceBuffer(),
                                          M, N, K, StrideA, StrideB, std::array<ck::index_t, 2>{StrideD0, StrideD1}, StrideE, PassThrough{},
                                          PassThrough{}, MultiplyMultiply{});
        },
        [&](auto&, auto*) { return 0; },
        [&]() {
            e_buf.FromDevice(e_dev.mData.data());
            if(do_log) LogRangeAsType<float>(std::cout << "e_device: ", e_dev.mData, ",") << std::endl;
            return ck::utils::check_err(e_dev, e_host, "Error: Incorrect results!", 0, 0); // int32 accumulators exact -> same bf16 bits
        },
        do_verification, time_kernel, n_warmup, n_iter, flop, bytes, 2.0);
    return pass ? 0 : 1;
}

REGISTER_PROFILER_OPERATION(OP_NAME, OP_DESC, profile_gemm_multiply_multiply);
