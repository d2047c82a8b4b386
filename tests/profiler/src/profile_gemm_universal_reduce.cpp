// ckProfiler gemm_universal_reduce -- argument list of profiler/src/profile_gemm_universal_reduce.cpp (data type, layout, verify,
// init, log, time, M N K StrideA StrideB StrideC KBatch [warm-up iters]); data type 0 here = bf16 A x int8 B -> bf16, layout 0
// (mk_kn_mn), the instances script/profile_mixed_gemm.sh sweeps.  Host check: ReferenceGemm<bf16, int8, bf16, float>.
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/tensor_operation_instance/gpu/gemm_universal_reduce.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "profiler/profile_util.hpp"
#include "profiler_operation_registry.hpp"

int profile_gemm_universal_reduce(int argc, char* argv[])
{
    if(argc != 15 && argc != 17)
    {
        printf("arg1: tensor operation (gemm_universal_reduce)\narg2: data type (0: bf16 A, int8 B, bf16 C)\n"
               "arg3: matrix layout (0: A[m, k] * B[k, n] = C[m, n])\narg4: verification (0: no; 1: yes)\n"
               "arg5: initialization (0: no init; 1: integer value; 2: decimal value)\narg6: print tensor value (0: no; 1: yes)\n"
               "arg7: time kernel (0=no, 1=yes)\narg8 to 13: M, N, K, StrideA, StrideB, StrideC (-1 = packed)\narg14: split k into mulitiple batch\n"
               "optional: arg15: warm-up cycles (default 1), arg16: iterations (default 10)\n");
        return 1;
    }
    using BF16 = ck::bhalf_t; using I8 = int8_t; using F32 = float;
    using Row = ck::tensor_layout::gemm::RowMajor;
    using namespace ck::tensor_operation::element_wise;
    using namespace ck::literals;
    const int data_type = std::stoi(argv[2]), layout = std::stoi(argv[3]);
    const bool do_verification = std::stoi(argv[4]); const int init_method = std::stoi(argv[5]);
    const bool do_log = std::stoi(argv[6]), time_kernel = std::stoi(argv[7]);
    const int M = std::stoi(argv[8]), N = std::stoi(argv[9]), K = std::stoi(argv[10]);
    auto dflt = [](int v, int d) { return v < 0 ? d : v; };
    const int StrideA = dflt(std::stoi(argv[11]), K), StrideB = dflt(std::stoi(argv[12]), N), StrideC = dflt(std::stoi(argv[13]), N);
    const int KBatch = std::stoi(argv[14]);
    const int n_warmup = argc == 17 ? std::stoi(argv[15]) : 1, n_iter = argc == 17 ? std::stoi(argv[16]) : 10;
    if(data_type != 0 || layout != 0)
    {
        std::cout << "this data_type & layout is not implemented" << std::endl;
        return 1;
    }
    const std::size_t m = M, n = N, k = K;
    Tensor<BF16> a(HostTensorDescriptor({m, k}, {std::size_t(StrideA), 1_uz}));
    Tensor<I8> b(HostTensorDescriptor({k, n}, {std::size_t(StrideB), 1_uz}));
    Tensor<BF16> c_host(HostTensorDescriptor({m, n}, {std::size_t(StrideC), 1_uz})), c_dev(c_host.mDesc);
    const bool ints = init_method == 1;
    if(init_method)
    {
        a.GenerateTensorValue([&](auto, auto) { return ck::type_convert<BF16>(ints ? float(std::rand() % 5 - 2) : float(std::rand()) / float(RAND_MAX) - 0.5f); });
        b.GenerateTensorValue([&](auto, auto) { return I8(ints ? std::rand() % 5 - 2 : std::rand() % 5 - 2); }); // unscaled int8: keep |c| moderate
    }
    DeviceMem a_buf(2 * a.GetElementSpaceSize()), b_buf(b.GetElementSpaceSize()), c_buf(2 * c_dev.GetElementSpaceSize());
    a_buf.ToDevice(a.mData.data()); b_buf.ToDevice(b.mData.data());
    if(do_verification)
    {
        using Ref = ck::tensor_operation::host::ReferenceGemm<BF16, I8, BF16, F32, PassThrough, PassThrough, PassThrough, BF16, BF16>;
        auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a, b, c_host, PassThrough{}, PassThrough{}, PassThrough{});
        inv.Run(arg);
    }
    using DeviceOp = ck::tensor_operation::device::DeviceGemmV2R1<Row, Row, ck::Tuple<>, Row, BF16, I8, ck::Tuple<>, BF16, PassThrough, PassThrough, PassThrough>;
    auto op_ptrs = ck::tensor_operation::device::instance::DeviceOperationInstanceFactory<DeviceOp>::GetInstances();
    const bool pass = ck::profiler::run_instances(
        op_ptrs,
        [&](auto& op) {
            return op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), std::array<const void*, 0>{}, c_buf.GetDeviceBuffer(), M, N, K,
                                          StrideA, StrideB, std::array<ck::index_t, 0>{}, StrideC, KBatch, PassThrough{}, PassThrough{}, PassThrough{});
        },
        [&](auto& op, auto* arg) {
            auto ws = std::make_shared<DeviceMem>(op.GetWorkSpaceSize(arg));
            op.SetWorkSpacePointer(arg, ws->GetDeviceBuffer());
            return ws;
        },
        [&]() {
            c_buf.FromDevice(c_dev.mData.data());
            if(do_log) LogRangeAsType<float>(std::cout << "c_device: ", c_dev.mData, ",") << std::endl;
            return ck::utils::check_err(c_dev, c_host, "Error: Incorrect results!", 1e-2, 1e-2);
        },
        do_verification, time_kernel, n_warmup, n_iter, 2.0 * M * N * K, 2.0 * M * K + 1.0 * K * N + 2.0 * M * N, 1.0);
    return pass ? 0 : 1;
}

REGISTER_PROFILER_OPERATION("gemm_universal_reduce", "Universal GEMM, split-K + reduce (bf16 A x int8 B)", profile_gemm_universal_reduce);
