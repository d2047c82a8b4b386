// ckProfiler gemm_multi_abd -- bf16 A x (int8 B0, bf16 per-channel B1) with the bias / FastGelu epilogues of
// client_example/30_gemm_bf16Aint8B.  The reference has no profiler entry for this op (its clients loop over the factory
// themselves, gemm_bias_xdl_bf16_i8.cpp:118-215); this one follows the common profile_*_impl recipe and verifies with the
// host composition of example/60_gemm_multi_ABD (Multiply into bf16, ReferenceGemm, CDE functor).
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/tensor_operation_instance/gpu/gemm_multi_abd.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "profiler/profile_util.hpp"
#include "profiler_operation_registry.hpp"

#define OP_NAME "gemm_multi_abd"
#define OP_DESC "GEMM bf16 A x int8 B (Multiply scale) + bias/FastGelu epilogue"

namespace {
using BF16 = ck::bhalf_t; using I8 = int8_t; using F32 = float;
using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
using namespace ck::tensor_operation::element_wise;
using namespace ck::literals;

struct Args
{
    bool do_verification, do_log, time_kernel;
    int init_method, M, N, K, StrideA, StrideB, StrideE, n_warmup, n_iter;
};

template <typename BLayout, typename CDEOp, bool HasBias>
bool profile(const Args& p)
{
    constexpr bool b_row = std::is_same_v<BLayout, Row>;
    const std::size_t m = p.M, n = p.N, k = p.K;
    Tensor<BF16> a(HostTensorDescriptor({m, k}, {std::size_t(p.StrideA), 1_uz}));
    Tensor<I8> b0(b_row ? HostTensorDescriptor({k, n}, {std::size_t(p.StrideB), 1_uz}) : HostTensorDescriptor({k, n}, {1_uz, std::size_t(p.StrideB)}));
    Tensor<BF16> b1(HostTensorDescriptor({k, n}, {0_uz, 1_uz})), d(HostTensorDescriptor({m, n}, {0_uz, 1_uz}));
    Tensor<BF16> e_host(HostTensorDescriptor({m, n}, {std::size_t(p.StrideE), 1_uz})), e_dev(e_host.mDesc);
    if(p.init_method == 1)
    {
        a.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(float(std::rand() % 5 - 2)); });
        b0.GenerateTensorValue([](auto, auto) { return I8(std::rand() % 5 - 2); });
        b1.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(float(1 + std::rand() % 2)); });
        d.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(float(std::rand() % 5 - 2)); });
    }
    else if(p.init_method == 2)
    {
        a.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(float(std::rand()) / float(RAND_MAX) - 0.5f); });
        b0.GenerateTensorValue([](auto, auto) { return I8(std::rand() % 255 - 127); });
        b1.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(0.001f + 0.01f * float(std::rand()) / float(RAND_MAX)); });
        d.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(float(std::rand()) / float(RAND_MAX) - 0.5f); });
    }
    DeviceMem a_buf(2 * a.GetElementSpaceSize()), b0_buf(b0.GetElementSpaceSize()), b1_buf(2 * b1.GetElementSpaceSize()),
        d_buf(2 * d.GetElementSpaceSize()), e_buf(2 * e_dev.GetElementSpaceSize());
    a_buf.ToDevice(a.mData.data()); b0_buf.ToDevice(b0.mData.data()); b1_buf.ToDevice(b1.mData.data()); d_buf.ToDevice(d.mData.data());
    if(p.do_verification)
    {
        Tensor<BF16> b(HostTensorDescriptor({k, n}, {n, 1_uz}));
        for(int j = 0; j < p.N; ++j) for(int i = 0; i < p.K; ++i) Multiply{}(b(i, j), b0(i, j), b1(i, j));
        Tensor<F32> c({m, n});
        using Ref = ck::tensor_operation::host::ReferenceGemm<BF16, BF16, F32, F32, PassThrough, PassThrough, PassThrough>;
        auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a, b, c, PassThrough{}, PassThrough{}, PassThrough{});
        inv.Run(arg);
        for(int i = 0; i < p.M; ++i)
            for(int j = 0; j < p.N; ++j)
            {
                if constexpr(HasBias) CDEOp{}(e_host(i, j), c(i, j), d(i, j));
                else CDEOp{}(e_host(i, j), c(i, j));
            }
    }
    using DsLayout = std::conditional_t<HasBias, ck::Tuple<Row>, ck::Tuple<>>;
    using DsType   = std::conditional_t<HasBias, ck::Tuple<BF16>, ck::Tuple<>>;
    using DeviceOp = ck::tensor_operation::device::DeviceGemmMultipleABD<ck::Tuple<Row>, ck::Tuple<BLayout, BLayout>, DsLayout, Row, ck::Tuple<BF16>,
                                                                         ck::Tuple<I8, BF16>, DsType, BF16, PassThrough, Multiply, CDEOp>;
    auto op_ptrs = ck::tensor_operation::device::instance::DeviceOperationInstanceFactory<DeviceOp>::GetInstances();
    constexpr std::size_t ND = HasBias ? 1 : 0;
    std::array<const void*, ND> ds{}; std::array<ck::index_t, ND> sds{};
    if constexpr(HasBias) { ds[0] = d_buf.GetDeviceBuffer(); sds[0] = 0; }
    const double flop = 2.0 * p.M * p.N * p.K, bytes = 2.0 * p.M * p.K + 1.0 * p.K * p.N + 2.0 * p.M * p.N;
    return ck::profiler::run_instances(
        op_ptrs,
        [&](auto& op) {
            return op.MakeArgumentPointer({a_buf.GetDeviceBuffer()}, {b0_buf.GetDeviceBuffer(), b1_buf.GetDeviceBuffer()}, ds, e_buf.GetDeviceBuffer(),
                                          p.M, p.N, p.K, {p.StrideA}, {p.StrideB, 0}, sds, p.StrideE, PassThrough{}, Multiply{}, CDEOp{});
        },
        [&](auto&, auto*) { return 0; },
        [&]() {
            e_buf.FromDevice(e_dev.mData.data());
            if(p.do_log) LogRangeAsType<float>(std::cout << "e_device: ", e_dev.mData, ",") << std::endl;
            return ck::utils::check_err(e_dev, e_host, "Error: Incorrect results!", 1e-2, 1e-2);
        },
        p.do_verification, p.time_kernel, p.n_warmup, p.n_iter, flop, bytes, 1.0);
}
} // namespace

int profile_gemm_multi_abd(int argc, char* argv[])
{
    if(argc != 15 && argc != 17)
    {
        printf("arg1: tensor operation (" OP_NAME ": " OP_DESC ")\n");
        printf("arg2: data type (0: bf16 A, int8 B0, bf16 B1 scale, bf16 E)\n");
        printf("arg3: matrix layout (0: A[m, k] * B[k, n] = E[m, n]; 1: A[m, k] * B[n, k] = E[m, n])\n");
        printf("arg4: epilogue (0: none; 1: bias; 2: FastGelu; 3: bias + FastGelu)\n");
        printf("arg5: verification (0: no; 1: yes)\n");
        printf("arg6: initialization (0: no init; 1: integer value; 2: decimal value)\n");
        printf("arg7: print tensor value (0: no; 1: yes)\n");
        printf("arg8: time kernel (0=no, 1=yes)\n");
        printf("arg9 to 14: M, N, K, StrideA, StrideB, StrideE (-1 = packed)\n");
        printf("optional: arg15: warm-up cycles (default 1), arg16: iterations (default 10)\n");
        return 1;
    }
    const int layout = std::stoi(argv[3]), epi = std::stoi(argv[4]);
    Args p{};
    p.do_verification = std::stoi(argv[5]); p.init_method = std::stoi(argv[6]); p.do_log = std::stoi(argv[7]); p.time_kernel = std::stoi(argv[8]);
    p.M = std::stoi(argv[9]); p.N = std::stoi(argv[10]); p.K = std::stoi(argv[11]);
    auto dflt = [](int v, int d) { return v < 0 ? d : v; };
    p.StrideA = dflt(std::stoi(argv[12]), p.K); p.StrideB = dflt(std::stoi(argv[13]), layout == 0 ? p.N : p.K); p.StrideE = dflt(std::stoi(argv[14]), p.N);
    p.n_warmup = 1; p.n_iter = 10;
    if(argc == 17) { p.n_warmup = std::stoi(argv[15]); p.n_iter = std::stoi(argv[16]); }
    if(std::stoi(argv[2]) != 0 || layout < 0 || layout > 1 || epi < 0 || epi > 3)
    {
        std::cout << "this data_type & layout is not implemented" << std::endl;
        return 1;
    }
    bool pass = false;
    auto by_layout = [&](auto cde, auto has_bias) {
        using CDE = decltype(cde);
        return layout == 0 ? profile<Row, CDE, decltype(has_bias)::value>(p) : profile<Col, CDE, decltype(has_bias)::value>(p);
    };
    switch(epi)
    {
    case 0: pass = by_layout(PassThrough{}, std::false_type{}); break;
    case 1: pass = by_layout(Add{}, std::true_type{}); break;
    case 2: pass = by_layout(FastGelu{}, std::false_type{}); break;
    default: pass = by_layout(AddFastGelu{}, std::true_type{}); break;
    }
    return pass ? 0 : 1;
}

REGISTER_PROFILER_OPERATION(OP_NAME, OP_DESC, profile_gemm_multi_abd);
