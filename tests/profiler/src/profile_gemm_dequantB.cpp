// ckProfiler fpAintB_gemm  -- example/64_fpAintB_gemm as a profiler op: fp16 A x biased-uint8 B x fp16 per-channel scale
//                             (DeviceFpAintBGemm_Wmma_CShuffle), verified with ReferencefpAintBGemm.
// ckProfiler gemm_w4a16    -- [EXT] bf16 A x packed int4 B, group-128 bf16 scales (+ optional uint8 zero points)
//                             (DeviceGemm_dequantB_GroupScale_Sm100), the Llama decode / prefill shapes of BASELINE.json.
// Neither op has a profiler entry in the reference; both follow the common profile_*_impl recipe.
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_fpAintB_gemm.hpp"
#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/fill.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_fpAintB_gemm_wmma.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_sm100.hpp"
#include "profiler/profile_util.hpp"
#include "profiler_operation_registry.hpp"

namespace {
using F16 = ck::half_t; using BF16 = ck::bhalf_t; using F32 = float; using I8 = int8_t; using U8 = uint8_t;
using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
using namespace ck::tensor_operation::element_wise;
using namespace ck::literals;

int usage(const char* name, const char* extra)
{
    printf("arg1: tensor operation (%s)\narg2: verification (0: no; 1: yes)\narg3: initialization (1: integer value; 2: decimal value)\n"
           "arg4: print tensor value (0: no; 1: yes)\narg5: time kernel (0=no, 1=yes)\narg6 to 8: M, N, K\n%s"
           "optional (after the above): warm-up cycles (default 1), iterations (default 10)\n", name, extra);
    return 1;
}
} // namespace

int profile_fpAintB_gemm(int argc, char* argv[])
{
    if(argc != 9 && argc != 11) return usage("fpAintB_gemm", "");
    const bool do_verification = std::stoi(argv[2]); const int init_method = std::stoi(argv[3]);
    const bool do_log = std::stoi(argv[4]), time_kernel = std::stoi(argv[5]);
    const int M = std::stoi(argv[6]), N = std::stoi(argv[7]), K = std::stoi(argv[8]);
    const int n_warmup = argc == 11 ? std::stoi(argv[9]) : 1, n_iter = argc == 11 ? std::stoi(argv[10]) : 10;
    const std::size_t m = M, n = N, k = K;
    Tensor<F16> a(HostTensorDescriptor({m, k}, {k, 1_uz})), s(HostTensorDescriptor({k, n}, {0_uz, 1_uz}));
    Tensor<I8> q(HostTensorDescriptor({k, n}, {1_uz, k}));
    Tensor<F16> c_host(HostTensorDescriptor({m, n}, {n, 1_uz})), c_dev(c_host.mDesc);
    if(init_method == 1)
    {
        ck::utils::FillUniformDistributionIntegerValue<F16>{-1.f, 1.f}(a);
        ck::utils::FillUniformDistributionIntegerValue<I8>{-1.f, 1.f}(q);
        ck::utils::FillUniformDistributionIntegerValue<F16>{-1.f, 1.f}(s);
    }
    else
    {
        ck::utils::FillUniformDistribution<F16>{-1.f, 1.f}(a);
        ck::utils::FillUniformDistribution<I8>{-127.f, 127.f}(q);
        ck::utils::FillUniformDistribution<F16>{0.001f, 0.02f}(s);
    }
    Tensor<U8> b(q.mDesc); // UnsignedWeightPreprocessor (example/64_fpAintB_gemm/common.hpp:56-83)
    for(std::size_t i = 0; i < b.mData.size(); ++i) b.mData[i] = U8(int(q.mData[i]) + 128);
    DeviceMem a_buf(2 * a.GetElementSpaceSize()), b_buf(b.GetElementSpaceSize()), s_buf(2 * s.GetElementSpaceSize()), c_buf(2 * c_dev.GetElementSpaceSize());
    a_buf.ToDevice(a.mData.data()); b_buf.ToDevice(b.mData.data()); s_buf.ToDevice(s.mData.data());
    if(do_verification)
    {
        using Ref = ck::tensor_operation::host::ReferencefpAintBGemm<F16, I8, F16, F16, F32, PassThrough, PassThrough, PassThrough>;
        auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a, q, s, c_host, PassThrough{}, PassThrough{}, PassThrough{});
        inv.Run(arg);
    }
    using Op = ck::tensor_operation::device::DeviceFpAintBGemm_Wmma_CShuffle<Row, Col, Row, F16, U8, F16, F16, F32, F16, PassThrough, PassThrough, PassThrough>;
    std::vector<std::unique_ptr<Op>> op_ptrs;
    op_ptrs.push_back(std::make_unique<Op>());
    const double tol = init_method == 1 ? 1e-3 : 1e-2;
    const bool pass = ck::profiler::run_instances(
        op_ptrs,
        [&](auto& op) {
            return op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), s_buf.GetDeviceBuffer(), c_buf.GetDeviceBuffer(), M, N, K, K, K,
                                          N, PassThrough{}, PassThrough{}, PassThrough{});
        },
        [&](auto&, auto*) { return 0; },
        [&]() {
            c_buf.FromDevice(c_dev.mData.data());
            if(do_log) LogRangeAsType<float>(std::cout << "c_device: ", c_dev.mData, ",") << std::endl;
            return ck::utils::check_err(c_dev, c_host, "Error: Incorrect results!", tol, tol);
        },
        do_verification, time_kernel, n_warmup, n_iter, 2.0 * M * N * K, 2.0 * M * K + 1.0 * K * N + 2.0 * N + 2.0 * M * N, 1.0);
    return pass ? 0 : 1;
}

int profile_gemm_w4a16(int argc, char* argv[])
{
    if(argc != 10 && argc != 12) return usage("gemm_w4a16", "arg9: zero points (0: constant 8; 1: uint8 [K/128, N])\n");
    constexpr ck::index_t G = 128;
    const bool do_verification = std::stoi(argv[2]); const int init_method = std::stoi(argv[3]);
    const bool do_log = std::stoi(argv[4]), time_kernel = std::stoi(argv[5]);
    const int M = std::stoi(argv[6]), N = std::stoi(argv[7]), K = std::stoi(argv[8]);
    const bool use_zp = std::stoi(argv[9]);
    const int n_warmup = argc == 12 ? std::stoi(argv[10]) : 1, n_iter = argc == 12 ? std::stoi(argv[11]) : 10;
    const std::size_t m = M, n = N, k = K, kg = (K + G - 1) / G;
    Tensor<BF16> a({m, k}), s({kg, n}), e_host({m, n}), e_dev(e_host.mDesc);
    Tensor<U8> q({k, n}), z({kg, n});
    const bool ints = init_method == 1;
    a.GenerateTensorValue([&](auto, auto) { return ck::type_convert<BF16>(ints ? float(std::rand() % 3 - 1) : 2.f * float(std::rand()) / float(RAND_MAX) - 1.f); });
    q.GenerateTensorValue([](auto, auto) { return U8(std::rand() & 15); });
    s.GenerateTensorValue([&](auto, auto) { return ck::type_convert<BF16>(ints ? 1.f : 0.005f + 0.015f * float(std::rand()) / float(RAND_MAX)); });
    z.GenerateTensorValue([](auto, auto) { return U8(std::rand() & 15); });
    DeviceMem a_buf(2 * a.GetElementSpaceSize()), q_buf(q.GetElementSpaceSize()), s_buf(2 * s.GetElementSpaceSize()), z_buf(z.GetElementSpaceSize());
    DeviceMem packed(ckb_prepack_b_int4_bytes(N, K)), e_buf(2 * e_dev.GetElementSpaceSize());
    a_buf.ToDevice(a.mData.data()); q_buf.ToDevice(q.mData.data()); s_buf.ToDevice(s.mData.data()); z_buf.ToDevice(z.mData.data());
    if(ckb_prepack_b_int4(q_buf.GetDeviceBuffer(), N, 1, N, K, packed.GetDeviceBuffer(), nullptr) != CKB_OK)
    {
        std::cerr << "prepack failed: " << ckb_last_error() << std::endl;
        return 1;
    }
    // both numerics contracts of the op (DequantizeGroup: ab_scale pattern, float reference operand; DequantizeGroupRound: one
    // rounding into A's type), each verified against its own host reference and timed
    auto run_contract = [&](auto bop, auto bref, const char* name) {
        using BOp = decltype(bop); using BRef = decltype(bref);
        std::cout << "B functor: " << name << std::endl;
        if(do_verification)
        {
            Tensor<BRef> b({k, n});
            for(int i = 0; i < K; ++i) for(int j = 0; j < N; ++j) BOp{}(b(i, j), int(q(i, j)) - int(use_zp ? z(i / G, j) : 8), s(i / G, j));
            Tensor<F32> c({m, n});
            using Ref = ck::tensor_operation::host::ReferenceGemm<BF16, BRef, F32, F32, PassThrough, PassThrough, PassThrough>;
            auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a, b, c, PassThrough{}, PassThrough{}, PassThrough{});
            inv.Run(arg);
            for(int i = 0; i < M; ++i) for(int j = 0; j < N; ++j) PassThrough{}(e_host(i, j), c(i, j));
        }
        using Op = ck::tensor_operation::device::DeviceGemm_dequantB_GroupScale_Sm100<Row, Row, BF16, BF16, G, PassThrough, BOp, PassThrough>;
        std::vector<std::unique_ptr<Op>> op_ptrs;
        op_ptrs.push_back(std::make_unique<Op>());
        const double bytes = 2.0 * M * K + 0.5 * N * K + 2.0 * kg * N + (use_zp ? 1.0 * kg * N : 0.0) + 2.0 * M * N;
        return ck::profiler::run_instances(
            op_ptrs,
            [&](auto& op) {
                return op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), packed.GetDeviceBuffer(), s_buf.GetDeviceBuffer(), use_zp ? z_buf.GetDeviceBuffer() : nullptr,
                                              e_buf.GetDeviceBuffer(), M, N, K, K, N, 8, PassThrough{}, BOp{}, PassThrough{});
            },
            [&](auto& op, auto* arg) {
                auto ws = std::make_shared<DeviceMem>(op.GetWorkSpaceSize(arg));
                op.SetWorkSpacePointer(arg, ws->GetDeviceBuffer());
                return ws;
            },
            [&]() {
                e_buf.FromDevice(e_dev.mData.data());
                if(do_log) LogRangeAsType<float>(std::cout << "e_device: ", e_dev.mData, ",") << std::endl;
                return ck::utils::check_err(e_dev, e_host, "Error: Incorrect results!", 1e-2, 1e-2);
            },
            do_verification, time_kernel, n_warmup, n_iter, 2.0 * M * N * K, bytes, 1.0);
    };
    const bool pass_f32 = run_contract(DequantizeGroup{}, F32{}, "DequantizeGroup (fp32 group scale)");
    const bool pass_rtn = run_contract(DequantizeGroupRound{}, BF16{}, "DequantizeGroupRound (one rounding into A's type)");
    const bool pass = pass_f32 && pass_rtn;
    return pass ? 0 : 1;
}

REGISTER_PROFILER_OPERATION("fpAintB_gemm", "GEMM fp16 A x int8 B, per-channel fp16 scale (example 64)", profile_fpAintB_gemm);
REGISTER_PROFILER_OPERATION("gemm_w4a16", "[EXT] GEMM bf16 A x packed int4 B, group-128 scales / zero points", profile_gemm_w4a16);
