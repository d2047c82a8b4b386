// [EXT] W4A16 group-128 GEMM through the CK-style API (no reference counterpart: the reference's "int4" examples keep one
// int4 per byte and have no dequant path).  bf16 A x packed int4 B with bf16 scales [K/128, N] and optional
// zero points; weights are pre-packed once with ckb_prepack_b_int4.  Two numerics contracts, selected by the B functor and
// honoured at every M: DequantizeGroup (ab_scale pattern: host check = float (q - z) * s into ReferenceGemm, the composition of
// gemm_multiply_multiply_xdl_fp8_ab_scale.cpp:258-313) and DequantizeGroupRound (host check = B pre-dequantised into A's
// type with ONE rounding, then ReferenceGemm).
// usage: example_w4a16_group_gemm [M N K] [use_zero_points] [time] [round: 0 = DequantizeGroup, 1 = DequantizeGroupRound]
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_sm100.hpp"

using BF16 = ck::bhalf_t; using F32 = float;
using Row = ck::tensor_layout::gemm::RowMajor;
using namespace ck::tensor_operation::element_wise;
using namespace ck::literals;
constexpr ck::index_t G = 128;

template <typename BOp, typename BRef>
int run(ck::index_t M, ck::index_t N, ck::index_t K, bool use_zp, bool time_kernel)
{
    using DeviceOp = ck::tensor_operation::device::DeviceGemm_dequantB_GroupScale_Sm100<Row, Row, BF16, BF16, G, PassThrough, BOp, PassThrough>;
    const ck::index_t KG = K / G;
    Tensor<BF16> a({std::size_t(M), std::size_t(K)}), s({std::size_t(KG), std::size_t(N)}), e_host({std::size_t(M), std::size_t(N)}), e_dev(e_host.mDesc);
    Tensor<uint8_t> q({std::size_t(K), std::size_t(N)}), z({std::size_t(KG), std::size_t(N)});
    a.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(2.f * float(std::rand()) / float(RAND_MAX) - 1.f); });
    q.GenerateTensorValue([](auto, auto) { return uint8_t(std::rand() & 15); });
    s.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(0.005f + 0.015f * float(std::rand()) / float(RAND_MAX)); });
    z.GenerateTensorValue([](auto, auto) { return uint8_t(std::rand() & 15); });

    DeviceMem a_buf(2 * a.GetElementSpaceSize()), q_buf(q.GetElementSpaceSize()), s_buf(2 * s.GetElementSpaceSize()), z_buf(z.GetElementSpaceSize());
    DeviceMem packed(ckb_prepack_b_int4_bytes(N, K)), e_buf(2 * e_dev.GetElementSpaceSize());
    a_buf.ToDevice(a.mData.data()); q_buf.ToDevice(q.mData.data()); s_buf.ToDevice(s.mData.data()); z_buf.ToDevice(z.mData.data());
    if(ckb_prepack_b_int4(q_buf.GetDeviceBuffer(), N, 1, N, K, packed.GetDeviceBuffer(), nullptr) != CKB_OK)
    {
        std::cerr << "prepack failed: " << ckb_last_error() << std::endl;
        return 2;
    }
    auto op = DeviceOp{}; auto invoker = op.MakeInvoker();
    auto arg = op.MakeArgument(a_buf.GetDeviceBuffer(), packed.GetDeviceBuffer(), s_buf.GetDeviceBuffer(), use_zp ? z_buf.GetDeviceBuffer() : nullptr,
                               e_buf.GetDeviceBuffer(), M, N, K, K, N, 8, PassThrough{}, BOp{}, PassThrough{});
    if(!op.IsSupportedArgument(arg)) { std::cerr << "unsupported: " << arg.why << std::endl; return 2; }
    DeviceMem ws(op.GetWorkSpaceSize(&arg));
    op.SetWorkSpacePointer(&arg, ws.GetDeviceBuffer());
    const float ms = invoker.Run(arg, StreamConfig{nullptr, time_kernel, 0, 5, 50});
    const double bytes = 2.0 * M * K + 0.5 * N * K + 2.0 * KG * N + (use_zp ? 1.0 * KG * N : 0.0) + 2.0 * M * N;
    if(time_kernel)
        std::cout << "Perf: " << ms << " ms, " << 2.0 * M * N * K / 1.E9 / ms << " TFlops, " << bytes / 1.E6 / ms << " GB/s, " << arg.plan.kernel_name << std::endl;
    e_buf.FromDevice(e_dev.mData.data());
    Tensor<BRef> b_k_n({std::size_t(K), std::size_t(N)});
    for(int k = 0; k < K; ++k)
        for(int n = 0; n < N; ++n) BOp{}(b_k_n(k, n), int(q(k, n)) - int(use_zp ? z(k / G, n) : 8), s(k / G, n));
    Tensor<F32> c({std::size_t(M), std::size_t(N)});
    using Ref = ck::tensor_operation::host::ReferenceGemm<BF16, BRef, F32, F32, PassThrough, PassThrough, PassThrough>;
    auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto rarg = ref.MakeArgument(a, b_k_n, c, PassThrough{}, PassThrough{}, PassThrough{});
    inv.Run(rarg);
    for(int m = 0; m < M; ++m) for(int n = 0; n < N; ++n) PassThrough{}(e_host(m, n), c(m, n));
    const bool ok = ck::utils::check_err(e_dev, e_host, "Error: Incorrect results!", 1e-2, 1e-2);
    std::cout << (ok ? "verification PASSED" : "verification FAILED") << std::endl;
    return ok ? 0 : 1;
}

int main(int argc, char* argv[])
{
    ck::index_t M = 16, N = 4096, K = 4096; bool use_zp = false, time_kernel = true, round = false;
    if(argc >= 4) { M = std::atoi(argv[1]); N = std::atoi(argv[2]); K = std::atoi(argv[3]); }
    if(argc >= 5) use_zp = std::atoi(argv[4]);
    if(argc >= 6) time_kernel = std::atoi(argv[5]);
    if(argc >= 7) round = std::atoi(argv[6]);
    return round ? run<DequantizeGroupRound, BF16>(M, N, K, use_zp, time_kernel) : run<DequantizeGroup, F32>(M, N, K, use_zp, time_kernel);
}
