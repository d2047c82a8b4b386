// Port of client_example/30_gemm_bf16Aint8B/gemm_bias_xdl_bf16_i8.cpp: instance-factory lookup of
// DeviceGemmMultipleABD<Row; {Row|Col,..}; Row; Row; BF16; {I8,BF16}; BF16; BF16; PassThrough, Multiply, Add>, run every
// instance, print the best.  The reference client does no verification; this port adds the host check of
// example/60_gemm_multi_ABD (b = Multiply(b0,b1) into bf16, ReferenceGemm, then Add) so it doubles as a test.
// usage: client_gemm_bias_bf16_i8 [M N K] [layout: 0 = mk_kn_mn, 1 = mk_nk_mn, 2 = km_kn_mn (column-major A)] [verify] [cde: add|add_gelu]
#include <cstring>
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/tensor_operation_instance/gpu/gemm_multi_abd.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/fill.hpp"
#include "ck/library/utility/host_tensor.hpp"

using BF16 = ck::bhalf_t; using I8 = int8_t; using F32 = float;
using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
using PassThrough = ck::tensor_operation::element_wise::PassThrough;
using Multiply = ck::tensor_operation::element_wise::Multiply;

template <typename BLayout, typename CDEOp, typename ALayout = Row>
int run(ck::index_t M, ck::index_t N, ck::index_t K, bool verify)
{
    using namespace ck::literals;
    constexpr bool b_row = std::is_same_v<BLayout, Row>;
    constexpr bool a_row = std::is_same_v<ALayout, Row>;
    const ck::index_t StrideA = a_row ? K : M, StrideB = b_row ? N : K, StrideD = 0, StrideE = N; // bias broadcast along M
    Tensor<BF16> a0_m_k(a_row ? HostTensorDescriptor({std::size_t(M), std::size_t(K)}, {std::size_t(StrideA), 1_uz})
                              : HostTensorDescriptor({std::size_t(M), std::size_t(K)}, {1_uz, std::size_t(StrideA)}));
    Tensor<I8> b0_k_n(b_row ? HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {std::size_t(StrideB), 1_uz})
                            : HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {1_uz, std::size_t(StrideB)}));
    Tensor<BF16> b1_k_n(HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {0_uz, 1_uz}));
    Tensor<BF16> d_m_n(HostTensorDescriptor({std::size_t(M), std::size_t(N)}, {std::size_t(StrideD), 1_uz}));
    Tensor<BF16> e_host(HostTensorDescriptor({std::size_t(M), std::size_t(N)}, {std::size_t(StrideE), 1_uz})), e_dev(e_host.mDesc);
    a0_m_k.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(float(std::rand() % 5 - 2)); });
    b0_k_n.GenerateTensorValue([](auto, auto) { return I8(std::rand() % 255 - 127); });
    b1_k_n.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(0.0078125f * float(1 + std::rand() % 4)); });
    d_m_n.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(float(std::rand() % 9 - 4)); });

    DeviceMem a_buf(2 * a0_m_k.GetElementSpaceSize()), b0_buf(b0_k_n.GetElementSpaceSize()), b1_buf(2 * b1_k_n.GetElementSpaceSize());
    DeviceMem d_buf(2 * d_m_n.GetElementSpaceSize()), e_buf(2 * e_dev.GetElementSpaceSize());
    a_buf.ToDevice(a0_m_k.mData.data()); b0_buf.ToDevice(b0_k_n.mData.data()); b1_buf.ToDevice(b1_k_n.mData.data());
    d_buf.ToDevice(d_m_n.mData.data());

    using DeviceOp = ck::tensor_operation::device::DeviceGemmMultipleABD<ck::Tuple<ALayout>, ck::Tuple<BLayout, BLayout>, ck::Tuple<Row>, Row,
                                                                         ck::Tuple<BF16>, ck::Tuple<I8, BF16>, ck::Tuple<BF16>, BF16,
                                                                         PassThrough, Multiply, CDEOp>;
    const auto op_ptrs = ck::tensor_operation::device::instance::DeviceOperationInstanceFactory<DeviceOp>::GetInstances();
    std::cout << "found " << op_ptrs.size() << " instances" << std::endl;
    std::string best_name; float best_tflops = 0, best_ms = 0, best_gbs = 0; int found = 0;
    for(auto& op_ptr : op_ptrs)
    {
        auto argument_ptr = op_ptr->MakeArgumentPointer({a_buf.GetDeviceBuffer()}, {b0_buf.GetDeviceBuffer(), b1_buf.GetDeviceBuffer()},
                                                        {d_buf.GetDeviceBuffer()}, e_buf.GetDeviceBuffer(), M, N, K, {StrideA},
                                                        {StrideB, 0}, {StrideD}, StrideE, PassThrough{}, Multiply{}, CDEOp{});
        auto invoker_ptr = op_ptr->MakeInvokerPointer();
        // row-major B and column-major A need workspace (K-major copy of the weight / row-major copy of the activations)
        DeviceMem ws(op_ptr->GetWorkSpaceSize(argument_ptr.get()));
        if(ws.GetDeviceBuffer()) op_ptr->SetWorkSpacePointer(argument_ptr.get(), ws.GetDeviceBuffer());
        if(!op_ptr->IsSupportedArgument(argument_ptr.get())) { std::cout << op_ptr->GetTypeString() << " does not support this problem" << std::endl; continue; }
        ++found;
        const float ave_time = invoker_ptr->Run(argument_ptr.get(), StreamConfig{nullptr, true, 0, 3, 10});
        const std::size_t flop = 2_uz * M * N * K, bytes = 2_uz * M * K + 1_uz * K * N + 2_uz * M * N;
        const float tflops = float(flop) / 1.E9 / ave_time, gbs = bytes / 1.E6 / ave_time;
        std::cout << "Perf: " << ave_time << " ms, " << tflops << " TFlops, " << gbs << " GB/s, " << op_ptr->GetTypeString() << std::endl;
        if(tflops > best_tflops) { best_tflops = tflops; best_ms = ave_time; best_gbs = gbs; best_name = op_ptr->GetTypeString(); }
    }
    if(!found) { std::cerr << "no instance supports this problem" << std::endl; return 2; }
    std::cout << "Best Perf: " << best_ms << " ms, " << best_tflops << " TFlops, " << best_gbs << " GB/s, " << best_name << std::endl;
    if(!verify) return 0;
    e_buf.FromDevice(e_dev.mData.data());
    // host check of example/60...:229-266 with CShuffleDataType = F32 (the library instances' choice)
    Tensor<BF16> b_k_n(HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {std::size_t(N), 1_uz}));
    for(int n = 0; n < N; ++n) for(int k = 0; k < K; ++k) Multiply{}(b_k_n(k, n), b0_k_n(k, n), b1_k_n(k, n));
    Tensor<F32> c_m_n({std::size_t(M), std::size_t(N)});
    using Ref = ck::tensor_operation::host::ReferenceGemm<BF16, BF16, F32, F32, PassThrough, PassThrough, PassThrough>;
    auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a0_m_k, b_k_n, c_m_n, PassThrough{}, PassThrough{}, PassThrough{});
    inv.Run(arg);
    for(int m = 0; m < M; ++m) for(int n = 0; n < N; ++n) CDEOp{}(e_host(m, n), c_m_n(m, n), d_m_n(m, n));
    const bool ok = ck::utils::check_err(e_dev, e_host, "Error: Incorrect results!", 1e-2, 1e-2);
    std::cout << (ok ? "verification PASSED" : "verification FAILED") << std::endl;
    return ok ? 0 : 1;
}

int main(int argc, char* argv[])
{
    ck::index_t M = 3840, N = 4096, K = 4096; int b_col = 0; bool verify = false; const char* cde = "add";
    if(argc >= 4) { M = std::atoi(argv[1]); N = std::atoi(argv[2]); K = std::atoi(argv[3]); }
    if(argc >= 5) b_col = std::atoi(argv[4]);
    if(argc >= 6) verify = std::atoi(argv[5]);
    if(argc >= 7) cde = argv[6];
    using Add = ck::tensor_operation::element_wise::Add; using AddFastGelu = ck::tensor_operation::element_wise::AddFastGelu;
    if(b_col == 2) return !std::strcmp(cde, "add_gelu") ? run<Row, AddFastGelu, Col>(M, N, K, verify) : run<Row, Add, Col>(M, N, K, verify);
    if(!std::strcmp(cde, "add_gelu")) return b_col ? run<Col, AddFastGelu>(M, N, K, verify) : run<Row, AddFastGelu>(M, N, K, verify);
    return b_col ? run<Col, Add>(M, N, K, verify) : run<Row, Add>(M, N, K, verify);
}
