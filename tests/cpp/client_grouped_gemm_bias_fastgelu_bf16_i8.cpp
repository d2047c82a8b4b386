// Port of client_example/31_grouped_gemm_bf16Aint8B/grouped_gemm_bias_fastgelu_xdl_bf16_i8.cpp: factory lookup of
// DeviceGroupedGemmMultiABDFixedNK<Row; {B,B}; Row; Row; BF16; {I8,BF16}; BF16; BF16; PassThrough, Multiply, AddFastGelu>,
// 16 groups with M = 32 + rand()%32, N = 1024, K = 512 (:267-286), the per-group arguments uploaded as a device array
// (SetDeviceKernelArgs, :214-223).  The reference client does not verify; this port adds the host check of
// example/59_grouped_gemm_multi_ABD (Multiply into bf16, ReferenceGemm, AddFastGelu) so it doubles as a test.
// usage: client_grouped_gemm_bias_fastgelu_bf16_i8 [groups N K] [b_layout 0 = Row (mk_kn_mn) | 1 = Col (mk_nk_mn)] [verify]
#include <cstring>
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/tensor_operation_instance/gpu/grouped_gemm_multi_abd_fixed_nk.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/host_tensor.hpp"

using BF16 = ck::bhalf_t; using I8 = int8_t; using F32 = float;
using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
using PassThrough = ck::tensor_operation::element_wise::PassThrough;
using Multiply = ck::tensor_operation::element_wise::Multiply;
using AddFastGelu = ck::tensor_operation::element_wise::AddFastGelu;

template <typename BLayout, typename ALayout = Row>
int run(int group_count, ck::index_t N, ck::index_t K, bool verify, ck::index_t a_pitch = 64)
{
    using namespace ck::literals;
    constexpr bool b_row = std::is_same_v<BLayout, Row>;
    constexpr ck::index_t NumATensor = 1, NumBTensor = 2, NumDTensor = 1;
    std::vector<ck::index_t> Ms;
    ck::index_t sum_of_m = 0;
    for(int i = 0; i < group_count; ++i) { Ms.push_back(32 + std::rand() % 32); sum_of_m += Ms.back(); }

    std::vector<Tensor<BF16>> a0, b1, d0, e_host, e_dev;
    std::vector<Tensor<I8>> b0;
    std::vector<std::unique_ptr<DeviceMem>> a0_buf, b0_buf, b1_buf, d0_buf, e_buf;
    std::vector<ck::tensor_operation::device::GemmMultiABDDesc> gemm_descs;
    std::vector<ck::tensor_operation::device::GroupedGemmMultiABDKernelArgument<NumATensor, NumBTensor, NumDTensor>> kargs;
    constexpr bool a_row = std::is_same_v<ALayout, Row>;
    // km_kn_mn: A[m,k] = a[k * StrideA + m]; one pitch >= max M for every group (64: 16-byte vector loads; 67: the element path)
    const ck::index_t StrideA = a_row ? K : a_pitch, StrideB = b_row ? N : K, StrideE = N;
    for(int i = 0; i < group_count; ++i)
    {
        const std::size_t M = Ms[i];
        a0.emplace_back(a_row ? HostTensorDescriptor({M, std::size_t(K)}, {std::size_t(StrideA), 1_uz})
                              : HostTensorDescriptor({M, std::size_t(K)}, {1_uz, std::size_t(StrideA)}));
        b0.emplace_back(b_row ? HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {std::size_t(StrideB), 1_uz})
                              : HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {1_uz, std::size_t(StrideB)}));
        b1.emplace_back(HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {0_uz, 1_uz}));
        d0.emplace_back(HostTensorDescriptor({M, std::size_t(N)}, {0_uz, 1_uz}));
        e_host.emplace_back(HostTensorDescriptor({M, std::size_t(N)}, {std::size_t(StrideE), 1_uz}));
        e_dev.emplace_back(e_host.back().mDesc);
        a0[i].GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(float(std::rand() % 5 - 2)); });
        b0[i].GenerateTensorValue([](auto, auto) { return I8(std::rand() % 255 - 127); });
        b1[i].GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(0.0078125f * float(1 + std::rand() % 4)); });
        d0[i].GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(float(std::rand() % 9 - 4)); });
        a0_buf.emplace_back(std::make_unique<DeviceMem>(2 * a0[i].GetElementSpaceSize()));
        b0_buf.emplace_back(std::make_unique<DeviceMem>(b0[i].GetElementSpaceSize()));
        b1_buf.emplace_back(std::make_unique<DeviceMem>(2 * b1[i].GetElementSpaceSize()));
        d0_buf.emplace_back(std::make_unique<DeviceMem>(2 * d0[i].GetElementSpaceSize()));
        e_buf.emplace_back(std::make_unique<DeviceMem>(2 * e_dev[i].GetElementSpaceSize()));
        a0_buf[i]->ToDevice(a0[i].mData.data()); b0_buf[i]->ToDevice(b0[i].mData.data());
        b1_buf[i]->ToDevice(b1[i].mData.data()); d0_buf[i]->ToDevice(d0[i].mData.data());
        gemm_descs.push_back({sum_of_m, N, K, {1}, {1, 1}, {0}, 1}); // as the reference client passes them (:150-151)
        gemm_descs.back().stride_As_ = {StrideA}; gemm_descs.back().stride_Bs_ = {StrideB, 0};
        gemm_descs.back().stride_Ds_ = {0}; gemm_descs.back().stride_C_ = StrideE;
        kargs.push_back({std::array<const void*, 1>{a0_buf[i]->GetDeviceBuffer()},
                         std::array<const void*, 2>{b0_buf[i]->GetDeviceBuffer(), b1_buf[i]->GetDeviceBuffer()},
                         std::array<const void*, 1>{d0_buf[i]->GetDeviceBuffer()}, e_buf[i]->GetDeviceBuffer(), Ms[i], N, K,
                         std::array<ck::index_t, 1>{StrideA}, std::array<ck::index_t, 2>{StrideB, 0}, std::array<ck::index_t, 1>{0}, StrideE});
    }

    using DeviceOp = ck::tensor_operation::device::DeviceGroupedGemmMultiABDFixedNK<ck::Tuple<ALayout>, ck::Tuple<BLayout, BLayout>, ck::Tuple<Row>,
                                                                                    Row, ck::Tuple<BF16>, ck::Tuple<I8, BF16>, ck::Tuple<BF16>,
                                                                                    BF16, PassThrough, Multiply, AddFastGelu>;
    const auto op_ptrs = ck::tensor_operation::device::instance::DeviceOperationInstanceFactory<DeviceOp>::GetInstances();
    std::cout << "found " << op_ptrs.size() << " instances" << std::endl;
    std::string best_name; float best_tflops = 0, best_ms = 0, best_gbs = 0; int found = 0;
    for(auto& op_ptr : op_ptrs)
    {
        std::vector<std::array<const void*, NumATensor>> p_As; std::vector<std::array<const void*, NumBTensor>> p_Bs;
        std::vector<std::array<const void*, NumDTensor>> p_Ds; std::vector<void*> p_Cs;
        auto argument_ptr = op_ptr->MakeArgumentPointer(p_As, p_Bs, p_Ds, p_Cs, gemm_descs);
        auto invoker_ptr  = op_ptr->MakeInvokerPointer();
        if(!op_ptr->IsSupportedArgument(argument_ptr.get())) { std::cout << op_ptr->GetTypeString() << " does not support this problem" << std::endl; continue; }
        ++found;
        // Run before SetDeviceKernelArgs must throw, as in the reference (:597-600)
        bool threw = false;
        try { invoker_ptr->Run(argument_ptr.get(), StreamConfig{nullptr, false}); } catch(const std::runtime_error&) { threw = true; }
        if(!threw) { std::cerr << "Run without device kernel args did not throw" << std::endl; return 1; }
        DeviceMem kargs_dev(op_ptr->GetDeviceKernelArgSize(argument_ptr.get()));
        if(kargs_dev.GetBufferSize() != kargs.size() * sizeof(kargs[0])) { std::cerr << "GetDeviceKernelArgSize mismatch" << std::endl; return 1; }
        kargs_dev.ToDevice(kargs.data());
        op_ptr->SetDeviceKernelArgs(argument_ptr.get(), kargs_dev.GetDeviceBuffer());
        op_ptr->SetElementwiseOps(argument_ptr.get(), PassThrough{}, Multiply{}, AddFastGelu{});
        op_ptr->SetKBatch(argument_ptr.get(), 1);
        const float ave_time = invoker_ptr->Run(argument_ptr.get(), StreamConfig{nullptr, true, 0, 3, 20});
        const std::size_t flop = 2_uz * sum_of_m * N * K;
        const std::size_t bytes = 2_uz * sum_of_m * K + 1_uz * K * N * group_count + 2_uz * sum_of_m * N;
        const float tflops = float(flop) / 1.E9 / ave_time, gbs = bytes / 1.E6 / ave_time;
        std::cout << "Perf: " << ave_time << " ms, " << tflops << " TFlops, " << gbs << " GB/s, " << op_ptr->GetTypeString() << std::endl;
        if(tflops > best_tflops) { best_tflops = tflops; best_ms = ave_time; best_gbs = gbs; best_name = op_ptr->GetTypeString(); }
    }
    if(!found) { std::cerr << "no instance supports this problem" << std::endl; return 2; }
    std::cout << "Best Perf: " << best_ms << " ms, " << best_tflops << " TFlops, " << best_gbs << " GB/s, " << best_name << std::endl;
    if(!verify) return 0;
    bool ok = true;
    for(int i = 0; i < group_count; ++i)
    {
        e_buf[i]->FromDevice(e_dev[i].mData.data());
        Tensor<BF16> b_k_n(HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {std::size_t(N), 1_uz}));
        for(int n = 0; n < N; ++n) for(int k = 0; k < K; ++k) Multiply{}(b_k_n(k, n), b0[i](k, n), b1[i](k, n));
        Tensor<F32> c_m_n({std::size_t(Ms[i]), std::size_t(N)});
        using Ref = ck::tensor_operation::host::ReferenceGemm<BF16, BF16, F32, F32, PassThrough, PassThrough, PassThrough>;
        auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a0[i], b_k_n, c_m_n, PassThrough{}, PassThrough{}, PassThrough{});
        inv.Run(arg);
        for(int m = 0; m < Ms[i]; ++m) for(int n = 0; n < N; ++n) AddFastGelu{}(e_host[i](m, n), c_m_n(m, n), d0[i](m, n));
        ok = ck::utils::check_err(e_dev[i], e_host[i], "Error: Incorrect results!", 1e-2, 1e-2) && ok;
    }
    std::cout << (ok ? "verification PASSED" : "verification FAILED") << std::endl;
    return ok ? 0 : 1;
}

int main(int argc, char* argv[])
{
    int groups = 16; ck::index_t N = 1024, K = 512; int b_col = 0; bool verify = true;
    if(argc >= 4) { groups = std::atoi(argv[1]); N = std::atoi(argv[2]); K = std::atoi(argv[3]); }
    if(argc >= 5) b_col = std::atoi(argv[4]);
    if(argc >= 6) verify = std::atoi(argv[5]);
    if(argc >= 7 && std::atoi(argv[6]) > 0) return run<Row, Col>(groups, N, K, verify, std::atoi(argv[6]));   // km_kn_mn with this pitch
    return b_col ? run<Col>(groups, N, K, verify) : run<Row>(groups, N, K, verify);
}
