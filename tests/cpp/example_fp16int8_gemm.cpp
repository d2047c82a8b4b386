// Port of example/64_fpAintB_gemm (fp16int8_gemm_wmma.cpp + run_gemm_example.inc + common.hpp) onto the B200 build:
// fp16 A (Row) x int8 B (Col, +128-biased on the device) x fp16 per-channel scale -> fp16 C, verified against
// ReferencefpAintBGemm with ck::utils::check_err (fp16 default 1e-3 for init_method 1; 1e-2 for real-valued init).
// usage: example_fp16int8_gemm [verify init time] [M N K [StrideA StrideB StrideC]]
#include <cstdlib>
#include <iostream>

#include "ck/library/reference_tensor_operation/cpu/reference_fpAintB_gemm.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/fill.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_fpAintB_gemm_wmma.hpp"

using ADataType = ck::half_t; using QuantDataType = int8_t; using BDataType = uint8_t; using ScaleDataType = ck::half_t;
using AccDataType = float; using CShuffleDataType = float; using CDataType = ck::half_t;
using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
using PassThrough = ck::tensor_operation::element_wise::PassThrough;
using DeviceGemmInstance = ck::tensor_operation::device::DeviceFpAintBGemm_Wmma_CShuffle<
    Row, Col, Row, ADataType, BDataType, ScaleDataType, CDataType, AccDataType, CShuffleDataType, PassThrough, PassThrough, PassThrough>;
using ReferenceGemmInstance = ck::tensor_operation::host::ReferencefpAintBGemm<ADataType, QuantDataType, ScaleDataType, CDataType,
                                                                              AccDataType, PassThrough, PassThrough, PassThrough>;

int main(int argc, char* argv[])
{
    bool do_verification = true; int init_method = 1; bool time_kernel = false;
    ck::index_t M = 3840, N = 4096, K = 4096, StrideA = 4096, StrideB = 4096, StrideC = 4096;
    if(argc >= 4) { do_verification = std::atoi(argv[1]); init_method = std::atoi(argv[2]); time_kernel = std::atoi(argv[3]); }
    if(argc >= 7) { M = std::atoi(argv[4]); N = std::atoi(argv[5]); K = std::atoi(argv[6]); StrideA = K; StrideB = K; StrideC = N; }
    if(argc >= 10) { StrideA = std::atoi(argv[7]); StrideB = std::atoi(argv[8]); StrideC = std::atoi(argv[9]); }

    using namespace ck::literals;
    Tensor<ADataType> a_m_k(HostTensorDescriptor({std::size_t(M), std::size_t(K)}, {std::size_t(StrideA), 1_uz}));
    Tensor<QuantDataType> quant_b_k_n(HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {1_uz, std::size_t(StrideB)}));
    Tensor<ScaleDataType> scale_k_n(HostTensorDescriptor({std::size_t(K), std::size_t(N)}, {0_uz, 1_uz})); // run_gemm_example.inc:31
    switch(init_method)
    {
    case 0: break;
    case 1:
        ck::utils::FillUniformDistributionIntegerValue<ADataType>{-1.f, 1.f}(a_m_k);
        ck::utils::FillUniformDistributionIntegerValue<QuantDataType>{-1.f, 1.f}(quant_b_k_n);
        ck::utils::FillUniformDistributionIntegerValue<ScaleDataType>{-1.f, 1.f}(scale_k_n);
        break;
    default: // real-valued data: weights over the whole int8 range (the reference's -1..1 fill truncates to ~0)
        ck::utils::FillUniformDistribution<ADataType>{-1.f, 1.f}(a_m_k);
        ck::utils::FillUniformDistribution<QuantDataType>{-127.f, 127.f}(quant_b_k_n);
        ck::utils::FillUniformDistribution<ScaleDataType>{0.001f, 0.02f}(scale_k_n);
    }
    // UnsignedWeightPreprocessor (common.hpp:56-83): u8 = int8 + 128
    Tensor<BDataType> b_k_n(quant_b_k_n.mDesc);
    for(std::size_t i = 0; i < b_k_n.mData.size(); ++i) b_k_n.mData[i] = BDataType(int(quant_b_k_n.mData[i]) + 128);
    Tensor<CDataType> c_host(HostTensorDescriptor({std::size_t(M), std::size_t(N)}, {std::size_t(StrideC), 1_uz})), c_dev(c_host.mDesc);
    std::cout << "a_m_k: " << a_m_k.mDesc << "\nb_k_n: " << b_k_n.mDesc << "\nscale_k_n: " << scale_k_n.mDesc << "\nc_m_n: " << c_host.mDesc << std::endl;

    DeviceMem a_buf(sizeof(ADataType) * a_m_k.GetElementSpaceSize()), b_buf(sizeof(BDataType) * b_k_n.GetElementSpaceSize());
    DeviceMem s_buf(sizeof(ScaleDataType) * scale_k_n.GetElementSpaceSize()), c_buf(sizeof(CDataType) * c_dev.GetElementSpaceSize());
    a_buf.ToDevice(a_m_k.mData.data()); b_buf.ToDevice(b_k_n.mData.data()); s_buf.ToDevice(scale_k_n.mData.data());

    auto gemm = DeviceGemmInstance{}; auto invoker = gemm.MakeInvoker();
    auto argument = gemm.MakeArgument(static_cast<ADataType*>(a_buf.GetDeviceBuffer()), static_cast<BDataType*>(b_buf.GetDeviceBuffer()),
                                      static_cast<ScaleDataType*>(s_buf.GetDeviceBuffer()), static_cast<CDataType*>(c_buf.GetDeviceBuffer()),
                                      M, N, K, StrideA, StrideB, StrideC, PassThrough{}, PassThrough{}, PassThrough{});
    if(!gemm.IsSupportedArgument(argument))
    {
        std::cerr << gemm.GetTypeString() << " does not support this problem: " << argument.why << std::endl;
        return 2;
    }
    float ave_time = invoker.Run(argument, StreamConfig{nullptr, time_kernel});
    std::size_t flop = 2_uz * M * N * K, num_btype = sizeof(ADataType) * M * K + sizeof(BDataType) * K * N + sizeof(CDataType) * M * N;
    if(time_kernel)
        std::cout << "Perf: " << ave_time << " ms, " << float(flop) / 1.E9 / ave_time << " TFlops, " << num_btype / 1.E6 / ave_time
                  << " GB/s, " << argument.plan.kernel_name << std::endl;
    c_buf.FromDevice(c_dev.mData.data());
    if(do_verification)
    {
        auto ref = ReferenceGemmInstance{}; auto ref_inv = ref.MakeInvoker();
        auto ref_arg = ref.MakeArgument(a_m_k, quant_b_k_n, scale_k_n, c_host, PassThrough{}, PassThrough{}, PassThrough{});
        ref_inv.Run(ref_arg);
        const double tol = init_method == 1 ? 1e-3 : 1e-2;
        const bool ok = ck::utils::check_err(c_dev, c_host, "Error: Incorrect results!", tol, tol);
        std::cout << (ok ? "verification PASSED" : "verification FAILED") << std::endl;
        return ok ? 0 : 1;
    }
    return 0;
}
