// Drop-in check of the concrete operator classes: the instance declarations below carry the reference's OWN template argument
// lists, typed exactly as the reference examples spell them (every tile-shape / vector-width / pipeline argument included):
//   example/64_fpAintB_gemm/fp16int8_gemm_wmma.cpp:38-80                      DeviceFpAintBGemm_Wmma_CShuffle
//   example/14_gemm_quantization/gemm_xdl_quantization_int8.cpp:49-92        DeviceGemmMultipleD_Xdl_CShuffle
//   example/60_gemm_multi_ABD/gemm_multi_ABD_xdl_bias_fastgelu_bf16_i8.cpp:64-70   DeviceGemmMultipleABD_Xdl_CShuffle
//   example/65_gemm_multiply_multiply/gemm_multiply_multiply_xdl_fp8.cpp (argument order of DeviceGemmMultiD_Xdl_CShuffle_V3,
//     with the [EXT] int8 data types of cfg4)
// The file only has to COMPILE against include/ck and construct the operators (no GPU needed): it prints GetTypeString() of each.
// With a GPU (argv[1] == "run") it also pushes one small problem through each operator's type-erased interface.
#include <iostream>
#include <vector>

#include "ck/library/utility/device_memory.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_fpAintB_gemm_wmma.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_multiple_abd_xdl_cshuffle.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_multiple_d_xdl_cshuffle.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_multiple_d_xdl_cshuffle_v3.hpp"

template <ck::index_t... Is>
using S = ck::Sequence<Is...>;

using F16  = ck::half_t;
using BF16 = ck::bhalf_t;
using F32  = float;
using I8   = int8_t;
using I32  = int32_t;
using Row  = ck::tensor_layout::gemm::RowMajor;
using Col  = ck::tensor_layout::gemm::ColumnMajor;

using PassThrough = ck::tensor_operation::element_wise::PassThrough;

// ---------------------------------------------------------------- example 64
namespace ex64 {
using ADataType        = ck::half_t;
using BDataType        = uint8_t;
using ScaleDataType    = ck::half_t;
using AccDataType      = float;
using CShuffleDataType = float;
using CDataType        = ck::half_t;
using ALayout = Row;
using BLayout = Col;
using CLayout = Row;
using AElementOp = PassThrough;
using BElementOp = PassThrough;
using CElementOp = PassThrough;
static constexpr auto GemmDefault = ck::tensor_operation::device::GemmSpecialization::MNKPadding;
// clang-format off
using DeviceGemmInstance = ck::tensor_operation::device::DeviceFpAintBGemm_Wmma_CShuffle
         < ALayout,
           BLayout,
           CLayout,
           ADataType,
           BDataType,
           ScaleDataType,
           CDataType,
           AccDataType,
           CShuffleDataType,
           AElementOp,
           BElementOp,
           CElementOp,
           GemmDefault,
           1,           // Prefetch stage
           128,         // BlockSize
           64,          // MPerBlock
           128,         // NPerBlock
           64,          // KPerBlock
           8,           // K1
           16,          // MPerWmma
           16,          // NPerWmma
           2,           // M-Repeat // M-PerWmma / M-Repeat = M-Wave
           4,           // N-Repeat // N-PerWmma / N-Repeat = N-Wave
           S<4, 32, 1>,
           S<1, 0, 2>,
           S<1, 0, 2>,
           2,
           8,
           8,
           true,
           S<4, 32, 1>,
           S<1, 0, 2>,
           S<1, 0, 2>,
           2,
           8,
           8,
           true,
           1,           // C shuffle (M Repeat) Per store
           1,           // C shuffle (N Repeat) Per store
           S<1, 32, 1,  4>,
           8>;
// clang-format on
} // namespace ex64

// ---------------------------------------------------------------- example 14
namespace ex14 {
using ActivationOp = PassThrough;
using CDEElementOp = ck::tensor_operation::element_wise::Activation_Mul_Clamp<ActivationOp>;
using ADataType        = I8;
using BDataType        = I8;
using AccDataType      = I32;
using CShuffleDataType = I32;
using DsDataType       = ck::Tuple<>;
using EDataType        = I8;
using ALayout  = Row;
using BLayout  = Col;
using DsLayout = ck::Tuple<>;
using ELayout  = Row;
static constexpr auto GemmDefault = ck::tensor_operation::device::GemmSpecialization::Default;
// clang-format off
using DeviceGemmInstance = ck::tensor_operation::device::DeviceGemmMultipleD_Xdl_CShuffle<
     ALayout,
     BLayout,
     DsLayout,
     ELayout,
     ADataType,
     BDataType,
     AccDataType,
     CShuffleDataType,
     DsDataType,
     EDataType,
     PassThrough,                // AElementwiseOperation,
     PassThrough,                // BElementwiseOperation,
     CDEElementOp,               // CDEElementwiseOperation,
     GemmDefault,                // GemmSpecialization GemmSpec,
     1,                          // NumGemmKPrefetchStage,
     256,                        // BlockSize,
     256,                        // MPerBlock,
     128,                        // NPerBlock,
     64,                         // KPerBlock,
     16,                         // AK1,
     16,                         // BK1,
     32,                         // MPerXDL,
     32,                         // NPerXDL,
     4,                          // MXdlPerWave,
     2,                          // NXdlPerWave,
     S<4, 64, 1>,                // ABlockTransferThreadClusterLengths_AK0_M_AK1,
     S<1, 0, 2>,                 // ABlockTransferThreadClusterArrangeOrder,
     S<1, 0, 2>,                 // ABlockTransferSrcAccessOrder,
     2,                          // index_t ABlockTransferSrcVectorDim,
     16,                         // index_t ABlockTransferSrcScalarPerVector,
     16,                         // index_t ABlockTransferDstScalarPerVector_AK1,
     1,                          // bool ABlockLdsExtraM,
     S<4, 64, 1>,                // typename BBlockTransferThreadClusterLengths_BK0_N_BK1,
     S<1, 0, 2>,                 // typename BBlockTransferThreadClusterArrangeOrder,
     S<1, 0, 2>,                 // typename BBlockTransferSrcAccessOrder,
     2,                          // index_t BBlockTransferSrcVectorDim,
     8,                          // index_t BBlockTransferSrcScalarPerVector,
     8,                          // index_t BBlockTransferDstScalarPerVector_BK1,
     1,                          // bool BBlockLdsExtraN,
     1,                          // index_t CShuffleMXdlPerWavePerShuffle,
     1,                          // index_t CShuffleNXdlPerWavePerShuffle,
     S<1, 64, 1, 4>,             // typename CShuffleBlockTransferClusterLengths_MBlock_MPerBlock_NBlock_NPerBlock,
     16>;                        // index_t CShuffleBlockTransferScalarPerVector_NPerBlock>
// clang-format on
} // namespace ex14

// ---------------------------------------------------------------- example 60 (bias + FastGelu, bf16 A x int8 B0 x bf16 B1)
namespace ex60 {
using A0DataType       = BF16;
using AsDataType       = ck::Tuple<A0DataType>;
using B0DataType       = I8;
using B1DataType       = BF16;
using BsDataType       = ck::Tuple<B0DataType, B1DataType>;
using AccDataType      = F32;
using CShuffleDataType = BF16;
using D0DataType       = BF16;
using DsDataType       = ck::Tuple<D0DataType>;
using EDataType        = BF16;
using A0Layout = Row;
using AsLayout = ck::Tuple<A0Layout>;
using B0Layout = Row;
using B1Layout = B0Layout;
using BsLayout = ck::Tuple<B0Layout, B1Layout>;
using D0Layout = Row;
using DsLayout = ck::Tuple<D0Layout>;
using ELayout  = Row;
using Multiply    = ck::tensor_operation::element_wise::Multiply;
using AddFastGelu = ck::tensor_operation::element_wise::AddFastGelu;
using AElementOp   = PassThrough;
using BElementOp   = Multiply;
using CDEElementOp = AddFastGelu;
static constexpr auto GemmSpec = ck::tensor_operation::device::GemmSpecialization::MNKPadding;
// clang-format off
using DeviceOpInstance = ck::tensor_operation::device::DeviceGemmMultipleABD_Xdl_CShuffle
         < AsLayout, BsLayout, DsLayout, ELayout, AsDataType, BsDataType, AccDataType, CShuffleDataType, DsDataType, EDataType,  AElementOp,  BElementOp, CDEElementOp,       GemmSpec,        1,   256,   128,   128,    64,   8,   4,  32,   32,    2,    2,     S<8, 32, 1>,     S<1, 0, 2>,    S<1, 0, 2>,               2,              8,              8,          0,    S<16, 16, 1>,    S<0, 2, 1>,     S<0, 2, 1>,             1,               8,              4,          0,          1,           1,               S<1, 32, 1, 8>,               8,  ck::BlockGemmPipelineScheduler::Intrawave, ck::BlockGemmPipelineVersion::v4>;
// clang-format on
} // namespace ex60

// ---------------------------------------------------------------- example 65 argument order (V3 class), int8 [EXT]
namespace ex65 {
using A0DataType       = I8;
using B0DataType       = I8;
using AccDataType      = I32;
using CShuffleDataType = I32;
using D0DataType       = F32;
using D1DataType       = F32;
using DsDataType       = ck::Tuple<D0DataType, D1DataType>;
using EDataType        = BF16;
using A0Layout = Row;
using B0Layout = Col;
using D0Layout = Row;
using D1Layout = Col;
using DsLayout = ck::Tuple<D0Layout, D1Layout>;
using ELayout  = Row;
using MultiplyMultiply = ck::tensor_operation::element_wise::MultiplyMultiply;
using AElementOp   = PassThrough;
using BElementOp   = PassThrough;
using CDEElementOp = MultiplyMultiply;
static constexpr auto GemmSpec = ck::tensor_operation::device::GemmSpecialization::Default;
// clang-format off
using DeviceOpInstance = ck::tensor_operation::device::DeviceGemmMultiD_Xdl_CShuffle_V3
    < Row, Col, DsLayout, ELayout, A0DataType, B0DataType, DsDataType, EDataType, AccDataType, CShuffleDataType,
      AElementOp,  BElementOp, CDEElementOp, GemmSpec, 256,
      256, 256, 64,
      16, 16,
      32, 32,
      4, 4,
      S<4, 64, 1>, S<1, 0, 2>, S<1, 0, 2>, 2, 16, 16, 0,
      S<4, 64, 1>, S<1, 0, 2>, S<1, 0, 2>, 2, 16, 16, 0,
      1, 1, S<1, 32, 1, 8>, S<8, 8, 1>,
      ck::BlockGemmPipelineScheduler::Intrawave, ck::BlockGemmPipelineVersion::v3, A0DataType>;
// clang-format on
} // namespace ex65

template <typename Op, typename Iface>
void show()
{
    static_assert(std::is_base_of_v<Iface, Op>, "the concrete operator must implement the reference's device interface");
    Op op;
    std::cout << op.GetTypeString() << std::endl;
}

int main(int argc, char* argv[])
{
    namespace dev = ck::tensor_operation::device;
    show<ex64::DeviceGemmInstance, dev::DeviceGemm_dequantB<Row, Col, Row, F16, uint8_t, F16, PassThrough, PassThrough, PassThrough>>();
    show<ex14::DeviceGemmInstance, dev::DeviceGemmMultipleD<Row, Col, ck::Tuple<>, Row, I8, I8, ck::Tuple<>, I8, PassThrough, PassThrough, ex14::CDEElementOp>>();
    show<ex60::DeviceOpInstance, dev::DeviceGemmMultipleABD<ex60::AsLayout, ex60::BsLayout, ex60::DsLayout, Row, ex60::AsDataType, ex60::BsDataType,
                                                           ex60::DsDataType, BF16, PassThrough, ex60::Multiply, ex60::AddFastGelu>>();
    show<ex65::DeviceOpInstance, dev::DeviceGemmMultipleD<Row, Col, ex65::DsLayout, Row, I8, I8, ex65::DsDataType, BF16, PassThrough, PassThrough,
                                                          ex65::MultiplyMultiply>>();
    if(argc > 1 && std::string(argv[1]) == "run")
    {
        // one small int8 problem through the example-14 instance (type-erased interface), result not checked here (the parity
        // tests do that): this exercises MakeArgumentPointer / IsSupportedArgument / Run / GetTypeString with a real plan
        const ck::index_t M = 256, N = 256, K = 256;
        std::vector<int8_t> a(M * K, 1), b(N * K, 2), e(M * N, 0);
        DeviceMem a_buf(a.size()), b_buf(b.size()), e_buf(e.size());
        a_buf.ToDevice(a.data());
        b_buf.ToDevice(b.data());
        ex14::DeviceGemmInstance op;
        auto arg = op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), {}, e_buf.GetDeviceBuffer(), M, N, K, K, K, {}, N,
                                          PassThrough{}, PassThrough{}, ex14::CDEElementOp{0.01f, ex14::ActivationOp{}});
        if(!op.IsSupportedArgument(arg.get())) { std::cerr << "unsupported" << std::endl; return 1; }
        op.MakeInvokerPointer()->Run(arg.get(), StreamConfig{nullptr, false});
        e_buf.FromDevice(e.data());
        std::cout << op.GetTypeString() << "  e[0] = " << int(e[0]) << std::endl;   // 256 * 1 * 2 * 0.01 = 5.12 -> 5
        if(e[0] != 5) return 1;
    }
    return 0;
}
