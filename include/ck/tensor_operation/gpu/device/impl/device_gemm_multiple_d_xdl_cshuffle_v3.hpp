// SPDX-License-Identifier: MIT
// The class name and template parameter list mirror Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence)
// so that client code compiles unchanged; the implementation is the B200 one (device_gemm_sm100.hpp).
// DeviceGemmMultiD_Xdl_CShuffle_V3 with the reference's full template parameter list
// (include/ck/tensor_operation/gpu/device/impl/device_gemm_multiple_d_xdl_cshuffle_v3.hpp:23-70).  NOTE the V3 order:
// ..., ADataType, BDataType, DsDataType, CDataType, GemmAccDataType, CShuffleDataType, ... (Ds / C before the accumulator
// types), unlike DeviceGemmMultipleD_Xdl_CShuffle.  A real class, not an alias: it re-orders into the implementation.
#pragma once
#include "ck/tensor_operation/gpu/device/impl/device_gemm_sm100.hpp"
namespace ck {
namespace tensor_operation {
namespace device {
template <typename ALayout, typename BLayout, typename DsLayout, typename CLayout, typename ADataType, typename BDataType,
          typename DsDataType, typename CDataType, typename GemmAccDataType, typename CShuffleDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation,
          GemmSpecialization GemmSpec = GemmSpecialization::Default,
          index_t BlockSize = 0, index_t MPerBlock = 0, index_t NPerBlock = 0, index_t KPerBlock = 0, index_t AK1 = 0,
          index_t BK1 = 0, index_t MPerXDL = 0, index_t NPerXDL = 0, index_t MXdlPerWave = 0, index_t NXdlPerWave = 0,
          typename ABlockTransferThreadClusterLengths_AK0_M_AK1 = void, typename ABlockTransferThreadClusterArrangeOrder = void,
          typename ABlockTransferSrcAccessOrder = void, index_t ABlockTransferSrcVectorDim = 0,
          index_t ABlockTransferSrcScalarPerVector = 0, index_t ABlockTransferDstScalarPerVector_AK1 = 0,
          bool ABlockLdsExtraM = false,
          typename BBlockTransferThreadClusterLengths_BK0_N_BK1 = void, typename BBlockTransferThreadClusterArrangeOrder = void,
          typename BBlockTransferSrcAccessOrder = void, index_t BBlockTransferSrcVectorDim = 0,
          index_t BBlockTransferSrcScalarPerVector = 0, index_t BBlockTransferDstScalarPerVector_BK1 = 0,
          bool BBlockLdsExtraN = false, index_t CShuffleMXdlPerWavePerShuffle = 0, index_t CShuffleNXdlPerWavePerShuffle = 0,
          typename CShuffleBlockTransferClusterLengths_MBlock_MPerBlock_NBlock_NPerBlock = void,
          typename CDEShuffleBlockTransferScalarPerVectors = void,
          BlockGemmPipelineScheduler BlkGemmPipeSched = BlockGemmPipelineScheduler::Intrawave,
          BlockGemmPipelineVersion BlkGemmPipelineVer = BlockGemmPipelineVersion::v1,
          typename ComputeTypeA = CDataType, typename ComputeTypeB = ComputeTypeA, typename LDSTypeA = ComputeTypeA,
          typename LDSTypeB = ComputeTypeB>
struct DeviceGemmMultiD_Xdl_CShuffle_V3
    : public DeviceGemmMultipleD_Sm100<ALayout, BLayout, DsLayout, CLayout, ADataType, BDataType, GemmAccDataType, CShuffleDataType,
                                       DsDataType, CDataType, AElementwiseOperation, BElementwiseOperation, CElementwiseOperation,
                                       GemmSpec>
{
    const char* ClassName() const override { return "DeviceGemmMultiD_Xdl_CShuffle_V3"; }
};
} // namespace device
} // namespace tensor_operation
} // namespace ck
