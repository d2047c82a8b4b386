// SPDX-License-Identifier: MIT
// The class name and template parameter list mirror Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence)
// so that client code compiles unchanged; the implementation is the B200 one (device_gemm_sm100.hpp).
// DeviceGemmMultipleD_Xdl_CShuffle with the reference's full template parameter list
// (include/ck/tensor_operation/gpu/device/impl/device_gemm_multiple_d_xdl_cshuffle.hpp:103-148); the XDL tiling parameters
// are accepted and ignored (defaulted).
#pragma once
#include "ck/tensor_operation/gpu/device/impl/device_gemm_sm100.hpp"
namespace ck {
namespace tensor_operation {
namespace device {
template <typename ALayout, typename BLayout, typename DsLayout, typename ELayout, typename ADataType, typename BDataType,
          typename AccDataType, typename CShuffleDataType, typename DsDataType, typename EDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CDEElementwiseOperation,
          GemmSpecialization GemmSpec = GemmSpecialization::Default,
          index_t NumGemmKPrefetchStage = 1, index_t BlockSize = 0, index_t MPerBlock = 0, index_t NPerBlock = 0,
          index_t KPerBlock = 0, index_t AK1 = 0, index_t BK1 = 0, index_t MPerXDL = 0, index_t NPerXDL = 0,
          index_t MXdlPerWave = 0, index_t NXdlPerWave = 0,
          typename ABlockTransferThreadClusterLengths_AK0_M_AK1 = void, typename ABlockTransferThreadClusterArrangeOrder = void,
          typename ABlockTransferSrcAccessOrder = void, index_t ABlockTransferSrcVectorDim = 0,
          index_t ABlockTransferSrcScalarPerVector = 0, index_t ABlockTransferDstScalarPerVector_AK1 = 0,
          index_t ABlockLdsExtraM = 0,
          typename BBlockTransferThreadClusterLengths_BK0_N_BK1 = void, typename BBlockTransferThreadClusterArrangeOrder = void,
          typename BBlockTransferSrcAccessOrder = void, index_t BBlockTransferSrcVectorDim = 0,
          index_t BBlockTransferSrcScalarPerVector = 0, index_t BBlockTransferDstScalarPerVector_BK1 = 0,
          index_t BBlockLdsExtraN = 0, index_t CShuffleMXdlPerWavePerShuffle = 0, index_t CShuffleNXdlPerWavePerShuffle = 0,
          typename CDEBlockTransferClusterLengths_MBlock_MPerBlock_NBlock_NPerBlock = void,
          index_t CDEBlockTransferScalarPerVector_NPerBlock = 0,
          LoopScheduler LoopSched = make_default_loop_scheduler(), PipelineVersion PipelineVer = PipelineVersion::v1,
          typename ComputeDataType = EDataType>
struct DeviceGemmMultipleD_Xdl_CShuffle
    : public DeviceGemmMultipleD_Sm100<ALayout, BLayout, DsLayout, ELayout, ADataType, BDataType, AccDataType, CShuffleDataType,
                                       DsDataType, EDataType, AElementwiseOperation, BElementwiseOperation,
                                       CDEElementwiseOperation, GemmSpec>
{
    const char* ClassName() const override { return "DeviceGemmMultipleD_Xdl_CShuffle"; }
};
} // namespace device
} // namespace tensor_operation
} // namespace ck
