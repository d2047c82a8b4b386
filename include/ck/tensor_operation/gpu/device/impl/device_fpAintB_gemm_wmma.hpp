// SPDX-License-Identifier: MIT
// The class name and template parameter list mirror Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence)
// so that client code compiles unchanged; the implementation is the B200 one (device_gemm_sm100.hpp).
// DeviceFpAintBGemm_Wmma_CShuffle with the reference's full template parameter list
// (include/ck/tensor_operation/gpu/device/impl/device_fpAintB_gemm_wmma.hpp:27-70).  Everything after GemmSpec describes the
// RDNA WMMA tiling; on sm_100a the tile shape is chosen per problem by ckb_gemm_query, so these are accepted and ignored
// (all defaulted, which also keeps the short form <layouts, types, ops[, GemmSpec]> valid).
#pragma once
#include "ck/tensor_operation/gpu/device/impl/device_gemm_sm100.hpp"
namespace ck {
namespace tensor_operation {
namespace device {
template <typename ALayout, typename BLayout, typename CLayout, typename ADataType, typename BDataType,
          typename ScaleDataType, typename CDataType, typename AccDataType, typename CShuffleDataType,
          typename AElementwiseOperation, typename BElementwiseOperation, typename CElementwiseOperation,
          GemmSpecialization GemmSpec = GemmSpecialization::Default,
          ck::index_t NumPrefetch = 1, ck::index_t BlockSize = 0, ck::index_t MPerBlock = 0, ck::index_t NPerBlock = 0,
          ck::index_t KPerBlock = 0, ck::index_t K1 = 0, ck::index_t MPerWmma = 0, ck::index_t NPerWmma = 0,
          ck::index_t MRepeat = 0, ck::index_t NRepeat = 0,
          typename ABlockTransferThreadClusterLengths_K0_M_K1 = void, typename ABlockTransferThreadClusterArrangeOrder = void,
          typename ABlockTransferSrcAccessOrder = void, ck::index_t ABlockTransferSrcVectorDim = 0,
          ck::index_t ABlockTransferSrcScalarPerVector = 0, ck::index_t ABlockTransferDstScalarPerVector_K1 = 0,
          bool ABlockLdsAddExtraM = false,
          typename BBlockTransferThreadClusterLengths_K0_N_K1 = void, typename BBlockTransferThreadClusterArrangeOrder = void,
          typename BBlockTransferSrcAccessOrder = void, ck::index_t BBlockTransferSrcVectorDim = 0,
          ck::index_t BBlockTransferSrcScalarPerVector = 0, ck::index_t BBlockTransferDstScalarPerVector_K1 = 0,
          bool BBlockLdsAddExtraN = false, index_t CShuffleMRepeatPerShuffle = 0, index_t CShuffleNRepeatPerShuffle = 0,
          typename CShuffleBlockTransferClusterLengths_MBlock_MPerBlock_NBlock_NPerBlock = void,
          index_t CShuffleBlockTransferScalarPerVector_NPerBlock = 0,
          ck::LoopScheduler LoopSched = make_default_loop_scheduler(),
          ck::PipelineVersion PipelineVer = ck::PipelineVersion::weight_only>
struct DeviceFpAintBGemm_Wmma_CShuffle
    : public DeviceFpAintBGemm_Sm100<ALayout, BLayout, CLayout, ADataType, BDataType, ScaleDataType, CDataType, AccDataType,
                                     CShuffleDataType, AElementwiseOperation, BElementwiseOperation, CElementwiseOperation, GemmSpec>
{
    const char* ClassName() const override { return "DeviceFpAintBGemm_Wmma_CShuffle"; }
};
} // namespace device
} // namespace tensor_operation
} // namespace ck
