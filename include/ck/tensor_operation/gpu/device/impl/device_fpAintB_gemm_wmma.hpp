// reference path kept for source compatibility; the sm_100a implementation lives in device_gemm_sm100.hpp
#pragma once
#include "ck/tensor_operation/gpu/device/impl/device_gemm_sm100.hpp"
