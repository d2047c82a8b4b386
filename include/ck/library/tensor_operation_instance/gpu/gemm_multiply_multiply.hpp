#pragma once
#include "ck/library/tensor_operation_instance/gpu/gemm_quantization.hpp"
