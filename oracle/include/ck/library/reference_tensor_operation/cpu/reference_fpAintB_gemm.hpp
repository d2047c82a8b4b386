#pragma once
#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
