#pragma once
#include "ck/tensor_operation/gpu/element/element_wise_operation.hpp"
