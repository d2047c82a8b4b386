#pragma once
#include "ck/library/utility/host_tensor.hpp"
