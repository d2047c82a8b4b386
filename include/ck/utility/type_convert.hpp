#pragma once
#include "ck/utility/data_type.hpp" // conversions live next to the types in this build
