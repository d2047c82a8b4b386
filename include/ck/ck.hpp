// ck/ck.hpp -- basic definitions of the CK-compatible host API (B200 build).
// Mirrors include/ck/ck.hpp of the reference: ck::index_t = int32_t (ck.hpp:267), CK_ENABLE_* switches
// (include/ck/config.h.in:33-134), CK_LOGGING (include/ck/utility/env.hpp:89-140).
#pragma once
#include <cstdint>
#include <cstdlib>

#define CK_ENABLE_INT8 1
#define CK_ENABLE_FP16 1
#define CK_ENABLE_BF16 1
#define CK_TIME_KERNEL 1
#define CK_B200 1

namespace ck {
using index_t      = int32_t;
using long_index_t = int64_t;

inline bool EnvIsEnabled(const char* name)
{
    const char* v = std::getenv(name);
    return v && v[0] != '\0' && v[0] != '0';
}
#define CK_ENV(name) #name
} // namespace ck
