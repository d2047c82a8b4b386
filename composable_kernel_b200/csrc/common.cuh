// common.cuh -- host-side helpers shared by the kernel launchers: error state, TMA descriptor
// encoding through the driver entry point, device properties.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/ckb.h"

namespace ckb {

std::string& last_error();
int set_error(int code, const char* fmt, ...);

#define CKB_CUDA_CHECK(expr)                                                                         \
    do {                                                                                             \
        cudaError_t _e = (expr);                                                                     \
        if(_e != cudaSuccess)                                                                        \
            return ::ckb::set_error(CKB_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                   \
                                    cudaGetErrorString(_e), __FILE__, __LINE__);                     \
    } while(0)

struct DeviceInfo
{
    int ok       = 0; // 1 = usable sm_100 device
    int sm_count = 0;
    int cc_major = 0, cc_minor = 0;
    int max_smem_optin = 0;
};
const DeviceInfo& device_info(); // of the CURRENT device (one record per device ordinal)

// Once-per-device guard for cudaFuncSetAttribute: function attributes are per device, so a process-wide flag set on the first
// device would leave the kernel without its shared-memory opt-in on the others.  Thread-safe; two racing threads may both set the
// (idempotent) attribute.
struct PerDeviceOnce
{
    std::atomic<uint64_t> done{0};
    bool need(uint64_t& bit)
    {
        int dev = 0;
        if(cudaGetDevice(&dev) != cudaSuccess) dev = 0;
        bit = uint64_t(1) << (dev & 63);
        return !(done.load(std::memory_order_acquire) & bit);
    }
    void mark(uint64_t bit) { done.fetch_or(bit, std::memory_order_release); }
};

// 2-D row-major tensor [rows][cols] with leading dimension ld (elements); box [box_rows][box_cols],
// 128-byte swizzle (box_cols * elem_bytes must be <= 128 for swizzled boxes; we always use == 128
// or a whole row of a narrow tensor), out-of-bounds elements read as zero.
int encode_tmap_2d(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, const void* base,
                   uint64_t rows, uint64_t cols, uint64_t ld_elems, uint32_t box_rows,
                   uint32_t box_cols, CUtensorMapSwizzle swz);

// general tiled map: dims/box innermost first, strides in bytes for dims 1..rank-1
int encode_tmap_nd(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz);

inline int dtype_bytes(int dt)
{
    switch(dt)
    {
    case CKB_F32: case CKB_I32: return 4;
    case CKB_F16: case CKB_BF16: return 2;
    case CKB_I8: case CKB_U8: return 1;
    default: return 0;
    }
}

} // namespace ckb
