// common.cu -- error state, device probing, TMA descriptor encoding.
#include "common.cuh"

#include <mutex>

namespace ckb {

std::string& last_error()
{
    static thread_local std::string s;
    return s;
}

int set_error(int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error() = buf;
    return code;
}

const DeviceInfo& device_info()
{
    static DeviceInfo infos[64];
    static std::once_flag once[64];
    static DeviceInfo none;
    int n = 0, dev = 0;
    if(cudaGetDeviceCount(&n) != cudaSuccess || n == 0)
    {
        cudaGetLastError();
        return none;
    }
    if(cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return none;
    std::call_once(once[dev], [dev] {
        DeviceInfo& info = infos[dev];
        cudaDeviceProp p{};
        if(cudaGetDeviceProperties(&p, dev) != cudaSuccess) return;
        info.sm_count = p.multiProcessorCount;
        info.cc_major = p.major;
        info.cc_minor = p.minor;
        cudaDeviceGetAttribute(&info.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        info.ok = (p.major == 10);
    });
    return infos[dev];
}

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
               cudaSuccess &&
           qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        else
            cudaGetLastError();
    });
    return fn;
}

int encode_tmap_2d(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, const void* base,
                   uint64_t rows, uint64_t cols, uint64_t ld_elems, uint32_t box_rows,
                   uint32_t box_cols, CUtensorMapSwizzle swz)
{
    EncodeTiledFn fn = get_encode_fn();
    if(!fn) return set_error(CKB_ERR_NO_DEVICE, "cuTensorMapEncodeTiled entry point unavailable");
    if((reinterpret_cast<uintptr_t>(base) & 15) != 0)
        return set_error(CKB_ERR_UNSUPPORTED, "TMA base pointer must be 16-byte aligned");
    if((ld_elems * elem_bytes) % 16 != 0)
        return set_error(CKB_ERR_UNSUPPORTED, "TMA row pitch must be a multiple of 16 bytes");
    cuuint64_t gdim[2]    = {cols, rows};
    cuuint64_t gstride[1] = {ld_elems * uint64_t(elem_bytes)};
    cuuint32_t box[2]     = {box_cols, box_rows};
    cuuint32_t estr[2]    = {1, 1};
    CUresult r = fn(out, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if(r != CUDA_SUCCESS)
        return set_error(CKB_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d "
                         "(rows %llu cols %llu ld %llu box %ux%u)", int(r),
                         (unsigned long long)rows, (unsigned long long)cols,
                         (unsigned long long)ld_elems, box_rows, box_cols);
    return CKB_OK;
}

int encode_tmap_nd(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz)
{
    EncodeTiledFn fn = get_encode_fn();
    if(!fn) return set_error(CKB_ERR_NO_DEVICE, "cuTensorMapEncodeTiled entry point unavailable");
    if((reinterpret_cast<uintptr_t>(base) & 15) != 0)
        return set_error(CKB_ERR_UNSUPPORTED, "TMA base pointer must be 16-byte aligned");
    cuuint64_t gdim[5], gstride[4];
    cuuint32_t bx[5], estr[5];
    for(int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; estr[i] = 1; }
    for(int i = 0; i + 1 < rank; ++i)
    {
        if(strides_bytes[i] % 16 != 0) return set_error(CKB_ERR_UNSUPPORTED, "TMA stride must be a multiple of 16 bytes");
        gstride[i] = strides_bytes[i];
    }
    CUresult r = fn(out, dt, rank, const_cast<void*>(base), gdim, gstride, bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if(r != CUDA_SUCCESS) return set_error(CKB_ERR_CUDA, "cuTensorMapEncodeTiled (rank %d) failed with CUresult %d", rank, int(r));
    return CKB_OK;
}

} // namespace ckb
