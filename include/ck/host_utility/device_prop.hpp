// ck/host_utility/device_prop.hpp -- device identification and architecture gates, B200 build.
// Role of include/ck/host_utility/device_prop.hpp:12-90 (get_device_name() maps the HIP gcnArchName to a gfx name and the
// is_*_supported() predicates gate IsSupportedArgument): here the name is "sm_<major><minor>" and the one gate that matters is
// is_sm100_supported() -- tcgen05 / TMEM / TMA kernels compiled for sm_100a only run on compute capability 10.x.
#pragma once
#include <cuda_runtime_api.h>

#include <string>

namespace ck {
inline std::string get_device_name()
{
    int dev = 0;
    cudaDeviceProp props{};
    if(cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&props, dev) != cudaSuccess) return std::string();
    return "sm_" + std::to_string(props.major) + std::to_string(props.minor);
}
inline int get_sm_count()
{
    int dev = 0, n = 0;
    if(cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return n;
}
inline bool is_sm100_supported()
{
    const std::string n = get_device_name();
    return n == "sm_100" || n == "sm_101" || n == "sm_103";
}
inline bool is_tcgen05_supported() { return is_sm100_supported(); }
// the reference's AMD gates: never true on this build (client code that consults them falls through to "not supported")
inline bool is_xdl_supported() { return false; }
inline bool is_lds_direct_load_supported() { return false; }
inline bool is_gfx101_supported() { return false; }
inline bool is_gfx103_supported() { return false; }
inline bool is_gfx11_supported() { return false; }
inline bool is_gfx12_supported() { return false; }
} // namespace ck
