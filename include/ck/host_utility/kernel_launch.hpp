// ck/host_utility/kernel_launch.hpp -- launch-and-time helpers, B200 build.
// Role of include/ck/host_utility/kernel_launch.hpp:13-162 (launch_and_time_kernel / _with_preprocess): warm up cold_niters_
// times, time nrepeat_ launches between two events on stream_id_, return the MEAN per launch in ms; with time_kernel_ == false
// launch once and return 0.  The reference templates take a __global__ function plus its arguments; kernels here sit behind the
// C-ABI, so the callable is any `void()` that enqueues the work on stream_config.stream_id_ (GemmInvoker passes ckb_gemm_run).
#pragma once
#include <cuda_runtime_api.h>

#include <cstdio>
#include <stdexcept>
#include <string>

#include "ck/ck.hpp"
#include "ck/stream_config.hpp"

namespace ck {
inline void cuda_check_error(cudaError_t e, const char* what = "CUDA runtime")
{
    if(e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}

template <typename Launch, typename Preprocess>
float launch_and_time_kernel_with_preprocess(const StreamConfig& sc, Preprocess&& preprocess, Launch&& launch)
{
    if(!sc.time_kernel_)
    {
        preprocess();
        launch();
        return 0.f;
    }
    if(EnvIsEnabled(CK_ENV(CK_LOGGING))) std::printf("Warm up %d times\n", sc.cold_niters_);
    preprocess();
    for(int i = 0; i < sc.cold_niters_; ++i) launch();
    if(sc.nrepeat_ == 0) return 0.f;
    if(EnvIsEnabled(CK_ENV(CK_LOGGING))) std::printf("Start running %d times...\n", sc.nrepeat_);
    cudaEvent_t start, stop;
    cuda_check_error(cudaEventCreate(&start));
    cuda_check_error(cudaEventCreate(&stop));
    cuda_check_error(cudaDeviceSynchronize());
    cuda_check_error(cudaEventRecord(start, sc.stream_id_));
    for(int i = 0; i < sc.nrepeat_; ++i)
    {
        preprocess();
        launch();
    }
    cuda_check_error(cudaEventRecord(stop, sc.stream_id_));
    cuda_check_error(cudaEventSynchronize(stop));
    float ms = 0.f;
    cuda_check_error(cudaEventElapsedTime(&ms, start, stop));
    cudaEventDestroy(start);
    cudaEventDestroy(stop);
    return ms / float(sc.nrepeat_);
}

template <typename Launch>
float launch_and_time_kernel(const StreamConfig& sc, Launch&& launch)
{
    return launch_and_time_kernel_with_preprocess(sc, [] {}, launch);
}
} // namespace ck
