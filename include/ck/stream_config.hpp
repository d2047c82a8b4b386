// ck/stream_config.hpp -- same aggregate as the reference (include/ck/stream_config.hpp:9-19), CUDA stream.
#pragma once
#include <cuda_runtime_api.h>

struct StreamConfig
{
    cudaStream_t stream_id_ = nullptr;
    bool time_kernel_       = false;
    int log_level_          = 0;
    int cold_niters_        = 5;
    int nrepeat_            = 50;

    bool flush_cache   = false;
    int rotating_count = 1;
};
