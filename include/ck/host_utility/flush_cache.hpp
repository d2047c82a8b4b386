// ck/host_utility/flush_cache.hpp -- cold-cache timing: rotating operand buffers + cache flush, B200 build.
// Role of include/ck/host_utility/flush_cache.hpp:18-235.  Why it matters here: B200 has a 126 MB L2, so a decode-shaped GEMM
// (weights of 8 - 120 MB) timed in a tight loop runs out of L2 and reports an L2 number, not the HBM number a serving step sees.
//   RotatingMem          (RotatingMemWrapper, :18-95)   rotating_count device copies of the A and B buffers; Next() points the
//                        argument at the next copy before each timed launch, the destructor restores the caller's pointers
//   flush_cache          (flush_icache, :97-105: the reference flushes the instruction cache of its AMD part; the cache that
//                        decides timings on this part is the L2) overwrites a scratch buffer larger than L2
//   launch_and_time_kernel_with_preprocess<TimePreprocess> (:107-235)  warm-ups, then nrepeat_ launches each timed with its own
//                        event pair after the preprocess step, result = the MEDIAN (the reference's `#define MEDIAN 1`)
// Used by GemmInvoker::Run when StreamConfig::flush_cache is set (StreamConfig::rotating_count copies).
#pragma once
#include <cuda_runtime_api.h>

#include <algorithm>
#include <cstdio>
#include <functional>
#include <vector>

#include "ck/host_utility/kernel_launch.hpp"

namespace ck {
namespace utility {

struct RotatingMem
{
    // `set` is called with (p_a, p_b) whenever the operand pointers change (the argument re-plans: TMA descriptors embed pointers)
    RotatingMem(const void* p_a, const void* p_b, std::size_t size_a, std::size_t size_b, std::size_t rotating_count,
                std::function<void(const void*, const void*)> set)
        : size_a_(size_a), size_b_(size_b), rotating_count_(std::max<std::size_t>(1, rotating_count)), set_(std::move(set))
    {
        p_a_.push_back(p_a);
        p_b_.push_back(p_b);
        for(std::size_t i = 1; i < rotating_count_; ++i)
        {
            void *a = nullptr, *b = nullptr;
            cuda_check_error(cudaMalloc(&a, std::max<std::size_t>(size_a, 16)), "cudaMalloc(rotating A)");
            cuda_check_error(cudaMalloc(&b, std::max<std::size_t>(size_b, 16)), "cudaMalloc(rotating B)");
            cuda_check_error(cudaMemcpy(a, p_a, size_a, cudaMemcpyDeviceToDevice));
            cuda_check_error(cudaMemcpy(b, p_b, size_b, cudaMemcpyDeviceToDevice));
            p_a_.push_back(a);
            p_b_.push_back(b);
        }
    }
    void Next()
    {
        if(rotating_count_ > 1)
        {
            const std::size_t idx = iter_++ % rotating_count_;
            set_(p_a_[idx], p_b_[idx]);
        }
    }
    void Print() const
    {
        std::printf("RotatingMemWrapper: { size_a: %zu, size_b: %zu, rotating_count: %zu}\n", size_a_, size_b_, rotating_count_);
    }
    ~RotatingMem()
    {
        if(rotating_count_ > 1)
        {
            set_(p_a_[0], p_b_[0]);
            for(std::size_t i = 1; i < rotating_count_; ++i)
            {
                cudaFree(const_cast<void*>(p_a_[i]));
                cudaFree(const_cast<void*>(p_b_[i]));
            }
        }
    }
    RotatingMem(const RotatingMem&)            = delete;
    RotatingMem& operator=(const RotatingMem&) = delete;

    private:
    std::size_t iter_ = 0, size_a_, size_b_, rotating_count_;
    std::function<void(const void*, const void*)> set_;
    std::vector<const void*> p_a_, p_b_;
};

// evict the L2: overwrite a scratch buffer 1.5x its size (allocated once per process)
inline void flush_cache(cudaStream_t stream = nullptr)
{
    static void* scratch     = nullptr;
    static std::size_t bytes = 0;
    if(!scratch)
    {
        int dev = 0, l2 = 0;
        cuda_check_error(cudaGetDevice(&dev));
        cuda_check_error(cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev));
        bytes = std::max<std::size_t>(std::size_t(l2) * 3 / 2, std::size_t(64) << 20);
        cuda_check_error(cudaMalloc(&scratch, bytes), "cudaMalloc(L2 flush scratch)");
    }
    cuda_check_error(cudaMemsetAsync(scratch, 0, bytes, stream), "cudaMemsetAsync(L2 flush)");
}
inline void flush_icache() { flush_cache(nullptr); } // the reference's name

template <bool TimePreprocess, typename Launch, typename Preprocess>
float launch_and_time_kernel_with_preprocess(const StreamConfig& sc, Preprocess&& preprocess, Launch&& launch)
{
    if(!sc.time_kernel_)
    {
        preprocess();
        launch();
        return 0.f;
    }
    for(int i = 0; i < sc.cold_niters_; ++i) launch();
    if(sc.nrepeat_ == 0) return 0.f;
    std::vector<float> times;
    cudaEvent_t start, stop;
    cuda_check_error(cudaEventCreate(&start));
    cuda_check_error(cudaEventCreate(&stop));
    for(int i = 0; i < sc.nrepeat_; ++i)
    {
        if constexpr(!TimePreprocess) preprocess();
        cuda_check_error(cudaDeviceSynchronize());
        cuda_check_error(cudaEventRecord(start, sc.stream_id_));
        if constexpr(TimePreprocess) preprocess();
        launch();
        cuda_check_error(cudaEventRecord(stop, sc.stream_id_));
        cuda_check_error(cudaEventSynchronize(stop));
        float t = 0.f;
        cuda_check_error(cudaEventElapsedTime(&t, start, stop));
        times.push_back(t);
        if(EnvIsEnabled(CK_ENV(CK_LOGGING))) std::printf("i: %d cur_time: %f\n", i, t);
    }
    cudaEventDestroy(start);
    cudaEventDestroy(stop);
    std::sort(times.begin(), times.end());
    const std::size_t n = times.size(), mid = (n - 1) / 2;
    return n % 2 == 1 ? times[mid] : 0.5f * (times[mid] + times[mid + 1]);
}

} // namespace utility
} // namespace ck
