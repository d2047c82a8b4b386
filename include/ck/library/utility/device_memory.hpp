// DeviceMem (library/include/ck/library/utility/device_memory.hpp:20-49) over the CUDA runtime.
#pragma once
#include <cuda_runtime_api.h>
#include <stdexcept>
#include <string>

struct DeviceMem
{
    DeviceMem() = default;
    explicit DeviceMem(std::size_t mem_size) : mMemSize(mem_size)
    {
        if(mem_size && cudaMalloc(&mpDeviceBuf, mem_size) != cudaSuccess) throw std::runtime_error("cudaMalloc failed");
    }
    DeviceMem(const DeviceMem&) = delete;
    DeviceMem& operator=(const DeviceMem&) = delete;
    void* GetDeviceBuffer() const { return mpDeviceBuf; }
    std::size_t GetBufferSize() const { return mMemSize; }
    void ToDevice(const void* p) const { check(cudaMemcpy(mpDeviceBuf, p, mMemSize, cudaMemcpyHostToDevice)); }
    void ToDevice(const void* p, std::size_t n) const { check(cudaMemcpy(mpDeviceBuf, p, n, cudaMemcpyHostToDevice)); }
    void FromDevice(void* p) const { check(cudaMemcpy(p, mpDeviceBuf, mMemSize, cudaMemcpyDeviceToHost)); }
    void FromDevice(void* p, std::size_t n) const { check(cudaMemcpy(p, mpDeviceBuf, n, cudaMemcpyDeviceToHost)); }
    void SetZero() const { if(mpDeviceBuf) check(cudaMemset(mpDeviceBuf, 0, mMemSize)); }
    ~DeviceMem() { if(mpDeviceBuf) cudaFree(mpDeviceBuf); }
    static void check(cudaError_t e) { if(e != cudaSuccess) throw std::runtime_error(std::string("CUDA: ") + cudaGetErrorString(e)); }
    void* mpDeviceBuf   = nullptr;
    std::size_t mMemSize = 0;
};
