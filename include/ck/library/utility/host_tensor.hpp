// ck/library/utility/host_tensor.hpp -- HostTensorDescriptor / Tensor<T> with the reference's surface
// (library/include/ck/library/utility/host_tensor.hpp:87-168,248-518): offset = sum(idx * stride),
// GenerateTensorValue, mData / mDesc members, ParallelTensorFunctor over contiguous chunks (:188-245).
#pragma once
#include <algorithm>
#include <cassert>
#include <functional>
#include <iostream>
#include <numeric>
#include <thread>
#include <string>
#include <vector>

#include "ck/utility/data_type.hpp"

namespace ck {
namespace literals {
inline constexpr std::size_t operator""_uz(unsigned long long v) { return std::size_t(v); }
} // namespace literals
} // namespace ck

struct HostTensorDescriptor
{
    HostTensorDescriptor() = default;
    template <typename X>
    HostTensorDescriptor(const std::initializer_list<X>& lens) : mLens(lens.begin(), lens.end()) { CalculateStrides(); }
    template <typename X, typename Y>
    HostTensorDescriptor(const std::initializer_list<X>& lens, const std::initializer_list<Y>& strides)
        : mLens(lens.begin(), lens.end()), mStrides(strides.begin(), strides.end()) {}
    HostTensorDescriptor(const std::vector<std::size_t>& lens) : mLens(lens) { CalculateStrides(); }
    HostTensorDescriptor(const std::vector<std::size_t>& lens, const std::vector<std::size_t>& strides) : mLens(lens), mStrides(strides) {}

    void CalculateStrides()
    {
        mStrides.assign(mLens.size(), 1);
        for(int i = int(mLens.size()) - 2; i >= 0; --i) mStrides[i] = mStrides[i + 1] * mLens[i + 1];
    }
    std::size_t GetNumOfDimension() const { return mLens.size(); }
    std::size_t GetElementSize() const { return std::accumulate(mLens.begin(), mLens.end(), std::size_t(1), std::multiplies<>()); }
    std::size_t GetElementSpaceSize() const
    {
        std::size_t space = 1;
        for(std::size_t i = 0; i < mLens.size(); ++i)
        {
            if(mLens[i] == 0) continue;
            space += (mLens[i] - 1) * mStrides[i];
        }
        return space;
    }
    const std::vector<std::size_t>& GetLengths() const { return mLens; }
    const std::vector<std::size_t>& GetStrides() const { return mStrides; }
    template <typename... Is>
    std::size_t GetOffsetFromMultiIndex(Is... is) const
    {
        const std::size_t idx[] = {static_cast<std::size_t>(is)...};
        std::size_t off = 0;
        for(std::size_t i = 0; i < sizeof...(Is); ++i) off += idx[i] * mStrides[i];
        return off;
    }
    friend std::ostream& operator<<(std::ostream& os, const HostTensorDescriptor& d)
    {
        os << "dim " << d.mLens.size() << ", lengths {";
        for(std::size_t i = 0; i < d.mLens.size(); ++i) os << (i ? ", " : "") << d.mLens[i];
        os << "}, strides {";
        for(std::size_t i = 0; i < d.mStrides.size(); ++i) os << (i ? ", " : "") << d.mStrides[i];
        return os << "}";
    }
    std::vector<std::size_t> mLens, mStrides;
};

// contiguous chunks of the flattened index space over `num_thread` std::threads (host_tensor.hpp:219-239)
template <typename F>
struct ParallelTensorFunctor2D
{
    F f;
    std::size_t d0, d1;
    void operator()(std::size_t num_thread = 1) const
    {
        const std::size_t total = d0 * d1;
        num_thread              = std::max<std::size_t>(1, std::min(num_thread, total ? total : 1));
        const std::size_t per   = (total + num_thread - 1) / num_thread;
        std::vector<std::thread> th;
        for(std::size_t t = 0; t < num_thread; ++t)
        {
            const std::size_t b = t * per, e = std::min(total, b + per);
            if(b >= e) break;
            th.emplace_back([this, b, e] { for(std::size_t i = b; i < e; ++i) f(i / d1, i % d1); });
        }
        for(auto& x : th) x.join();
    }
};
template <typename F>
auto make_ParallelTensorFunctor(F f, std::size_t d0, std::size_t d1) { return ParallelTensorFunctor2D<F>{f, d0, d1}; }

template <typename T>
struct Tensor
{
    using Data = std::vector<T>;
    template <typename X>
    Tensor(std::initializer_list<X> lens) : mDesc(lens), mData(mDesc.GetElementSpaceSize()) {}
    Tensor(const HostTensorDescriptor& desc) : mDesc(desc), mData(mDesc.GetElementSpaceSize()) {}
    std::size_t GetElementSize() const { return mDesc.GetElementSize(); }
    std::size_t GetElementSpaceSize() const { return mDesc.GetElementSpaceSize(); }
    std::size_t GetElementSpaceSizeInBytes() const { return sizeof(T) * GetElementSpaceSize(); }
    void SetZero() { std::fill(mData.begin(), mData.end(), T{}); }
    template <typename G>
    void GenerateTensorValue(G g, std::size_t = 1)
    {
        // visit every logical element once in row-major index order (the reference's ParallelTensorFunctor order)
        const auto& L = mDesc.GetLengths();
        assert(L.size() == 2);
        for(std::size_t i = 0; i < L[0]; ++i)
            for(std::size_t j = 0; j < L[1]; ++j) (*this)(i, j) = g(i, j);
    }
    template <typename... Is>
    T& operator()(Is... is) { return mData[mDesc.GetOffsetFromMultiIndex(is...)]; }
    template <typename... Is>
    const T& operator()(Is... is) const { return mData[mDesc.GetOffsetFromMultiIndex(is...)]; }
    typename Data::iterator begin() { return mData.begin(); }
    typename Data::iterator end() { return mData.end(); }
    typename Data::const_iterator begin() const { return mData.begin(); }
    typename Data::const_iterator end() const { return mData.end(); }
    std::size_t size() const { return mData.size(); }
    T* data() { return mData.data(); }
    const T* data() const { return mData.data(); }
    HostTensorDescriptor mDesc;
    Data mData;
};

// Print a range converted to T with a delimiter (host_tensor.hpp LogRange / LogRangeAsType in the reference).
template <typename T, typename Range>
std::ostream& LogRangeAsType(std::ostream& os, Range&& range, const std::string& delim)
{
    bool first = true;
    for(auto&& v : range)
    {
        if(!first) os << delim;
        first = false;
        os << ck::type_convert<T>(v);
    }
    return os;
}
template <typename Range>
std::ostream& LogRange(std::ostream& os, Range&& range, const std::string& delim)
{
    bool first = true;
    for(auto&& v : range)
    {
        if(!first) os << delim;
        first = false;
        os << v;
    }
    return os;
}
