// ck::utils::Fill* (library/include/ck/library/utility/fill.hpp:19-172): every call seeds a FRESH std::mt19937(11939).
#pragma once
#include <algorithm>
#include <cmath>
#include <random>
#include "ck/utility/data_type.hpp"
namespace ck {
namespace utils {
template <typename T>
struct FillUniformDistribution
{
    float a_{-5.f};
    float b_{5.f};
    template <typename It>
    void operator()(It first, It last) const
    {
        std::mt19937 gen(11939);
        std::uniform_real_distribution<float> dis(a_, b_);
        std::generate(first, last, [&dis, &gen]() { return ck::type_convert<T>(dis(gen)); });
    }
    template <typename R>
    void operator()(R&& r) const { (*this)(std::begin(r), std::end(r)); }
};
template <typename T>
struct FillUniformDistributionIntegerValue
{
    float a_{-5.f};
    float b_{5.f};
    template <typename It>
    void operator()(It first, It last) const
    {
        std::mt19937 gen(11939);
        std::uniform_real_distribution<float> dis(a_, b_);
        std::generate(first, last, [&dis, &gen]() { return ck::type_convert<T>(std::round(dis(gen))); });
    }
    template <typename R>
    void operator()(R&& r) const { (*this)(std::begin(r), std::end(r)); }
};
template <typename T>
struct FillConstant
{
    T value_{0};
    template <typename It>
    void operator()(It first, It last) const { std::fill(first, last, value_); }
    template <typename R>
    void operator()(R&& r) const { (*this)(std::begin(r), std::end(r)); }
};
} // namespace utils
} // namespace ck
