// GeneratorTensor_* (library/include/ck/library/utility/host_tensor_generator.hpp:13-287); _2 / _3 draw from
// the un-seeded std::rand() exactly like the reference (:93).
#pragma once
#include <cstdlib>
#include "ck/utility/data_type.hpp"
template <typename T>
struct GeneratorTensor_1
{
    T value = 1;
    template <typename... Is>
    T operator()(Is...) { return value; }
};
template <typename T>
struct GeneratorTensor_2
{
    int min_value = 0;
    int max_value = 1;
    template <typename... Is>
    T operator()(Is...)
    {
        const float tmp = float((std::rand() % (max_value - min_value)) + min_value);
        return ck::type_convert<T>(tmp);
    }
};
template <typename T>
struct GeneratorTensor_3
{
    float min_value = 0;
    float max_value = 1;
    template <typename... Is>
    T operator()(Is...)
    {
        const float tmp = float(std::rand()) / float(RAND_MAX);
        return ck::type_convert<T>(min_value + tmp * (max_value - min_value));
    }
};
