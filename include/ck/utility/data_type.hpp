// ck/utility/data_type.hpp + type_convert.hpp -- host-side scalar types and conversions with CK's exact
// bit rules (include/ck/utility/data_type.hpp:10-14, include/ck/utility/type_convert.hpp:17-100,508-549):
//   half_t = _Float16, bhalf_t = unsigned short (a plain integer type!), fp32 -> bf16 TRUNCATES.
#pragma once
#include <cstdint>
#include <cstring>
#include <type_traits>

#include "ck/ck.hpp"

namespace ck {
using half_t  = _Float16;
using bhalf_t = unsigned short;
using int4_t  = int8_t; // unpacked, sizeof == 1 as in the reference (run_gemm_example.inc:8-10)

template <typename Y, typename X>
inline constexpr bool is_same_v = std::is_same<Y, X>::value;

template <typename Y, typename X>
inline Y type_convert(X x)
{
    using NY = std::remove_const_t<Y>;
    using NX = std::remove_const_t<X>;
    if constexpr(std::is_same_v<NY, NX> && !std::is_same_v<NY, bhalf_t>) return x;
    else if constexpr(std::is_same_v<NY, float> && std::is_same_v<NX, bhalf_t>)
    {
        uint32_t u = uint32_t(x) << 16; // type_convert.hpp:42-52
        float f;
        std::memcpy(&f, &u, 4);
        return f;
    }
    else if constexpr(std::is_same_v<NY, bhalf_t> && std::is_same_v<NX, float>)
    {
        uint32_t u; // type_convert.hpp:53-64: truncation
        std::memcpy(&u, &x, 4);
        return bhalf_t(u >> 16);
    }
    else if constexpr(std::is_same_v<NY, half_t> && std::is_same_v<NX, bhalf_t>)
        return static_cast<half_t>(type_convert<float>(x));
    else if constexpr(std::is_same_v<NY, bhalf_t> && std::is_same_v<NX, half_t>)
        return type_convert<bhalf_t>(static_cast<float>(x));
    else if constexpr(std::is_same_v<NY, int8_t> && std::is_same_v<NX, bhalf_t>)
        return static_cast<int8_t>(type_convert<float>(x));
    else if constexpr(std::is_same_v<NY, bhalf_t> && std::is_same_v<NX, int8_t>)
        return type_convert<bhalf_t>(static_cast<float>(x));
    else
        return static_cast<NY>(x);
}

template <typename Y, typename X>
inline Y bf16_convert_rtn(X x);
template <>
inline bhalf_t bf16_convert_rtn<bhalf_t, float>(float x)
{
    uint32_t u; // type_convert.hpp:508-549
    std::memcpy(&u, &x, 4);
    const bool flag0 = (~u & 0x7f800000u) != 0;
    const bool flag1 = !flag0 && (u & 0xffffu);
    u += flag0 ? 0x7fffu + ((u >> 16) & 1u) : 0u;
    u |= flag1 ? 0x10000u : 0u;
    return bhalf_t(u >> 16);
}
template <>
inline bhalf_t bf16_convert_rtn<bhalf_t, half_t>(half_t x) { return bf16_convert_rtn<bhalf_t>(static_cast<float>(x)); }

// ---- minimal compile-time containers so client code compiles unchanged -----------------------
template <typename... Ts>
struct Tuple
{
    static constexpr index_t Size() { return sizeof...(Ts); }
};
template <index_t... Is>
struct Sequence
{
    static constexpr index_t Size() { return sizeof...(Is); }
};
template <index_t N>
struct Number
{
    static constexpr index_t value = N;
    constexpr operator index_t() const { return N; }
};

// element type i of a ck::Tuple
template <index_t I, typename T>
struct tuple_element;
template <index_t I, typename H, typename... Ts>
struct tuple_element<I, Tuple<H, Ts...>> : tuple_element<I - 1, Tuple<Ts...>> {};
template <typename H, typename... Ts>
struct tuple_element<0, Tuple<H, Ts...>> { using type = H; };
template <index_t I, typename T>
using tuple_element_t = typename tuple_element<I, T>::type;

enum struct BlockGemmPipelineScheduler { Intrawave, Interwave };
enum struct BlockGemmPipelineVersion { v1, v2, v3, v4, v5 };
enum struct PipelineVersion { v1, v2, v4, weight_only };
enum struct LoopScheduler { Default, Interwave };
constexpr LoopScheduler make_default_loop_scheduler() { return LoopScheduler::Default; } // gridwise_gemm_pipeline_selector.hpp
} // namespace ck
