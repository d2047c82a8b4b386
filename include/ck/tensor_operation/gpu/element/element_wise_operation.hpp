// ck/tensor_operation/gpu/element/element_wise_operation.hpp -- the A/B/CDE element-wise functors of the
// mixed-precision GEMM path, same names and host arithmetic as the reference:
//   PassThrough, Scale, Relu, FastGelu           element/unary_element_wise_operation.hpp:28-225,289-332,409,435-548
//   Add, Multiply, AddFastGelu, MultiplyFastGelu  element/binary_element_wise_operation.hpp:13-93,117-207,461-589
//   MultiplyAdd, MultiplyMultiply, MultiplyAddFastGelu   element/element_wise_operation.hpp:200-296
//   Activation_Mul_Clamp, Add_Activation_Mul_Clamp, Activation_Mul2_Clamp, Add_Activation_Mul2_Clamp
//                                                 element/quantization_operation.hpp:25-201
// The reference spells every dtype combination as an in-class explicit specialisation (clang-only);
// here one templated operator() dispatches with `if constexpr`.  On the device the functor is only a TAG:
// kCkbCde / kCkbBOp select the fused prologue/epilogue inside the sm_100a kernels (include/ckb.h).
#pragma once
#include <algorithm>
#include <cmath>

#include "ck/utility/data_type.hpp"
#include "ckb.h"

namespace ck {
namespace tensor_operation {
namespace element_wise {

namespace detail {
template <typename T>
inline float to_f32(const T& x)
{
    if constexpr(std::is_same_v<T, bhalf_t>) return type_convert<float>(x);
    else return static_cast<float>(x);
}
template <typename Y>
inline void from_f32(Y& y, float v)
{
    y = type_convert<Y>(v);
}
} // namespace detail

struct PassThrough
{
    static constexpr int kCkbBOp = CKB_BOP_PASSTHROUGH;
    static constexpr int kCkbCde = CKB_CDE_PASSTHROUGH;
    template <typename Y, typename X>
    void operator()(Y& y, const X& x) const
    {
        if constexpr(std::is_same_v<Y, bhalf_t> && std::is_same_v<X, bhalf_t>) y = x;
        else y = type_convert<Y>(x);
    }
};
struct ConvertBF16RTN
{
    template <typename Y, typename X>
    void operator()(Y& y, const X& x) const { y = bf16_convert_rtn<Y>(x); }
};
struct Scale
{
    explicit Scale(float s = 1.f) : scale_(s) {}
    template <typename Y, typename X>
    void operator()(Y& y, const X& x) const { detail::from_f32(y, scale_ * detail::to_f32(x)); }
    float scale_;
};
struct Relu
{
    static constexpr int kCkbAct = CKB_ACT_RELU;
    template <typename T>
    void operator()(T& y, const T& x) const
    {
        if constexpr(std::is_same_v<T, bhalf_t>) { const float f = type_convert<float>(x); y = type_convert<bhalf_t>(f > 0 ? f : 0.f); }
        else y = x > 0 ? x : T(0);
    }
};
struct FastGelu
{
    static constexpr int kCkbCde = CKB_CDE_FASTGELU;
    // y = x / (1 + exp(x * (-0.071354 x^2 - 1.59577)))   (host form: exp + true divide, :443-452)
    static float gelu(float x)
    {
        const float c1  = -2.0 * 0.035677f;
        const float c2  = -2.0 * 0.797885f;
        const float u   = x * (c1 * x * x + c2);
        const float emu = std::exp(u);
        return x / (1.f + emu);
    }
    template <typename Y, typename X>
    void operator()(Y& y, const X& x) const { detail::from_f32(y, gelu(detail::to_f32(x))); }
};

struct Add
{
    static constexpr int kCkbCde = CKB_CDE_ADD;
    template <typename Y, typename X0, typename X1>
    void operator()(Y& y, const X0& x0, const X1& x1) const
    {
        if constexpr(std::is_same_v<Y, half_t> && std::is_same_v<X0, float> && std::is_same_v<X1, half_t>)
            y = type_convert<half_t>(x0) + x1; // :47-52
        else if constexpr(std::is_same_v<Y, int8_t>) y = x0 + x1;
        else detail::from_f32(y, detail::to_f32(x0) + detail::to_f32(x1));
    }
};
struct Multiply
{
    static constexpr int kCkbBOp = CKB_BOP_MULTIPLY; // as B functor: Multiply<bhalf_t,int8_t,bhalf_t> :184-190
    static constexpr int kCkbCde = CKB_CDE_MULTIPLY; // as CDE functor: E = C * D0
    template <typename Y, typename X0, typename X1>
    void operator()(Y& y, const X0& x0, const X1& x1) const
    {
        if constexpr(std::is_same_v<Y, half_t> && std::is_same_v<X0, float> && std::is_same_v<X1, half_t>)
            y = type_convert<half_t>(x0) * x1; // :153-158
        else if constexpr(std::is_same_v<Y, int8_t>) y = x0 * x1;
        else detail::from_f32(y, detail::to_f32(x0) * detail::to_f32(x1));
    }
};
struct AddFastGelu
{
    static constexpr int kCkbCde = CKB_CDE_ADD_FASTGELU;
    template <typename E, typename C, typename D>
    void operator()(E& e, const C& c, const D& d) const { detail::from_f32(e, FastGelu::gelu(detail::to_f32(c) + detail::to_f32(d))); }
};
struct MultiplyFastGelu
{
    static constexpr int kCkbCde = CKB_CDE_MULTIPLY_FASTGELU;
    template <typename E, typename C, typename D>
    void operator()(E& e, const C& c, const D& d) const { detail::from_f32(e, FastGelu::gelu(detail::to_f32(c) * detail::to_f32(d))); }
};
struct MultiplyAdd
{
    // element_wise_operation.hpp:224-231 multiplies raw ushort bit patterns for bf16 (a reference
    // bug); this build implements the intended c*d0 + d1 in float, like the fp16/float variants :232-249.
    static constexpr int kCkbCde = CKB_CDE_MULTIPLY_ADD;
    template <typename E, typename C, typename D0, typename D1>
    void operator()(E& e, const C& c, const D0& d0, const D1& d1) const
    {
        detail::from_f32(e, detail::to_f32(c) * detail::to_f32(d0) + detail::to_f32(d1));
    }
};
struct MultiplyMultiply
{
    static constexpr int kCkbCde = CKB_CDE_MULTIPLY_MULTIPLY;
    template <typename E, typename C, typename D0, typename D1>
    void operator()(E& e, const C& c, const D0& d0, const D1& d1) const
    {
        const float x0_f = detail::to_f32(c) * detail::to_f32(d0) * detail::to_f32(d1); // :258-274
        detail::from_f32(e, x0_f);
    }
};
struct MultiplyAddFastGelu
{
    static constexpr int kCkbCde = CKB_CDE_MULTIPLY_ADD_FASTGELU;
    template <typename E, typename C, typename D0, typename D1>
    void operator()(E& e, const C& c, const D0& d0, const D1& d1) const
    {
        detail::from_f32(e, FastGelu::gelu(detail::to_f32(c) * detail::to_f32(d0) + detail::to_f32(d1)));
    }
};

namespace detail {
template <typename Act> struct act_tag { static constexpr int value = CKB_ACT_PASSTHROUGH; };
template <> struct act_tag<Relu> { static constexpr int value = CKB_ACT_RELU; };
inline float clamp_i8(float v) { return std::min(std::max(v, -128.f), 127.f); }
} // namespace detail

template <typename Activation>
struct Activation_Mul_Clamp
{
    static constexpr int kCkbCde = CKB_CDE_ACT_MUL_CLAMP;
    static constexpr int kCkbAct = detail::act_tag<Activation>::value;
    Activation_Mul_Clamp(float requantScale, Activation activationOp) : requantScale_(requantScale), activationOp_(activationOp) {}
    template <typename Y>
    void operator()(Y& y, const int32_t& x) const
    {
        float y_fp32 = type_convert<float>(x);
        activationOp_(y_fp32, y_fp32);
        y_fp32 = detail::clamp_i8(requantScale_ * y_fp32);
        y      = type_convert<Y>(y_fp32);
    }
    float ckb_scale() const { return requantScale_; }
    float requantScale_;
    Activation activationOp_;
};
template <typename Activation>
struct Add_Activation_Mul_Clamp
{
    static constexpr int kCkbCde = CKB_CDE_ADD_ACT_MUL_CLAMP;
    static constexpr int kCkbAct = detail::act_tag<Activation>::value;
    Add_Activation_Mul_Clamp(float requantScale, Activation activationOp) : requantScale_(requantScale), activationOp_(activationOp) {}
    template <typename Y>
    void operator()(Y& y, const int32_t& x, const int32_t& bias) const
    {
        float y_fp32 = type_convert<float>(x + bias);
        activationOp_(y_fp32, y_fp32);
        y_fp32 = detail::clamp_i8(requantScale_ * y_fp32);
        y      = type_convert<Y>(y_fp32);
    }
    float ckb_scale() const { return requantScale_; }
    float requantScale_;
    Activation activationOp_;
};
template <typename Activation>
struct Activation_Mul2_Clamp
{
    static constexpr int kCkbCde = CKB_CDE_ACT_MUL2_CLAMP;
    static constexpr int kCkbAct = detail::act_tag<Activation>::value;
    explicit Activation_Mul2_Clamp(Activation activationOp) : activationOp_(activationOp) {}
    template <typename Y>
    void operator()(Y& y, const int32_t& x, const float& requantScale) const
    {
        float y_fp32 = type_convert<float>(x);
        activationOp_(y_fp32, y_fp32);
        y = type_convert<Y>(detail::clamp_i8(requantScale * y_fp32));
    }
    Activation activationOp_;
};
template <typename Activation>
struct Add_Activation_Mul2_Clamp
{
    static constexpr int kCkbCde = CKB_CDE_ADD_ACT_MUL2_CLAMP;
    static constexpr int kCkbAct = detail::act_tag<Activation>::value;
    explicit Add_Activation_Mul2_Clamp(Activation activationOp) : activationOp_(activationOp) {}
    template <typename Y>
    void operator()(Y& y, const int32_t& x, const int32_t& bias, const float& requantScale) const
    {
        float y_fp32 = type_convert<float>(x + bias);
        activationOp_(y_fp32, y_fp32);
        y = type_convert<Y>(detail::clamp_i8(requantScale * y_fp32));
    }
    Activation activationOp_;
};

// [EXT] B functors of the W4A16 path.  Two numerics contracts, each valid at every M (the device op never switches between
// them by problem size):
//  * DequantizeGroup -- ab_scale pattern (device/device_gemm_multiple_d_ab_scale.hpp:22-59; host loop
//    example/65_gemm_multiply_multiply/gemm_multiply_multiply_xdl_fp8_ab_scale.cpp:258-290): the operand is the exact
//    (q - zero[k/G,n]); scale[k/G,n] multiplies in float.  The functor returns float.
//  * DequantizeGroupRound -- dequant-in-B-copy pattern (threadwise_tensor_slice_transfer_v3r1_dequant.hpp:509-547): the
//    operand is (q - zero) * scale rounded once (nearest-even) into A's type.
struct DequantizeGroup
{
    static constexpr int kCkbBOp = CKB_BOP_GROUP;
    void operator()(float& y, const int& q_minus_zero, const float& scale) const { y = float(q_minus_zero) * scale; }
    template <typename Y>
    void operator()(float& y, const int& q_minus_zero, const Y& scale) const
    {
        y = float(q_minus_zero) * detail::to_f32(scale);
    }
};
struct DequantizeGroupRound
{
    static constexpr int kCkbBOp = CKB_BOP_GROUP_RTN;
    template <typename Y>
    void operator()(Y& y, const int& q_minus_zero, const Y& scale) const
    {
        const float v = float(q_minus_zero) * detail::to_f32(scale);
        if constexpr(std::is_same_v<Y, bhalf_t>) y = bf16_convert_rtn<bhalf_t>(v);
        else y = static_cast<Y>(v);
    }
};

} // namespace element_wise
} // namespace tensor_operation
} // namespace ck
