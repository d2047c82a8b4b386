// epilogue.cuh -- device side of CK's CDE element-wise functors, applied to accumulator rows read
// back from TMEM.  Arithmetic follows the functor specialisations the oracle restates:
//   Add / Multiply / AddFastGelu / MultiplyFastGelu   binary_element_wise_operation.hpp:13-93,117-207,461-589
//   MultiplyAdd / MultiplyMultiply / MultiplyAddFastGelu   element_wise_operation.hpp:200-296
//   FastGelu                                            unary_element_wise_operation.hpp:435-548
//   Activation_Mul_Clamp family                         quantization_operation.hpp:25-201
// bf16 results are TRUNCATED (type_convert.hpp:53-64) unless CKB_FLAG_BF16_RNE is set.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdint>

#include "../../include/ckb.h"

namespace ckb {

struct EpiParams
{
    int32_t cde_op, act;
    float scale;
    uint32_t flags;
    int32_t num_d;
    int32_t d_dtype[CKB_MAX_D];
    const void* d[CKB_MAX_D];
    long long d_sm[CKB_MAX_D], d_sn[CKB_MAX_D];
    int32_t e_dtype;
    void* e;
    long long lde;
    int32_t M, N;
};

__device__ __forceinline__ float bf16_bits_to_f32(uint16_t b) { return __uint_as_float(uint32_t(b) << 16); }

__device__ __forceinline__ uint16_t f32_to_bf16_bits(float v, uint32_t flags)
{
    uint32_t u = __float_as_uint(v);
    if(flags & CKB_FLAG_BF16_RNE)
    {
        if((~u & 0x7f800000u) != 0) u += 0x7fffu + ((u >> 16) & 1u);
        else if(u & 0xffffu) u |= 0x10000u;
    }
    return uint16_t(u >> 16);
}

// host FastGelu formula (exp + true divide) so that device and oracle agree to ~1 ulp of fp32
__device__ __forceinline__ float fast_gelu_f32(float x)
{
    const float c1 = -2.0f * 0.035677f;
    const float c2 = -2.0f * 0.797885f;
    const float u  = x * (c1 * x * x + c2);
    return x / (1.f + expf(u));
}

__device__ __forceinline__ float load_d_f32(int dtype, const void* p, long long off)
{
    switch(dtype)
    {
    case CKB_F32: return static_cast<const float*>(p)[off];
    case CKB_F16: return __half2float(static_cast<const __half*>(p)[off]);
    case CKB_BF16: return bf16_bits_to_f32(static_cast<const uint16_t*>(p)[off]);
    case CKB_I32: return float(static_cast<const int32_t*>(p)[off]);
    case CKB_I8: return float(static_cast<const int8_t*>(p)[off]);
    default: return 0.f;
    }
}

// float-accumulator CDE functors: c is the fp32 accumulator (CShuffleDataType = F32)
__device__ __forceinline__ float cde_apply_f32(const EpiParams& ep, float c, float d0, float d1)
{
    switch(ep.cde_op)
    {
    case CKB_CDE_ADD:
        if(ep.e_dtype == CKB_F16 && ep.d_dtype[0] == CKB_F16) // Add<half_t,float,half_t>: half(c) + d in fp16
            return __half2float(__hadd(__float2half_rn(c), __float2half_rn(d0)));
        return c + d0;
    case CKB_CDE_FASTGELU: return fast_gelu_f32(c);
    case CKB_CDE_ADD_FASTGELU: return fast_gelu_f32(c + d0);
    case CKB_CDE_MULTIPLY:
        if(ep.e_dtype == CKB_F16 && ep.d_dtype[0] == CKB_F16)
            return __half2float(__hmul(__float2half_rn(c), __float2half_rn(d0)));
        return c * d0;
    case CKB_CDE_MULTIPLY_ADD: return __fadd_rn(__fmul_rn(c, d0), d1);
    case CKB_CDE_MULTIPLY_FASTGELU: return fast_gelu_f32(c * d0);
    case CKB_CDE_MULTIPLY_ADD_FASTGELU: return fast_gelu_f32(__fadd_rn(__fmul_rn(c, d0), d1));
    case CKB_CDE_MULTIPLY_MULTIPLY: return __fmul_rn(__fmul_rn(c, d0), d1);
    default: return c;
    }
}

// int32-accumulator functors -> float result to be truncated (static_cast) into int8/int32, or a
// float-like E for MultiplyMultiply / PassThrough-to-float.
__device__ __forceinline__ float
cde_apply_i32(const EpiParams& ep, int32_t acc, const void* const* dptr, const long long* doff)
{
    const float act_in_scale = ep.scale;
    auto act = [&](float v) { return ep.act == CKB_ACT_RELU ? (v > 0.f ? v : 0.f) : v; };
    switch(ep.cde_op)
    {
    case CKB_CDE_ACT_MUL_CLAMP:
        return fminf(fmaxf(__fmul_rn(act_in_scale, act(float(acc))), -128.f), 127.f);
    case CKB_CDE_ADD_ACT_MUL_CLAMP: {
        const int32_t b = static_cast<const int32_t*>(dptr[0])[doff[0]];
        return fminf(fmaxf(__fmul_rn(act_in_scale, act(float(acc + b))), -128.f), 127.f);
    }
    case CKB_CDE_ACT_MUL2_CLAMP: {
        const float s = load_d_f32(ep.d_dtype[0], dptr[0], doff[0]);
        return fminf(fmaxf(__fmul_rn(s, act(float(acc))), -128.f), 127.f);
    }
    case CKB_CDE_ADD_ACT_MUL2_CLAMP: {
        const int32_t b = static_cast<const int32_t*>(dptr[0])[doff[0]];
        const float s   = load_d_f32(ep.d_dtype[1], dptr[1], doff[1]);
        return fminf(fmaxf(__fmul_rn(s, act(float(acc + b))), -128.f), 127.f);
    }
    case CKB_CDE_MULTIPLY_MULTIPLY: {
        const float d0 = load_d_f32(ep.d_dtype[0], dptr[0], doff[0]);
        const float d1 = load_d_f32(ep.d_dtype[1], dptr[1], doff[1]);
        return __fmul_rn(__fmul_rn(float(acc), d0), d1);
    }
    default: return float(acc);
    }
}

// Convert NV consecutive float results (columns n0 .. n0+NV) of row m and store them.
template <int NV>
__device__ __forceinline__ void
store_row_f32(const EpiParams& ep, int m, int n0, const float* v)
{
    if(m >= ep.M) return;
    const long long base = (long long)m * ep.lde + n0;
    const bool full      = (n0 + NV <= ep.N);
    if(ep.e_dtype == CKB_BF16 || ep.e_dtype == CKB_F16)
    {
        uint16_t* e = static_cast<uint16_t*>(ep.e) + base;
        uint16_t h[NV];
#pragma unroll
        for(int i = 0; i < NV; ++i)
            h[i] = ep.e_dtype == CKB_BF16 ? f32_to_bf16_bits(v[i], ep.flags)
                                          : __half_as_ushort(__float2half_rn(v[i]));
        if(full && ((reinterpret_cast<uintptr_t>(e) & 15) == 0))
        {
#pragma unroll
            for(int i = 0; i < NV; i += 8)
            {
                uint4 q;
                q.x = h[i] | (uint32_t(h[i + 1]) << 16);
                q.y = h[i + 2] | (uint32_t(h[i + 3]) << 16);
                q.z = h[i + 4] | (uint32_t(h[i + 5]) << 16);
                q.w = h[i + 6] | (uint32_t(h[i + 7]) << 16);
                *reinterpret_cast<uint4*>(e + i) = q;
            }
        }
        else
        {
#pragma unroll
            for(int i = 0; i < NV; ++i)
                if(n0 + i < ep.N) e[i] = h[i];
        }
    }
    else if(ep.e_dtype == CKB_F32)
    {
        float* e = static_cast<float*>(ep.e) + base;
        if(full && ((reinterpret_cast<uintptr_t>(e) & 15) == 0))
        {
#pragma unroll
            for(int i = 0; i < NV; i += 4)
                *reinterpret_cast<float4*>(e + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        }
        else
        {
#pragma unroll
            for(int i = 0; i < NV; ++i)
                if(n0 + i < ep.N) e[i] = v[i];
        }
    }
    else if(ep.e_dtype == CKB_I8)
    {
        int8_t* e = static_cast<int8_t*>(ep.e) + base;
#pragma unroll
        for(int i = 0; i < NV; ++i)
            if(n0 + i < ep.N) e[i] = static_cast<int8_t>(v[i]); // static_cast: toward zero
    }
    else if(ep.e_dtype == CKB_I32)
    {
        int32_t* e = static_cast<int32_t*>(ep.e) + base;
#pragma unroll
        for(int i = 0; i < NV; ++i)
            if(n0 + i < ep.N) e[i] = static_cast<int32_t>(v[i]);
    }
}

// D(m, n0..n0+31) as fp32.  Fast paths: per-column vector (d_sm == 0, d_sn == 1: bias / column scale, 16-byte loads),
// per-row scalar (d_sn == 0: row scale); anything else element by element.
__device__ __forceinline__ void load_d_chunk32(const EpiParams& ep, int i, int m, int n0, float* d)
{
    const int dt = ep.d_dtype[i];
    const long long sm = ep.d_sm[i], sn = ep.d_sn[i];
    const char* base = static_cast<const char*>(ep.d[i]);
    if(sn == 0)
    {
        const float v = load_d_f32(dt, base, (long long)m * sm);
#pragma unroll
        for(int j = 0; j < 32; ++j) d[j] = v;
        return;
    }
    const bool full = (n0 + 32 <= ep.N);
    if(sn == 1 && full && (dt == CKB_BF16 || dt == CKB_F16))
    {
        const uint16_t* p = reinterpret_cast<const uint16_t*>(base) + (long long)m * sm + n0;
        if((reinterpret_cast<uintptr_t>(p) & 15) == 0)
        {
#pragma unroll
            for(int j = 0; j < 4; ++j)
            {
                const uint4 q = __ldg(reinterpret_cast<const uint4*>(p) + j);
                const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for(int t = 0; t < 4; ++t)
                {
                    if(dt == CKB_BF16)
                    {
                        d[8 * j + 2 * t]     = __uint_as_float(w[t] << 16);
                        d[8 * j + 2 * t + 1] = __uint_as_float(w[t] & 0xffff0000u);
                    }
                    else
                    {
                        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[t]));
                        d[8 * j + 2 * t] = f.x; d[8 * j + 2 * t + 1] = f.y;
                    }
                }
            }
            return;
        }
    }
    if(sn == 1 && full && (dt == CKB_F32 || dt == CKB_I32))
    {
        const uint32_t* p = reinterpret_cast<const uint32_t*>(base) + (long long)m * sm + n0;
        if((reinterpret_cast<uintptr_t>(p) & 15) == 0)
        {
#pragma unroll
            for(int j = 0; j < 8; ++j)
            {
                const uint4 q = __ldg(reinterpret_cast<const uint4*>(p) + j);
                const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for(int t = 0; t < 4; ++t) d[4 * j + t] = dt == CKB_F32 ? __uint_as_float(w[t]) : float(int32_t(w[t]));
            }
            return;
        }
    }
#pragma unroll
    for(int j = 0; j < 32; ++j) d[j] = (n0 + j < ep.N) ? load_d_f32(dt, base, (long long)m * sm + (long long)(n0 + j) * sn) : 0.f;
}

// One accumulator row chunk (32 fp32 columns of row m, bit patterns in r) through the float CDE functors.  The functor is
// selected ONCE per chunk so the element loops are branch-free.
__device__ __forceinline__ void cde_chunk32_f32(const EpiParams& ep, int m, int n0, const uint32_t* r, float* v)
{
    float d0[32], d1[32];
    if(ep.num_d > 0) load_d_chunk32(ep, 0, m, n0, d0);
    if(ep.num_d > 1) load_d_chunk32(ep, 1, m, n0, d1);
    const bool half_arith = ep.e_dtype == CKB_F16 && ep.num_d > 0 && ep.d_dtype[0] == CKB_F16;
    switch(ep.cde_op)
    {
    case CKB_CDE_ADD:
#pragma unroll
        for(int i = 0; i < 32; ++i)
            v[i] = half_arith ? __half2float(__hadd(__float2half_rn(__uint_as_float(r[i])), __float2half_rn(d0[i])))
                              : __uint_as_float(r[i]) + d0[i];
        break;
    case CKB_CDE_FASTGELU:
#pragma unroll
        for(int i = 0; i < 32; ++i) v[i] = fast_gelu_f32(__uint_as_float(r[i]));
        break;
    case CKB_CDE_ADD_FASTGELU:
#pragma unroll
        for(int i = 0; i < 32; ++i) v[i] = fast_gelu_f32(__uint_as_float(r[i]) + d0[i]);
        break;
    case CKB_CDE_MULTIPLY:
#pragma unroll
        for(int i = 0; i < 32; ++i)
            v[i] = half_arith ? __half2float(__hmul(__float2half_rn(__uint_as_float(r[i])), __float2half_rn(d0[i])))
                              : __uint_as_float(r[i]) * d0[i];
        break;
    case CKB_CDE_MULTIPLY_ADD:
#pragma unroll
        for(int i = 0; i < 32; ++i) v[i] = __fadd_rn(__fmul_rn(__uint_as_float(r[i]), d0[i]), d1[i]);
        break;
    case CKB_CDE_MULTIPLY_FASTGELU:
#pragma unroll
        for(int i = 0; i < 32; ++i) v[i] = fast_gelu_f32(__uint_as_float(r[i]) * d0[i]);
        break;
    case CKB_CDE_MULTIPLY_ADD_FASTGELU:
#pragma unroll
        for(int i = 0; i < 32; ++i) v[i] = fast_gelu_f32(__fadd_rn(__fmul_rn(__uint_as_float(r[i]), d0[i]), d1[i]));
        break;
    case CKB_CDE_MULTIPLY_MULTIPLY:
#pragma unroll
        for(int i = 0; i < 32; ++i) v[i] = __fmul_rn(__fmul_rn(__uint_as_float(r[i]), d0[i]), d1[i]);
        break;
    default:
#pragma unroll
        for(int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
    }
}

} // namespace ckb
