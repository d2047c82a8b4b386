// dequant.cuh -- in-register weight unpack/dequant for the TMEM-A GEMM kernels (K1/K2).
// The role of CK's ThreadwiseTensorSliceTransfer_v3r1_dequant + FastNumericArrayConverter<uint8_t,half_t,N>
// (thread/threadwise_tensor_slice_transfer_v3r1_dequant.hpp:429-549,
//  element/unary_element_wise_operation.hpp:1056-1122) and of the Multiply B-functor inside
// ThreadwiseTensorSliceTransfer_v7r2 (binary_element_wise_operation.hpp:184-190), done with
// prmt / lop3 magic numbers.  Every routine produces 16-bit pairs packed (k even -> low half) in
// the order the tcgen05 TMEM A operand wants: one 32-bit column = two consecutive k.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdint>

namespace ckb {

enum WFmt : int {
    WF_W4_GROUP   = 0, // pre-packed 4-bit codes, (q - zp) exact in A's type, group scale applied to fp32 partial sums
    WF_W8_MUL     = 1, // int8 (signed) * bf16 scale[n] -> bf16 by truncation (Multiply<bhalf_t,int8_t,bhalf_t>)
    WF_W8_FPAINTB = 2, // biased uint8 -> fp16(u8 - 128) * fp16 scale[n], one fp16 rounding (fpAintB)
    WF_W8_PASS    = 3  // int8 (signed) -> A's type exactly (PassThrough)
};

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}
template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
    return r;
}
__device__ __forceinline__ uint32_t hsub2_f16(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ uint32_t hmul2_f16(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ uint32_t hsub2_bf16(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

// ---- W4: one 32-bit word = 8 codes stored in the interleaved order [k0,k2,k4,k6,k1,k3,k5,k7]
// (nibble 0 = bits 0..3).  (w >> 4i) & 0x000F000F puts (k_2i, k_2i+1) in the two halves.
//   bf16: 0x4300 | q = 128 + q exactly;  fp16: 0x6400 | q = 1024 + q exactly
// `zmagic` = packed (magic + zero_point) in A's type; the subtraction is exact.
template <bool F16>
__device__ __forceinline__ void dequant_w4_word(uint32_t w, uint32_t zmagic, uint32_t* out4)
{
    constexpr uint32_t MAGIC = F16 ? 0x64006400u : 0x43004300u;
#pragma unroll
    for(int i = 0; i < 4; ++i)
    {
        // (x & 0x000F000F) | MAGIC   -> lop3 LUT (a & b) | c = 0xEA
        const uint32_t t = lop3<0xEA>(w >> (4 * i), 0x000F000Fu, MAGIC);
        out4[i]          = F16 ? hsub2_f16(t, zmagic) : hsub2_bf16(t, zmagic);
    }
}

__device__ __forceinline__ uint32_t hmul2_bf16(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
// W4 with the (k-group, n) scale folded in before the MMA: round_to_nearest_A( (q - zp) * s ).
// (q - zp) is exact in A's type and the product of two 8/11-bit significands is exact in fp32, so the
// packed multiply performs exactly ONE rounding -- the [EXT] analogue of CK's Multiply B-functor /
// fpAintB converter, which also round q*s once into A's type before the MMA.
template <bool F16>
__device__ __forceinline__ void dequant_w4_word_scaled(uint32_t w, uint32_t zmagic, uint32_t scale2, uint32_t* out4)
{
    constexpr uint32_t MAGIC = F16 ? 0x64006400u : 0x43004300u;
#pragma unroll
    for(int i = 0; i < 4; ++i)
    {
        const uint32_t t = lop3<0xEA>(w >> (4 * i), 0x000F000Fu, MAGIC);
        out4[i] = F16 ? hmul2_f16(hsub2_f16(t, zmagic), scale2) : hmul2_bf16(hsub2_bf16(t, zmagic), scale2);
    }
}

// Fragment-major W4 word (ckb_prepack_b_int4): pairs 0,2 belong to row r_lo, pairs 1,3 to row r_hi = r_lo + 8, each row with
// its own (magic + zero point) and scale.  out4 = {(lo;k0,k1), (hi;k0,k1), (lo;k2,k3), (hi;k2,k3)} = m16n8k16 A fragment.
template <bool F16>
__device__ __forceinline__ void dequant_w4_word_scaled2(uint32_t w, uint32_t zm_lo, uint32_t zm_hi, uint32_t sc_lo,
                                                        uint32_t sc_hi, uint32_t* out4)
{
    constexpr uint32_t MAGIC = F16 ? 0x64006400u : 0x43004300u;
#pragma unroll
    for(int i = 0; i < 4; ++i)
    {
        const uint32_t t  = lop3<0xEA>(w >> (4 * i), 0x000F000Fu, MAGIC);
        const uint32_t zm = (i & 1) ? zm_hi : zm_lo;
        const uint32_t sc = (i & 1) ? sc_hi : sc_lo;
        out4[i] = F16 ? hmul2_f16(hsub2_f16(t, zm), sc) : hmul2_bf16(hsub2_bf16(t, zm), sc);
    }
}

// Same word, exact part only: (q - zp) in A's type (no rounding at all); the caller applies the group scale to the
// fp32 partial sums (CK's ab_scale host pattern scales in float: gemm_multiply_multiply_xdl_fp8_ab_scale.cpp:258-290).
template <bool F16>
__device__ __forceinline__ void dequant_w4_word_exact2(uint32_t w, uint32_t zm_lo, uint32_t zm_hi, uint32_t* out4)
{
    constexpr uint32_t MAGIC = F16 ? 0x64006400u : 0x43004300u;
#pragma unroll
    for(int i = 0; i < 4; ++i)
    {
        const uint32_t t  = lop3<0xEA>(w >> (4 * i), 0x000F000Fu, MAGIC);
        const uint32_t zm = (i & 1) ? zm_hi : zm_lo;
        out4[i] = F16 ? hsub2_f16(t, zm) : hsub2_bf16(t, zm);
    }
}

// Same word with the zero point left in: magic + q (128 + q in bf16, 1024 + q in fp16), exact, 3 shifts + 4 lop3.  The caller
// removes (magic + zp) * sum_k(a_k) from the partial sums (the activation sums come from one extra HMMA per k16 step with an
// all-ones A fragment, shared by all channel groups of the warp).
template <bool F16>
__device__ __forceinline__ void dequant_w4_word_magic(uint32_t w, uint32_t* out4)
{
    constexpr uint32_t MAGIC = F16 ? 0x64006400u : 0x43004300u;
#pragma unroll
    for(int i = 0; i < 4; ++i) out4[i] = lop3<0xEA>(w >> (4 * i), 0x000F000Fu, MAGIC);
}

// ---- W8 fpAintB (fp16): u8 -> 0x6400|u8 = 1024 + u8; minus 1152 (= 0x6480) removes the +128 bias
// (the reference's own trick, unary_element_wise_operation.hpp:1061-1091); times scale: ONE rounding.
__device__ __forceinline__ void dequant_u8x4_fpaintb(uint32_t w, uint32_t scale2, uint32_t* out2)
{
    const uint32_t lo = prmt(w, 0x64646464u, 0x4140); // bytes: [b0, 0x64, b1, 0x64]
    const uint32_t hi = prmt(w, 0x64646464u, 0x4342); // bytes: [b2, 0x64, b3, 0x64]
    out2[0]           = hmul2_f16(hsub2_f16(lo, 0x64806480u), scale2);
    out2[1]           = hmul2_f16(hsub2_f16(hi, 0x64806480u), scale2);
}

// ---- W8 signed int8 -> exact value in fp32: flip the sign bit (int8 + 128 as u8), place the byte in
// the mantissa of 2^23 (0x4B000000), subtract 2^23 + 128.
__device__ __forceinline__ float s8_byte_to_f32(uint32_t w_biased, int byte)
{
    const uint32_t sel = 0x7650u | uint32_t(byte); // [byte, 0x00, 0x00, 0x4B] of (w, 0x4B000000)
    return __uint_as_float(prmt(w_biased, 0x4B000000u, sel)) - 8388736.0f;
}

// q * s in ONE FFMA per code, exactly: the biased byte b = q + 128 is placed in mantissa bits 8..15 of 2^15 (0x4700bb00 = 2^15 + b,
// one PRMT), and fma(2^15 + b, s, -(2^15 + 128) s) = (b - 128) s = q s.  Exact because the FFMA rounds the exact sum once and q s has
// <= 16 significant bits (s is a bf16 value); the addend -(2^15 + 128) s = -2^7 * 257 * s has <= 17 bits, so it is exact as well.
// (3 instead of 4 instructions per code with the final pack; these kernels are issue-bound on the conversion.)  The only
// difference from (float(q)) * s is the SIGN of a zero weight under a negative scale (+0 instead of -0), which cannot reach a GEMM
// output: the accumulators start at +0 and (+0) + (-0) = +0.  Domain: |s| < 2^112 (beyond it the addend overflows; checked exhaustively
// over all finite bf16 scales x all 256 codes, tests/test_host_logic.py::test_int8_scale_fma_identity).
__device__ __forceinline__ float s8_biased_byte_times_scale(uint32_t w_biased, int byte, float scale, float neg_bias_scale)
{
    const uint32_t sel = 0x7404u | (uint32_t(byte) << 4); // [0x47, 0x00, byte, 0x00] of (w, 0x47000000)
    return fmaf(__uint_as_float(prmt(w_biased, 0x47000000u, sel)), scale, neg_bias_scale);
}

// Multiply<bhalf_t,int8_t,bhalf_t>: truncate(float(q) * float(s)) -- q*s is exact in fp32 (8 x 8 bits)
__device__ __forceinline__ void dequant_s8x4_mul_bf16(uint32_t w, float scale, uint32_t* out2)
{
    const uint32_t wb = w ^ 0x80808080u;
    const float nb = scale * -32896.0f;
    const float f0 = s8_biased_byte_times_scale(wb, 0, scale, nb), f1 = s8_biased_byte_times_scale(wb, 1, scale, nb);
    const float f2 = s8_biased_byte_times_scale(wb, 2, scale, nb), f3 = s8_biased_byte_times_scale(wb, 3, scale, nb);
    out2[0] = prmt(__float_as_uint(f0), __float_as_uint(f1), 0x7632); // high halves: truncation
    out2[1] = prmt(__float_as_uint(f2), __float_as_uint(f3), 0x7632);
}

// PassThrough<AType,int8_t>: exact (|q| <= 128 fits both 16-bit types)
template <bool F16>
__device__ __forceinline__ void dequant_s8x4_pass(uint32_t w, uint32_t* out2)
{
    const uint32_t wb = w ^ 0x80808080u;
    if(F16)
    {
        const uint32_t lo = prmt(wb, 0x64646464u, 0x4140);
        const uint32_t hi = prmt(wb, 0x64646464u, 0x4342);
        out2[0]           = hsub2_f16(lo, 0x64806480u);
        out2[1]           = hsub2_f16(hi, 0x64806480u);
    }
    else
    {
        const float f0 = s8_byte_to_f32(wb, 0), f1 = s8_byte_to_f32(wb, 1);
        const float f2 = s8_byte_to_f32(wb, 2), f3 = s8_byte_to_f32(wb, 3);
        out2[0] = prmt(__float_as_uint(f0), __float_as_uint(f1), 0x7632);
        out2[1] = prmt(__float_as_uint(f2), __float_as_uint(f3), 0x7632);
    }
}

} // namespace ckb
