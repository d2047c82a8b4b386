/*
 * ck_oracle.h -- C ABI of the CPU oracle (TEST INFRASTRUCTURE, not product code).
 *
 * A plain C++ restatement of Composable Kernel's host reference GEMMs:
 *   - ReferenceGemm            library/include/ck/library/reference_tensor_operation/cpu/reference_gemm.hpp:60-108
 *   - ReferencefpAintBGemm     .../cpu/reference_fpAintB_gemm.hpp:62-127
 *   - ReferenceGemmMultipleD   .../cpu/reference_gemm_multiple_d.hpp:66-125
 * and of the example-level compositions that pin the mixed-precision path
 * (example/60_gemm_multi_ABD/gemm_multi_ABD_xdl_bias_fastgelu_bf16_i8.cpp:229-266,
 *  example/14_gemm_quantization/gemm_xdl_bias_relu_quantization_int8.cpp:97-103,213-229,
 *  example/65_gemm_multiply_multiply/gemm_multiply_multiply_xdl_fp8.cpp:243-266,
 *  example/65_gemm_multiply_multiply/gemm_multiply_multiply_xdl_fp8_ab_scale.cpp:258-313).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (composable_kernel_b200/, include/) never does.
 *
 * PARITY PIN STATUS: the reference's own host code cannot be compiled in this image (clang/HIP-only
 * constructs: _BitInt, ext_vector_type, in-class explicit specialisations, hard <hip/hip_runtime.h>
 * include) and the reference stores NO golden vectors for this path.  The oracle is
 * pinned against (a) the known-answer constants in the reference's test/data_type/type_convert_const.cpp
 * and test/data_type/test_int4.cpp, (b) an independent float64/int64 numpy evaluation, and
 * (c) committed fixtures under tests/golden/ produced by oracle/make_golden.py.
 * GEMM-level parity is therefore "pinned on conversion KATs + independent re-derivation", not on
 * reference-produced outputs.
 */
#ifndef CK_ORACLE_H
#define CK_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum orc_dtype { ORC_F32 = 0, ORC_F16 = 1, ORC_BF16 = 2, ORC_I8 = 3, ORC_U8 = 4, ORC_I32 = 5 };

/* how B(k,n) becomes a GEMM operand */
enum orc_b_op {
    ORC_BOP_PASSTHROUGH = 0, /* b = type_convert<AType>(B0)            unary_element_wise_operation.hpp:112-121 */
    ORC_BOP_MULTIPLY    = 1, /* Multiply<bhalf_t,int8_t,bhalf_t>        binary_element_wise_operation.hpp:184-190 (truncating bf16) */
    ORC_BOP_FPAINTB     = 2, /* type_convert<A>(b) * scale in A's type  reference_fpAintB_gemm.hpp:110 (fp16, one rounding) */
    ORC_BOP_GROUP_F32   = 3, /* [EXT] float(q - zp[k/g,n]) * float(s[k/g,n]); pattern: gemm_multiply_multiply_xdl_fp8_ab_scale.cpp:258-290 */
    ORC_BOP_GROUP_RTN   = 4  /* [EXT] the same product rounded ONCE (nearest-even) into A's type before the GEMM, as CK's
                                dequant B-functors do (reference_fpAintB_gemm.hpp:110; bf16 through bf16_convert_rtn,
                                type_convert.hpp:508-549).  This is the definition the W4A16 kernels implement. */
};

/* CDE (epilogue) functor */
enum orc_cde_op {
    ORC_CDE_PASSTHROUGH           = 0,
    ORC_CDE_ADD                   = 1,  /* binary_element_wise_operation.hpp:13-93   */
    ORC_CDE_FASTGELU              = 2,  /* unary_element_wise_operation.hpp:435-548  */
    ORC_CDE_ADD_FASTGELU          = 3,  /* binary_element_wise_operation.hpp:461-524 */
    ORC_CDE_MULTIPLY              = 4,  /* binary_element_wise_operation.hpp:117-207 */
    ORC_CDE_MULTIPLY_ADD          = 5,  /* element_wise_operation.hpp:200-250 (bf16 variant restated as intended) */
    ORC_CDE_MULTIPLY_FASTGELU     = 6,  /* binary_element_wise_operation.hpp:526-589 */
    ORC_CDE_MULTIPLY_ADD_FASTGELU = 7,  /* element_wise_operation.hpp:277-296 */
    ORC_CDE_MULTIPLY_MULTIPLY     = 8,  /* element_wise_operation.hpp:252-275 */
    ORC_CDE_ACT_MUL_CLAMP         = 9,  /* quantization_operation.hpp:25-66   */
    ORC_CDE_ADD_ACT_MUL_CLAMP     = 10, /* quantization_operation.hpp:132-171 */
    ORC_CDE_ACT_MUL2_CLAMP        = 11, /* quantization_operation.hpp:102-127 */
    ORC_CDE_ADD_ACT_MUL2_CLAMP    = 12  /* quantization_operation.hpp:176-201 */
};

enum orc_act { ORC_ACT_PASSTHROUGH = 0, ORC_ACT_RELU = 1 };

typedef struct orc_problem {
    int32_t M, N, K;

    /* A(m,k) = a[m*a_sm + k*a_sk] */
    int32_t a_dtype; /* ORC_F16 | ORC_BF16 | ORC_I8 */
    const void* a;
    int64_t a_sm, a_sk;

    /* B(k,n) = b[k*b_sk + n*b_sn]; ORC_I8 (signed) or ORC_U8 (unpacked 4-bit codes 0..15 for GROUP_F32) */
    int32_t b_dtype;
    const void* b;
    int64_t b_sk, b_sn;
    int32_t b_op;

    /* scale: MULTIPLY/FPAINTB: scale(k,n) = s[k*s_sk + n*s_sn]; GROUP_F32: s[(k/group)*s_sk + n*s_sn] */
    int32_t s_dtype; /* ORC_F16 | ORC_BF16 | ORC_F32 */
    const void* scale;
    int64_t s_sk, s_sn;

    /* GROUP_F32 only: zero points u8 [(k/group)*z_sk + n*z_sn], or NULL -> zero_const */
    const void* zero;
    int64_t z_sk, z_sn;
    int32_t zero_const;
    int32_t group_size;

    /* accumulator: float (k ascending, sequential) or int32 when a_dtype == ORC_I8 */
    int32_t cshuffle_dtype; /* type C is rounded to before the CDE op: ORC_F32 | ORC_BF16 | ORC_F16 | ORC_I32 */

    int32_t num_d;
    int32_t d_dtype[2];
    const void* d[2];
    int64_t d_sm[2], d_sn[2];

    int32_t cde_op;
    float cde_scale; /* requantScale_ for *_MUL_CLAMP */
    int32_t act;     /* activation inside *_MUL*_CLAMP */

    int32_t e_dtype; /* ORC_F16 | ORC_BF16 | ORC_F32 | ORC_I8 | ORC_I32 */
    void* e;
    int64_t e_sm, e_sn;

    /* evaluate rows [m_begin, m_end) only (0,0 = all rows): row-block sampling of big shapes */
    int32_t m_begin, m_end;
    int32_t num_threads; /* 0 = std::thread::hardware_concurrency() (reference_gemm.hpp:103-105) */
} orc_problem;

/* returns 0, or a negative code for an unsupported dtype/op combination (message via orc_last_error) */
int orc_gemm(const orc_problem* p);
const char* orc_last_error(void);

/* conversions (include/ck/utility/type_convert.hpp:42-100, 508-549) exposed for KAT tests */
uint16_t orc_f32_to_bf16_trunc(float x);
uint16_t orc_f32_to_bf16_rtn(float x);
float orc_bf16_to_f32(uint16_t x);
uint16_t orc_f32_to_f16(float x);
float orc_f16_to_f32(uint16_t x);
uint16_t orc_i8_to_bf16(int8_t x);
int8_t orc_bf16_to_i8(uint16_t x);
float orc_fast_gelu(float x);

/* std::mt19937(11939)-seeded fills of library/include/ck/library/utility/fill.hpp:19-95 */
void orc_fill_uniform_f32(float* dst, int64_t n, float a, float b, int round_to_int);

/* element-wise helper used by tests: UnsignedWeightPreprocessor (example/64_fpAintB_gemm/common.hpp:56-83) */
void orc_int8_to_biased_u8(const int8_t* src, uint8_t* dst, int64_t n);

int orc_hardware_threads(void);

#ifdef __cplusplus
}
#endif
#endif
