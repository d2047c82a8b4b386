// ck_oracle.cpp -- CPU oracle (TEST INFRASTRUCTURE ONLY; see ck_oracle.h for scope and pin status).
//
// Restates, in plain C++17 for g++, the arithmetic of Composable Kernel's host reference GEMMs.
// Every function names the reference file:line it follows.  Build flags matter for parity:
// -ffp-contract=off (the reference host build has no FMA contraction on baseline x86-64) and
// no -ffast-math.  See oracle/Makefile.
#include "ck_oracle.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

namespace {

thread_local std::string g_err;

using half_t  = _Float16; // include/ck/utility/data_type.hpp:11  (half_t = _Float16)
using bhalf_t = uint16_t; // include/ck/utility/data_type.hpp:10  (bhalf_t = ushort)

// ---- type_convert (include/ck/utility/type_convert.hpp) -------------------------------------

// NOTE: g++ 13.3 -O1+ miscompiles `cond ? bf16_to_f32(f32_to_bf16_trunc(v)) : v` into `v` when both
// helpers are inlined (phi-merge of a bit-mask round trip); the helpers are therefore noinline.  The
// numpy re-derivation in tests/test_oracle.py guards this.
#define ORC_NOINLINE __attribute__((noinline))

// :42-52  bf16 -> fp32: bits << 16
ORC_NOINLINE float bf16_to_f32(bhalf_t x)
{
    uint32_t u = uint32_t(x) << 16;
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
// :53-64  fp32 -> bf16: TRUNCATION (bits >> 16), not round-to-nearest
ORC_NOINLINE bhalf_t f32_to_bf16_trunc(float x)
{
    uint32_t u;
    std::memcpy(&u, &x, 4);
    return uint16_t(u >> 16);
}
// :508-549  bf16_convert_rtn<bhalf_t,float>: round-to-nearest-even, NaN preserved
ORC_NOINLINE bhalf_t f32_to_bf16_rtn(float x)
{
    uint32_t u;
    std::memcpy(&u, &x, 4);
    const bool flag0 = (~u & 0x7f800000u) != 0;
    const bool flag1 = !flag0 && (u & 0xffffu);
    u += flag0 ? 0x7fffu + ((u >> 16) & 1u) : 0u;
    u |= flag1 ? 0x10000u : 0u;
    return uint16_t(u >> 16);
}
ORC_NOINLINE float f16_to_f32(half_t x) { return static_cast<float>(x); }
ORC_NOINLINE half_t f32_to_f16(float x) { return static_cast<half_t>(x); } // :17-25 static_cast (RNE)
// :93-100  int8 -> bf16 via fp32
inline bhalf_t i8_to_bf16(int8_t x) { return f32_to_bf16_trunc(static_cast<float>(x)); }
// :84-91  bf16 -> int8 via fp32
inline int8_t bf16_to_i8(bhalf_t x) { return static_cast<int8_t>(bf16_to_f32(x)); }

// ---- FastGelu host specialisation (unary_element_wise_operation.hpp:443-452) ----------------
inline float fast_gelu(float x)
{
    const float c1  = -2.0 * 0.035677f;
    const float c2  = -2.0 * 0.797885f;
    const float u   = x * (c1 * x * x + c2);
    const float emu = std::exp(u);
    return x / (1.f + emu);
}

inline float clampf(float v, float lo, float hi) { return std::min(std::max(v, lo), hi); }

inline float load_as_f32(int dtype, const void* p, int64_t off)
{
    switch(dtype)
    {
    case ORC_F32: return static_cast<const float*>(p)[off];
    case ORC_F16: return f16_to_f32(static_cast<const half_t*>(p)[off]);
    case ORC_BF16: return bf16_to_f32(static_cast<const bhalf_t*>(p)[off]);
    case ORC_I8: return float(static_cast<const int8_t*>(p)[off]);
    case ORC_U8: return float(static_cast<const uint8_t*>(p)[off]);
    case ORC_I32: return float(static_cast<const int32_t*>(p)[off]);
    }
    return 0.f;
}

// PassThrough<Y,float> into the output / CShuffle type: bf16 = truncation (type_convert.hpp:53-64),
// fp16 = static_cast.
inline float round_through(int dtype, float v)
{
    switch(dtype)
    {
    case ORC_BF16: return bf16_to_f32(f32_to_bf16_trunc(v));
    case ORC_F16: return f16_to_f32(f32_to_f16(v));
    default: return v;
    }
}

inline void store_float(int dtype, void* p, int64_t off, float v)
{
    switch(dtype)
    {
    case ORC_F32: static_cast<float*>(p)[off] = v; break;
    case ORC_F16: static_cast<half_t*>(p)[off] = f32_to_f16(v); break;
    case ORC_BF16: static_cast<bhalf_t*>(p)[off] = f32_to_bf16_trunc(v); break;
    default: break;
    }
}

struct Ctx
{
    const orc_problem* p;
    std::vector<float> bf;   // dequantised B as float, [n][k] (k contiguous)   -- float path
    std::vector<float> af;   // A rows as float [m - m_begin][k]
    std::vector<int8_t> bi;  // int8 B [n][k]                                   -- int path
};

// B operand value, restating the per-element B functors.
int build_b(Ctx& c)
{
    const orc_problem& p = *c.p;
    const int K = p.K, N = p.N;
    if(p.a_dtype == ORC_I8)
    {
        if(p.b_dtype != ORC_I8 || p.b_op != ORC_BOP_PASSTHROUGH)
        {
            g_err = "int8 A requires int8 B with PassThrough";
            return -1;
        }
        c.bi.resize(size_t(N) * K);
        const int8_t* b = static_cast<const int8_t*>(p.b);
        for(int n = 0; n < N; ++n)
            for(int k = 0; k < K; ++k)
                c.bi[size_t(n) * K + k] = b[k * p.b_sk + n * p.b_sn];
        return 0;
    }
    c.bf.resize(size_t(N) * K);
    std::atomic<int> rc_all{0};
    auto do_cols = [&](int n_begin, int n_end) -> void {
    for(int n = n_begin; n < n_end; ++n)
    {
        for(int k = 0; k < K; ++k)
        {
            float v = 0.f;
            const int64_t boff = k * p.b_sk + n * p.b_sn;
            switch(p.b_op)
            {
            case ORC_BOP_PASSTHROUGH: {
                // PassThrough<half_t,int8_t> / <bhalf_t,int8_t> (unary_element_wise_operation.hpp:112-121):
                // exact for |q| <= 128 in both 16-bit types.
                if(p.b_dtype == ORC_I8)
                    v = float(static_cast<const int8_t*>(p.b)[boff]);
                else if(p.b_dtype == ORC_U8)
                    v = float(static_cast<const uint8_t*>(p.b)[boff]);
                else
                    v = load_as_f32(p.b_dtype, p.b, boff);
                v = round_through(p.a_dtype, v);
                break;
            }
            case ORC_BOP_MULTIPLY: {
                // Multiply<bhalf_t,int8_t,bhalf_t> (binary_element_wise_operation.hpp:184-190):
                //   y = type_convert<bhalf_t>(float(x0) * float(x1))   -> truncation
                if(p.b_dtype != ORC_I8 || p.s_dtype != ORC_BF16 || p.a_dtype != ORC_BF16)
                {
                    g_err = "Multiply B-op is defined for int8 B0, bf16 B1, bf16 A only";
                    rc_all = -1; return;
                }
                const float x0 = float(static_cast<const int8_t*>(p.b)[boff]);
                const float x1 =
                    bf16_to_f32(static_cast<const bhalf_t*>(p.scale)[k * p.s_sk + n * p.s_sn]);
                v = bf16_to_f32(f32_to_bf16_trunc(x0 * x1));
                break;
            }
            case ORC_BOP_FPAINTB: {
                // reference_fpAintB_gemm.hpp:110   v_converted_b = type_convert<ADataType>(v_b) * v_scale
                // evaluated in ADataType (= fp16): one rounding of q*s.
                if(p.b_dtype != ORC_I8 || p.s_dtype != ORC_F16 || p.a_dtype != ORC_F16)
                {
                    g_err = "fpAintB is defined for fp16 A, int8 B, fp16 scale only";
                    rc_all = -1; return;
                }
                const half_t q = static_cast<half_t>(static_cast<const int8_t*>(p.b)[boff]);
                const half_t s = static_cast<const half_t*>(p.scale)[k * p.s_sk + n * p.s_sn];
                const half_t r = static_cast<half_t>(q * s);
                v              = f16_to_f32(r);
                break;
            }
            case ORC_BOP_GROUP_F32:
            case ORC_BOP_GROUP_RTN: {
                // [EXT] K-blocked scales; host pattern of gemm_multiply_multiply_xdl_fp8_ab_scale.cpp:258-290:
                // pre-scale B into a FLOAT tensor with b1(k / ScaleBlockK, n), then ReferenceGemm<float,...>.
                if(p.group_size <= 0)
                {
                    g_err = "group_size must be > 0";
                    rc_all = -1; return;
                }
                const int g = k / p.group_size;
                int q       = (p.b_dtype == ORC_U8) ? int(static_cast<const uint8_t*>(p.b)[boff])
                                                    : int(static_cast<const int8_t*>(p.b)[boff]);
                const int z = p.zero ? int(static_cast<const uint8_t*>(p.zero)[g * p.z_sk + n * p.z_sn])
                                     : p.zero_const;
                const float s = load_as_f32(p.s_dtype, p.scale, g * p.s_sk + n * p.s_sn);
                v             = float(q - z) * s;
                if(p.b_op == ORC_BOP_GROUP_RTN) // one rounding into A's type (RNE)
                    v = p.a_dtype == ORC_BF16 ? bf16_to_f32(f32_to_bf16_rtn(v)) : f16_to_f32(f32_to_f16(v));
                break;
            }
            default: rc_all = -1; return;
            }
            c.bf[size_t(n) * K + k] = v;
        }
    }
    };
    int nt = p.num_threads > 0 ? p.num_threads : int(std::thread::hardware_concurrency());
    nt     = std::max(1, std::min(nt, N / 64 + 1));
    const int per = (N + nt - 1) / nt;
    std::vector<std::thread> th;
    for(int t = 0; t < nt; ++t)
    {
        const int b = t * per, e = std::min(N, b + per);
        if(b < e) th.emplace_back(do_cols, b, e);
    }
    for(auto& t : th) t.join();
    if(rc_all != 0)
    {
        g_err = "unsupported dtype / b_op combination for the B operand (see ck_oracle.h)";
        return -1;
    }
    return 0;
}


// CDE functors on a float C (already rounded through the CShuffle type).
inline bool cde_float(const orc_problem& p, float c, float d0, float d1, float& e)
{
    const int E = p.e_dtype;
    switch(p.cde_op)
    {
    case ORC_CDE_PASSTHROUGH: e = c; return true;
    case ORC_CDE_ADD:
        // Add<bhalf_t,float,bhalf_t> :80-85 -> trunc(c + float(d)); Add<half_t,float,half_t> :47-52 -> half(c) + d in fp16
        if(E == ORC_F16 && p.d_dtype[0] == ORC_F16)
            e = f16_to_f32(static_cast<half_t>(f32_to_f16(c) + f32_to_f16(d0)));
        else
            e = c + d0;
        return true;
    case ORC_CDE_FASTGELU: e = fast_gelu(c); return true;
    case ORC_CDE_ADD_FASTGELU: e = fast_gelu(c + d0); return true; // :501-524
    case ORC_CDE_MULTIPLY:
        // Multiply<bhalf_t,float,bhalf_t> :193-199; Multiply<half_t,float,half_t> :153-158 -> half(c) * d in fp16
        if(E == ORC_F16 && p.d_dtype[0] == ORC_F16)
            e = f16_to_f32(static_cast<half_t>(f32_to_f16(c) * f32_to_f16(d0)));
        else
            e = c * d0;
        return true;
    case ORC_CDE_MULTIPLY_ADD:
        // element_wise_operation.hpp:224-231 multiplies raw ushort bit patterns for bf16 (a
        // reference bug); restated as the intended float arithmetic c*d0 + d1 (as :232-249 do for fp16/float).
        e = c * d0 + d1;
        return true;
    case ORC_CDE_MULTIPLY_FASTGELU: e = fast_gelu(c * d0); return true;          // :566-589
    case ORC_CDE_MULTIPLY_ADD_FASTGELU: e = fast_gelu(c * d0 + d1); return true; // :283-295
    case ORC_CDE_MULTIPLY_MULTIPLY: e = c * d0 * d1; return true;                // :258-274
    default: return false;
    }
}

void run_rows_float(const Ctx& c, int m0, int m1)
{
    const orc_problem& p = *c.p;
    const int K = p.K, N = p.N;
    std::vector<float> arow(K);
    for(int m = m0; m < m1; ++m)
    {
        for(int k = 0; k < K; ++k)
            arow[k] = load_as_f32(p.a_dtype, p.a, m * p.a_sm + k * p.a_sk);
        for(int n = 0; n < N; ++n)
        {
            const float* b = &c.bf[size_t(n) * K];
            // reference_gemm.hpp:64-95: AccDataType v_acc = 0; for k ascending: v_acc += float(a)*float(b)
            float acc = 0.f;
            for(int k = 0; k < K; ++k)
                acc += arow[k] * b[k];
            // c_element_op into CShuffleDataType (PassThrough<C,float>), then the CDE functor
            const float cs = round_through(p.cshuffle_dtype, acc);
            float d0 = 0.f, d1 = 0.f;
            if(p.num_d > 0) d0 = load_as_f32(p.d_dtype[0], p.d[0], m * p.d_sm[0] + n * p.d_sn[0]);
            if(p.num_d > 1) d1 = load_as_f32(p.d_dtype[1], p.d[1], m * p.d_sm[1] + n * p.d_sn[1]);
            float e = 0.f;
            cde_float(p, cs, d0, d1, e);
            store_float(p.e_dtype, p.e, m * p.e_sm + n * p.e_sn, e);
        }
    }
}

void run_rows_int(const Ctx& c, int m0, int m1)
{
    const orc_problem& p = *c.p;
    const int K = p.K, N = p.N;
    const int8_t* a = static_cast<const int8_t*>(p.a);
    std::vector<int32_t> arow(K);
    for(int m = m0; m < m1; ++m)
    {
        for(int k = 0; k < K; ++k)
            arow[k] = a[m * p.a_sm + k * p.a_sk];
        for(int n = 0; n < N; ++n)
        {
            const int8_t* b = &c.bi[size_t(n) * K];
            // ReferenceGemm<int8,int8,int32,int32> (gemm_xdl_bias_relu_quantization_int8.cpp:97-103): exact
            int32_t acc = 0;
            for(int k = 0; k < K; ++k)
                acc += arow[k] * int32_t(b[k]);
            const int64_t eoff = m * p.e_sm + n * p.e_sn;
            auto d_i32 = [&](int i) {
                return static_cast<const int32_t*>(p.d[i])[m * p.d_sm[i] + n * p.d_sn[i]];
            };
            auto d_f32 = [&](int i) { return load_as_f32(p.d_dtype[i], p.d[i], m * p.d_sm[i] + n * p.d_sn[i]); };
            auto act = [&](float v) { return p.act == ORC_ACT_RELU ? (v > 0 ? v : 0.f) : v; };
            auto put_q = [&](float y) {
                // quantization_operation.hpp:39-45: clamp(scale * act(float(x)), -128, 127) then static_cast
                if(p.e_dtype == ORC_I8)
                    static_cast<int8_t*>(p.e)[eoff] = static_cast<int8_t>(y);
                else
                    static_cast<int32_t*>(p.e)[eoff] = static_cast<int32_t>(y);
            };
            switch(p.cde_op)
            {
            case ORC_CDE_PASSTHROUGH:
                if(p.e_dtype == ORC_I32)
                    static_cast<int32_t*>(p.e)[eoff] = acc;
                else if(p.e_dtype == ORC_I8) // PassThrough<int8_t,int32_t> :130-133
                    static_cast<int8_t*>(p.e)[eoff] = static_cast<int8_t>(acc);
                else
                    store_float(p.e_dtype, p.e, eoff, float(acc));
                break;
            case ORC_CDE_ACT_MUL_CLAMP:
                put_q(clampf(p.cde_scale * act(float(acc)), -128.f, 127.f));
                break;
            case ORC_CDE_ADD_ACT_MUL_CLAMP: // :153-160  float(x + bias)
                put_q(clampf(p.cde_scale * act(float(acc + d_i32(0))), -128.f, 127.f));
                break;
            case ORC_CDE_ACT_MUL2_CLAMP: // :107-114
                put_q(clampf(d_f32(0) * act(float(acc)), -128.f, 127.f));
                break;
            case ORC_CDE_ADD_ACT_MUL2_CLAMP: // :181-188
                put_q(clampf(d_f32(1) * act(float(acc + d_i32(0))), -128.f, 127.f));
                break;
            case ORC_CDE_MULTIPLY_MULTIPLY: {
                // [EXT dtype] MultiplyMultiply<E,int32->float,float,float>: e = convert<E>(float(c)*d0*d1)
                const float x = float(acc) * d_f32(0) * d_f32(1);
                store_float(p.e_dtype, p.e, eoff, x);
                break;
            }
            default: break;
            }
        }
    }
}

} // namespace

extern "C" {

const char* orc_last_error(void) { return g_err.c_str(); }
int orc_hardware_threads(void) { return int(std::thread::hardware_concurrency()); }

int orc_gemm(const orc_problem* pp)
{
    if(!pp) { g_err = "null problem"; return -1; }
    const orc_problem& p = *pp;
    if(p.M < 0 || p.N < 0 || p.K < 0) { g_err = "negative size"; return -1; }
    const bool is_int = p.a_dtype == ORC_I8;
    if(is_int)
    {
        switch(p.cde_op)
        {
        case ORC_CDE_PASSTHROUGH: case ORC_CDE_ACT_MUL_CLAMP: case ORC_CDE_ADD_ACT_MUL_CLAMP:
        case ORC_CDE_ACT_MUL2_CLAMP: case ORC_CDE_ADD_ACT_MUL2_CLAMP: case ORC_CDE_MULTIPLY_MULTIPLY: break;
        default: g_err = "cde_op not defined for int32 accumulators"; return -2;
        }
    }
    else
    {
        if(p.a_dtype != ORC_F16 && p.a_dtype != ORC_BF16) { g_err = "A must be f16/bf16/i8"; return -2; }
        if(p.cde_op > ORC_CDE_MULTIPLY_MULTIPLY) { g_err = "quantization CDE ops need int8 inputs"; return -2; }
        if(p.e_dtype != ORC_F16 && p.e_dtype != ORC_BF16 && p.e_dtype != ORC_F32) { g_err = "E must be float-like"; return -2; }
    }
    static const int need_d[] = {0, 1, 0, 1, 1, 2, 1, 2, 2, 0, 1, 1, 2};
    if(p.num_d < need_d[p.cde_op]) { g_err = "too few D tensors for cde_op"; return -2; }

    Ctx c;
    c.p = pp;
    if(int rc = build_b(c)) return rc;

    int m0 = 0, m1 = p.M;
    if(p.m_end > p.m_begin) { m0 = p.m_begin; m1 = std::min(p.m_end, p.M); }
    if(m1 <= m0 || p.N == 0) return 0;

    // make_ParallelTensorFunctor (library/include/ck/library/utility/host_tensor.hpp:219-239):
    // contiguous chunks over threads; here chunks of rows.
    int nt = p.num_threads > 0 ? p.num_threads : int(std::thread::hardware_concurrency());
    nt     = std::max(1, std::min(nt, m1 - m0));
    const int rows = m1 - m0, per = (rows + nt - 1) / nt;
    std::vector<std::thread> th;
    for(int t = 0; t < nt; ++t)
    {
        const int r0 = m0 + t * per, r1 = std::min(m1, r0 + per);
        if(r0 >= r1) break;
        if(is_int) th.emplace_back([&c, r0, r1] { run_rows_int(c, r0, r1); });
        else       th.emplace_back([&c, r0, r1] { run_rows_float(c, r0, r1); });
    }
    for(auto& t : th) t.join();
    return 0;
}

uint16_t orc_f32_to_bf16_trunc(float x) { return f32_to_bf16_trunc(x); }
uint16_t orc_f32_to_bf16_rtn(float x) { return f32_to_bf16_rtn(x); }
float orc_bf16_to_f32(uint16_t x) { return bf16_to_f32(x); }
uint16_t orc_f32_to_f16(float x)
{
    half_t h = f32_to_f16(x);
    uint16_t u;
    std::memcpy(&u, &h, 2);
    return u;
}
float orc_f16_to_f32(uint16_t x)
{
    half_t h;
    std::memcpy(&h, &x, 2);
    return f16_to_f32(h);
}
uint16_t orc_i8_to_bf16(int8_t x) { return i8_to_bf16(x); }
int8_t orc_bf16_to_i8(uint16_t x) { return bf16_to_i8(x); }
float orc_fast_gelu(float x) { return fast_gelu(x); }

// FillUniformDistribution / FillUniformDistributionIntegerValue (fill.hpp:19-95): a FRESH
// std::mt19937(11939) per call, std::uniform_real_distribution<float>(a,b), optional std::round.
void orc_fill_uniform_f32(float* dst, int64_t n, float a, float b, int round_to_int)
{
    std::mt19937 gen(11939);
    std::uniform_real_distribution<float> dis(a, b);
    for(int64_t i = 0; i < n; ++i)
    {
        float v = dis(gen);
        dst[i]  = round_to_int ? std::round(v) : v;
    }
}

// UnsignedWeightPreprocessor<int8_t> (example/64_fpAintB_gemm/common.hpp:56-83): u8 = int8 + 128
void orc_int8_to_biased_u8(const int8_t* src, uint8_t* dst, int64_t n)
{
    for(int64_t i = 0; i < n; ++i) dst[i] = uint8_t(int(src[i]) + 128);
}

} // extern "C"
