// ASan + UBSan pass over the oracle and over the host-only parts of the C++ operator layer (sanitizers row; the
// reference's -fsanitize=address build switch, CMakeLists.txt:111-118).  Test infrastructure: built by `make -C oracle sanitize`
// into oracle/_san/ and run by tests/test_host_logic.py (no GPU, no libckb.so: only header-only host code is exercised).
// Every weight format / CDE op of the oracle on ragged shapes with strided and broadcast operands, m_range sampling, threading;
// Tensor / HostTensorDescriptor / element-wise functors / type conversions of include/ck.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ck/library/utility/host_tensor.hpp"
#include "ck/tensor_operation/gpu/element/element_wise_operation.hpp"
#include "ck/utility/data_type.hpp"
#include "ck_oracle.h"

static int fails = 0;
#define CHECK(x) do { if(!(x)) { std::printf("FAILED: %s (line %d)\n", #x, __LINE__); ++fails; } } while(0)

static void run(int M, int N, int K, int a_dt, int b_dt, int b_op, int e_dt, int cde, int num_d, int threads, bool sample)
{
    const int G = 128, KG = (K + G - 1) / G;
    const size_t m_ = size_t(M), n_ = size_t(N), k_ = size_t(K), kg_ = size_t(KG);
    std::vector<uint16_t> a16(m_ * k_), s16(kg_ * n_), d16(n_), e16(m_ * n_);
    std::vector<int8_t> a8(m_ * k_), b8(k_ * n_), e8(m_ * n_);
    std::vector<uint8_t> z(kg_ * n_);
    std::vector<float> ef(m_ * n_ + 1), df(m_ + 1);
    std::vector<int32_t> ei(m_ * n_);
    for(auto& v : a16) v = orc_f32_to_bf16_trunc(float(std::rand() % 17 - 8) * 0.125f);
    for(auto& v : s16) v = a_dt == ORC_F16 ? orc_f32_to_f16(0.01f + 0.001f * float(std::rand() % 7)) : orc_f32_to_bf16_trunc(0.01f + 0.001f * float(std::rand() % 7));
    if(a_dt == ORC_F16) for(auto& v : a16) v = orc_f32_to_f16(float(std::rand() % 17 - 8) * 0.125f);
    for(auto& v : d16) v = orc_f32_to_bf16_trunc(float(std::rand() % 5));
    for(auto& v : a8) v = int8_t(std::rand() % 255 - 127);
    for(auto& v : b8) v = b_op == ORC_BOP_GROUP_F32 || b_op == ORC_BOP_GROUP_RTN ? int8_t(std::rand() & 15) : int8_t(std::rand() % 255 - 127);
    for(auto& v : z) v = uint8_t(std::rand() & 15);
    for(auto& v : df) v = 1e-3f;
    orc_problem p;
    std::memset(&p, 0, sizeof p);
    p.M = M; p.N = N; p.K = K;
    p.a_dtype = a_dt; p.a = a_dt == ORC_I8 ? (const void*)a8.data() : (const void*)a16.data(); p.a_sm = K; p.a_sk = 1;
    p.b_dtype = b_dt; p.b = b8.data(); p.b_sk = 1; p.b_sn = K;            // K-major (column-major B)
    p.b_op = b_op;
    if(b_op != ORC_BOP_PASSTHROUGH)
    {
        p.s_dtype = a_dt; p.scale = s16.data();
        const bool grp = b_op == ORC_BOP_GROUP_F32 || b_op == ORC_BOP_GROUP_RTN;
        p.s_sk = grp ? N : 0; p.s_sn = 1;                                  // per-channel scales broadcast along K (stride 0)
        if(grp) { p.zero = z.data(); p.z_sk = N; p.z_sn = 1; p.zero_const = 8; p.group_size = G; }
    }
    p.cshuffle_dtype = a_dt == ORC_I8 ? ORC_I32 : ORC_F32;
    p.num_d = num_d;
    if(num_d >= 1) { p.d_dtype[0] = a_dt == ORC_I8 ? ORC_F32 : ORC_BF16; p.d[0] = a_dt == ORC_I8 ? (const void*)ef.data() : (const void*)d16.data(); p.d_sm[0] = 0; p.d_sn[0] = 1; }
    if(num_d >= 2) { p.d_dtype[1] = ORC_F32; p.d[1] = df.data(); p.d_sm[1] = 1; p.d_sn[1] = 0; }
    p.cde_op = cde; p.cde_scale = 0.01f; p.act = ORC_ACT_RELU;
    p.e_dtype = e_dt;
    p.e = e_dt == ORC_F32 ? (void*)ef.data() : e_dt == ORC_I32 ? (void*)ei.data() : e_dt == ORC_I8 ? (void*)e8.data() : (void*)e16.data();
    if(a_dt == ORC_I8 && num_d >= 1) { p.d[0] = df.data(); p.d_sm[0] = 0; p.d_sn[0] = 0; }   // one broadcast scalar
    p.e_sm = N; p.e_sn = 1;
    if(sample) { p.m_begin = M / 2; p.m_end = M; }
    p.num_threads = threads;
    const int rc = orc_gemm(&p);
    if(rc != 0) std::printf("orc_gemm rc %d: %s\n", rc, orc_last_error());
    CHECK(rc == 0);
}

int main()
{
    // oracle: every weight format, ragged shapes, 1 and many threads, row sampling
    run(5, 37, 130, ORC_BF16, ORC_I8, ORC_BOP_MULTIPLY, ORC_BF16, ORC_CDE_ADD, 1, 1, false);
    run(33, 64, 256, ORC_BF16, ORC_I8, ORC_BOP_MULTIPLY, ORC_BF16, ORC_CDE_ADD_FASTGELU, 1, 0, true);
    run(7, 48, 128, ORC_F16, ORC_I8, ORC_BOP_FPAINTB, ORC_F16, ORC_CDE_PASSTHROUGH, 0, 3, false);
    run(3, 32, 256, ORC_BF16, ORC_U8, ORC_BOP_GROUP_F32, ORC_F32, ORC_CDE_PASSTHROUGH, 0, 0, false);
    run(9, 32, 384, ORC_BF16, ORC_U8, ORC_BOP_GROUP_RTN, ORC_BF16, ORC_CDE_PASSTHROUGH, 0, 2, true);
    run(6, 24, 100, ORC_BF16, ORC_I8, ORC_BOP_PASSTHROUGH, ORC_BF16, ORC_CDE_FASTGELU, 0, 0, false);
    run(11, 40, 96, ORC_I8, ORC_I8, ORC_BOP_PASSTHROUGH, ORC_I32, ORC_CDE_PASSTHROUGH, 0, 0, false);
    run(11, 40, 96, ORC_I8, ORC_I8, ORC_BOP_PASSTHROUGH, ORC_I8, ORC_CDE_ACT_MUL_CLAMP, 0, 4, false);
    run(0, 16, 64, ORC_BF16, ORC_I8, ORC_BOP_PASSTHROUGH, ORC_BF16, ORC_CDE_PASSTHROUGH, 0, 0, false);   // empty problem
    // conversions at the edges (UBSan: no out-of-range float -> int casts, no signed overflow)
    CHECK(orc_bf16_to_f32(orc_f32_to_bf16_trunc(1.0f)) == 1.0f);
    (void)orc_f32_to_bf16_rtn(3.4e38f); (void)orc_f32_to_bf16_trunc(-3.4e38f); (void)orc_f32_to_f16(1e9f); (void)orc_f32_to_f16(-1e-9f);
    (void)orc_bf16_to_i8(orc_f32_to_bf16_trunc(127.0f)); (void)orc_fast_gelu(80.f); (void)orc_fast_gelu(-80.f);
    // header-only host layer: Tensor with strides, functors, type_convert
    {
        using namespace ck::literals;
        Tensor<ck::bhalf_t> t(HostTensorDescriptor({5_uz, 7_uz}, {1_uz, 5_uz}));   // column-major view
        t.GenerateTensorValue([](auto m, auto n) { return ck::type_convert<ck::bhalf_t>(float(m) - float(n)); });
        CHECK(ck::type_convert<float>(t(4, 6)) == -2.0f);
        float y = 0;
        ck::tensor_operation::element_wise::AddFastGelu{}(y, 1.0f, 2.0f);
        CHECK(y > 2.9f && y < 3.1f);
        ck::bhalf_t yb{};
        ck::tensor_operation::element_wise::Multiply{}(yb, int8_t(-128), ck::type_convert<ck::bhalf_t>(0.5f));
        CHECK(ck::type_convert<float>(yb) == -64.0f);
        int8_t q = 0;
        ck::tensor_operation::element_wise::Activation_Mul_Clamp<ck::tensor_operation::element_wise::PassThrough> amc{1000.f, {}};
        amc(q, int32_t(2000000));
        CHECK(q == 127);
    }
    std::printf(fails ? "sanitize_main: %d failures\n" : "sanitize_main: OK\n", fails);
    return fails ? 1 : 0;
}
