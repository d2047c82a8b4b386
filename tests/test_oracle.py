"""CPU tests of the oracle itself: conversion KATs from the reference's own tests, the seeded fills,
and an independent numpy re-derivation of every GEMM mode (bit-exact, same accumulation order)."""
import struct

import numpy as np
import pytest

from oracle import ck_oracle as orc

L = orc.lib()


def f32_bits(x):
    return struct.unpack("<I", struct.pack("<f", x))[0]


# ---- reference KATs: test/data_type/type_convert_const.cpp:17-37 --------------------------------
@pytest.mark.parametrize("x", [0.0, -123.0, 3.981323, 0.2429])
def test_type_convert_const_cases(x):
    bf16_epsilon = 0.0078125
    abs_tol = abs(2 * bf16_epsilon * x)
    y = L.orc_f32_to_bf16_trunc(x)
    assert abs(L.orc_bf16_to_f32(y) - x) <= abs_tol
    # truncation is exactly "bits >> 16" (include/ck/utility/type_convert.hpp:53-64)
    assert y == f32_bits(x) >> 16
    assert f32_bits(L.orc_bf16_to_f32(y)) == (f32_bits(x) >> 16) << 16


def test_bf16_known_bit_patterns():
    assert L.orc_f32_to_bf16_trunc(1.0) == 0x3F80
    assert L.orc_f32_to_bf16_trunc(-123.0) == 0xC2F6
    assert L.orc_f32_to_bf16_trunc(3.981323) == 0x407E      # truncates (RNE would give 0x407F)
    assert L.orc_f32_to_bf16_rtn(3.981323) == 0x407F
    assert L.orc_f32_to_bf16_rtn(1.00390625) == 0x3F80       # tie -> even
    assert L.orc_f32_to_bf16_rtn(1.01171875) == 0x3F82       # tie -> even (up)
    assert L.orc_f32_to_bf16_rtn(float("inf")) == 0x7F80
    assert L.orc_i8_to_bf16(-128) == 0xC300 and L.orc_i8_to_bf16(127) == 0x42FE
    assert L.orc_bf16_to_i8(0xC300) == -128


def test_bf16_helpers_match_oracle():
    rng = np.random.default_rng(0)
    x = rng.standard_normal(4096).astype(np.float32) * 37
    t = orc.bf16_trunc(x)
    r = orc.bf16_rtn(x)
    for i in range(0, 4096, 97):
        assert t[i] == L.orc_f32_to_bf16_trunc(float(x[i]))
        assert r[i] == L.orc_f32_to_bf16_rtn(float(x[i]))
    assert np.array_equal(orc.bf16_trunc(orc.bf16_to_f32(t)), t)


def test_f16_matches_numpy():
    rng = np.random.default_rng(1)
    for v in rng.standard_normal(200).astype(np.float32) * 100:
        assert L.orc_f32_to_f16(float(v)) == int(np.float16(v).view(np.uint16))


def test_fast_gelu_formula():
    # y = x / (1 + exp(x * (-0.071354 x^2 - 1.59577)))  (unary_element_wise_operation.hpp:443-452)
    for x in [-3.0, -0.5, 0.0, 0.25, 1.0, 4.0]:
        ref = x / (1.0 + np.exp(x * (-2 * 0.035677 * x * x - 2 * 0.797885)))
        assert abs(L.orc_fast_gelu(x) - ref) < 1e-6 * max(1.0, abs(ref))


# ---- seeded fill: std::mt19937(11939) + uniform_real_distribution<float> (fill.hpp:27,71) -------
def test_fill_uniform_matches_independent_mt19937():
    n = 1000
    raw = np.frombuffer(np.random.RandomState(11939).bytes(4 * n), dtype="<u4")  # init_genrand(seed)
    canon = raw.astype(np.float32) / np.float32(4294967296.0)
    canon = np.where(canon >= 1.0, np.nextafter(np.float32(1), np.float32(0)), canon)
    ref = canon * np.float32(2.0) + np.float32(-1.0)
    got = orc.fill_uniform(n, -1.0, 1.0)
    assert np.array_equal(got, ref.astype(np.float32))
    goti = orc.fill_uniform(n, -1.0, 1.0, integer=True)
    assert set(np.unique(goti)) <= {-1.0, 0.0, 1.0}
    # std::round = half away from zero
    assert np.array_equal(goti, np.where(ref >= 0, np.floor(ref + 0.5), np.ceil(ref - 0.5)))


# ---- independent numpy evaluation, same k-ascending fp32 accumulation ---------------------------
def seq_gemm_f32(af, bf):
    M, K = af.shape
    acc = np.zeros((M, bf.shape[1]), np.float32)
    for k in range(K):
        acc = (acc + af[:, k:k + 1] * bf[k:k + 1, :]).astype(np.float32)
    return acc


def np_gelu(x):
    x = x.astype(np.float32)
    c1 = np.float32(-2.0 * np.float32(0.035677))
    c2 = np.float32(-2.0 * np.float32(0.797885))
    u = (x * (c1 * x * x + c2)).astype(np.float32)
    return (x / (np.float32(1) + np.exp(u, dtype=np.float32))).astype(np.float32)


def rand_bf16(rng, shape, scale=1.0):
    return orc.bf16_trunc((rng.standard_normal(shape) * scale).astype(np.float32))


@pytest.mark.parametrize("layout_b", ["row", "col"])
@pytest.mark.parametrize("cde", ["pass", "add", "gelu", "add_gelu"])
def test_w8a16_bf16_multiply_vs_numpy(layout_b, cde):
    rng = np.random.default_rng(5)
    M, N, K = 9, 24, 40
    a = rand_bf16(rng, (M, K))
    q = rng.integers(-128, 128, (K, N)).astype(np.int8)
    b = q if layout_b == "row" else np.ascontiguousarray(q.T).T
    s = rand_bf16(rng, (N,), 0.02)
    sv = np.broadcast_to(s, (K, N))       # StrideBs = {StrideB, 0}
    bias = rand_bf16(rng, (N,), 0.5)
    biasv = np.broadcast_to(bias, (M, N))
    op = {"pass": orc.CDE_PASSTHROUGH, "add": orc.CDE_ADD, "gelu": orc.CDE_FASTGELU,
          "add_gelu": orc.CDE_ADD_FASTGELU}[cde]
    ds = (biasv,) if "add" in cde else ()
    e = orc.gemm(a, orc.BF16, b, orc.I8, orc.BF16, b_op=orc.BOP_MULTIPLY, scale=sv, s_dtype=orc.BF16,
                 ds=ds, d_dtypes=(orc.BF16,) * len(ds), cde_op=op)
    bdq = orc.bf16_to_f32(orc.bf16_trunc(q.astype(np.float32) * orc.bf16_to_f32(s)[None, :]))
    c = seq_gemm_f32(orc.bf16_to_f32(a), bdq)
    if "add" in cde:
        c = (c + orc.bf16_to_f32(bias)[None, :]).astype(np.float32)
    if "gelu" in cde:
        ref = orc.bf16_trunc(np_gelu(c))
        # exp() may differ by an ulp between libm and numpy: allow 1 bf16 ulp on a few elements
        assert np.max(np.abs(e.astype(np.int32) - ref.astype(np.int32))) <= 1
    else:
        assert np.array_equal(e, orc.bf16_trunc(c))


def test_cshuffle_bf16_rounds_before_epilogue():
    rng = np.random.default_rng(6)
    M, N, K = 5, 16, 32
    a = rand_bf16(rng, (M, K))
    q = rng.integers(-5, 5, (K, N)).astype(np.int8)
    s = rand_bf16(rng, (N,), 0.3)
    bias = rand_bf16(rng, (N,), 0.5)
    e = orc.gemm(a, orc.BF16, q, orc.I8, orc.BF16, b_op=orc.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)),
                 s_dtype=orc.BF16, ds=(np.broadcast_to(bias, (M, N)),), d_dtypes=(orc.BF16,),
                 cde_op=orc.CDE_ADD, cshuffle_dtype=orc.BF16)
    bdq = orc.bf16_to_f32(orc.bf16_trunc(q.astype(np.float32) * orc.bf16_to_f32(s)[None, :]))
    c = orc.bf16_to_f32(orc.bf16_trunc(seq_gemm_f32(orc.bf16_to_f32(a), bdq)))
    assert np.array_equal(e, orc.bf16_trunc(c + orc.bf16_to_f32(bias)[None, :]))


def test_fpAintB_fp16_vs_numpy():
    # reference_fpAintB_gemm.hpp:110-112 ; example/64 layout: B column-major, scale desc {K,N},{0,1}
    rng = np.random.default_rng(7)
    M, N, K = 7, 24, 48
    a = rng.standard_normal((M, K)).astype(np.float16)
    q = np.ascontiguousarray(rng.integers(-128, 128, (N, K)).astype(np.int8)).T   # [K,N] col-major
    s = (rng.standard_normal(N) * 0.05).astype(np.float16)
    e = orc.gemm(a, orc.F16, q, orc.I8, orc.F16, b_op=orc.BOP_FPAINTB, scale=np.broadcast_to(s, (K, N)),
                 s_dtype=orc.F16)
    bdq = (q.astype(np.float16) * s[None, :]).astype(np.float16)          # one fp16 rounding
    c = seq_gemm_f32(a.astype(np.float32), bdq.astype(np.float32))
    assert np.array_equal(e.view(np.uint16), c.astype(np.float16).view(np.uint16))


@pytest.mark.parametrize("with_zero", [False, True])
def test_w4a16_group_vs_numpy(with_zero):
    rng = np.random.default_rng(8)
    M, N, K, G = 4, 16, 256, 128
    a = rand_bf16(rng, (M, K))
    q = rng.integers(0, 16, (K, N)).astype(np.uint8)
    s = rand_bf16(rng, (K // G, N), 0.01)
    z = rng.integers(0, 16, (K // G, N)).astype(np.uint8) if with_zero else None
    e = orc.gemm(a, orc.BF16, q, orc.U8, orc.BF16, b_op=orc.BOP_GROUP_F32, scale=s, s_dtype=orc.BF16,
                 zero=z, zero_const=8, group_size=G)
    zz = np.repeat(z, G, axis=0).astype(np.float32) if with_zero else np.float32(8)
    bdq = ((q.astype(np.float32) - zz) * np.repeat(orc.bf16_to_f32(s), G, axis=0)).astype(np.float32)
    c = seq_gemm_f32(orc.bf16_to_f32(a), bdq)
    assert np.array_equal(e, orc.bf16_trunc(c))


@pytest.mark.parametrize("cde", ["i32", "mul_clamp", "relu_add_mul_clamp", "mulmul_bf16", "mul2"])
def test_int8_paths_exact(cde):
    rng = np.random.default_rng(9)
    M, N, K = 11, 20, 300
    a = rng.integers(-128, 128, (M, K)).astype(np.int8)
    b = np.ascontiguousarray(rng.integers(-128, 128, (N, K)).astype(np.int8)).T   # col-major B
    acc = a.astype(np.int64) @ b.astype(np.int64)
    assert np.abs(acc).max() < 2 ** 31
    if cde == "i32":
        e = orc.gemm(a, orc.I8, b, orc.I8, orc.I32, cshuffle_dtype=orc.I32)
        assert np.array_equal(e, acc.astype(np.int32))
    elif cde == "mul_clamp":
        e = orc.gemm(a, orc.I8, b, orc.I8, orc.I8, cshuffle_dtype=orc.I32, cde_op=orc.CDE_ACT_MUL_CLAMP,
                     cde_scale=0.03)
        y = np.clip(np.float32(0.03) * acc.astype(np.float32), -128, 127)
        assert np.array_equal(e, np.trunc(y).astype(np.int8))
    elif cde == "relu_add_mul_clamp":
        bias = rng.integers(-5000, 5000, N).astype(np.int32)
        e = orc.gemm(a, orc.I8, b, orc.I8, orc.I8, cshuffle_dtype=orc.I32, ds=(np.broadcast_to(bias, (M, N)),),
                     d_dtypes=(orc.I32,), cde_op=orc.CDE_ADD_ACT_MUL_CLAMP, cde_scale=0.01, act=orc.ACT_RELU)
        y = np.maximum((acc + bias[None, :]).astype(np.float32), 0)
        y = np.clip(np.float32(0.01) * y, -128, 127)
        assert np.array_equal(e, np.trunc(y).astype(np.int8))
    elif cde == "mul2":
        rs = (rng.random(N) * 0.02).astype(np.float32)
        e = orc.gemm(a, orc.I8, b, orc.I8, orc.I8, cshuffle_dtype=orc.I32, ds=(np.broadcast_to(rs, (M, N)),),
                     d_dtypes=(orc.F32,), cde_op=orc.CDE_ACT_MUL2_CLAMP)
        y = np.clip(rs[None, :] * acc.astype(np.float32), -128, 127)
        assert np.array_equal(e, np.trunc(y).astype(np.int8))
    else:
        d0 = (rng.random(N) * 1e-3 + 5e-4).astype(np.float32)   # Row layout, stride 0 -> indexed by n
        d1 = (rng.random(M) * 1e-3 + 5e-4).astype(np.float32)   # Col layout, stride 0 -> indexed by m
        e = orc.gemm(a, orc.I8, b, orc.I8, orc.BF16, cshuffle_dtype=orc.I32,
                     ds=(np.broadcast_to(d0, (M, N)), np.broadcast_to(d1[:, None], (M, N))),
                     d_dtypes=(orc.F32, orc.F32), cde_op=orc.CDE_MULTIPLY_MULTIPLY)
        x = (acc.astype(np.float32) * d0[None, :]).astype(np.float32) * d1[:, None]
        assert np.array_equal(e, orc.bf16_trunc(x.astype(np.float32)))


def test_row_range_and_threads_agree():
    rng = np.random.default_rng(10)
    M, N, K = 33, 16, 64
    a = rand_bf16(rng, (M, K))
    q = rng.integers(-128, 128, (K, N)).astype(np.int8)
    full = orc.gemm(a, orc.BF16, q, orc.I8, orc.BF16, num_threads=1)
    mt = orc.gemm(a, orc.BF16, q, orc.I8, orc.BF16, num_threads=5)
    assert np.array_equal(full, mt)
    part = orc.gemm(a, orc.BF16, q, orc.I8, orc.BF16, m_range=(8, 20))
    assert np.array_equal(part[8:20], full[8:20]) and not part[:8].any() and not part[20:].any()


def test_empty_and_errors():
    a = np.zeros((0, 8), np.uint16)
    q = np.zeros((8, 4), np.int8)
    assert orc.gemm(a, orc.BF16, q, orc.I8, orc.BF16).shape == (0, 4)
    with pytest.raises(ValueError):
        orc.gemm(np.zeros((2, 8), np.float16), orc.F16, q, orc.I8, orc.F16, b_op=orc.BOP_MULTIPLY,
                 scale=np.zeros((8, 4), np.float16), s_dtype=orc.F16)
