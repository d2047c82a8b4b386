"""GPU parity tests (pytest -m gpu): every GEMM goes through the C-ABI (libckb.so) and is compared with
the CPU oracle / the committed golden vectors.

Tolerances (written here, as the spec asks):
  * int8 x int8: int32 accumulators and everything derived from them -> BIT-EXACT.
  * fp paths (fp32 accumulate, fp16/bf16 output): |o - r| <= atol + rtol*|r| with rtol = atol = 1e-2
    (check_err criterion, library/include/ck/library/utility/check_err.hpp:54).  In practice the kernels
    agree with the oracle to ~1e-5 before the final 16-bit rounding; the 1e-2 budget only has to absorb
    one-ulp flips of the truncating bf16 store.  The tighter figure is asserted on fp32 outputs.
"""
import glob
import os

import numpy as np
import pytest

from tests import helpers as H

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RTOL = ATOL = 1e-2


def orc():
    from oracle import ck_oracle
    return ck_oracle


def rand_bf16(rng, shape, scale=1.0):
    return orc().bf16_trunc((rng.standard_normal(shape) * scale).astype(np.float32))


# ------------------------------------------------------------------------------------------ golden
@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "*.npz"))), ids=os.path.basename)
def test_golden(path):
    g = np.load(path)
    kind = str(g["kind"])
    if kind == "w4a16":
        f16 = bool(g["f16"])
        z = g["z"] if g["z"].size else None
        got, _ = H.run_w4a16(g["a"], g["q"], g["s"], z, f16=f16)
        bad, err = (H.f_close(got, g["e"], RTOL, ATOL) if f16 else H.bf16_close(got, g["e"], RTOL, ATOL))
        assert bad == 0, f"max err {err}"
        got32, _ = H.run_w4a16(g["a"], g["q"], g["s"], z, f16=f16, out_f32=True)
        # the decode kernel IS the fp32-dequant definition (ORC_BOP_GROUP_F32: exact (q-zp), group scale applied in fp32,
        # CK's ab_scale host pattern): only fp32 accumulation-order noise is left
        err = float(np.abs(got32 - g["e_f32"]).max())
        assert err <= 2e-5 * float(np.abs(g["e_f32"]).max()) + 1e-5, f"max err vs fp32-dequant oracle {err}"
    elif kind == "w8a16_bf16":
        got, _ = H.run_w8a16_bf16(g["a"], g["qn"], g["s"], g["bias"], str(g["cde"]))
        bad, err = H.bf16_close(got, g["e"], RTOL, ATOL)
        assert bad == 0, f"max err {err}"
    elif kind == "fpaintb":
        got, _ = H.run_fpaintb(g["a"], g["qn"], g["s"])
        bad, err = H.f_close(got, g["e"], RTOL, ATOL)
        assert bad == 0, f"max err {err}"
    elif kind == "int8":
        from composable_kernel_b200 import capi
        M, N = g["acc"].shape
        acc, _ = H.run_int8(g["a"], g["bn"], "i32")
        assert np.array_equal(acc, g["acc"])
        e_q, _ = H.run_int8(g["a"], g["bn"], "i8", cde_op=capi.CDE_ACT_MUL_CLAMP, cde_scale=0.03)
        assert np.array_equal(e_q, g["e_q"])
        e_bq, _ = H.run_int8(g["a"], g["bn"], "i8", cde_op=capi.CDE_ADD_ACT_MUL_CLAMP, ds=(g["bias"],),
                             d_strides=((0, 1),), cde_scale=0.01, act=capi.ACT_RELU)
        assert np.array_equal(e_bq, g["e_bq"])
        e_mm, _ = H.run_int8(g["a"], g["bn"], "bf16", cde_op=capi.CDE_MULTIPLY_MULTIPLY, ds=(g["d0"], g["d1"]),
                             d_strides=((0, 1), (1, 0)))
        assert np.array_equal(e_mm, g["e_mm"])
    else:
        raise AssertionError(kind)


# ------------------------------------------------------------------------------------------ int8 (bit exact)
@pytest.mark.parametrize("M,N,K", [(128, 256, 128), (1, 16, 16), (200, 300, 144), (640, 1280, 512), (1000, 520, 2064)])
def test_int8_accumulators_bit_exact(M, N, K):
    rng = np.random.default_rng(M * 7 + N)
    a = rng.integers(-128, 128, (M, K)).astype(np.int8)
    bn = rng.integers(-128, 128, (N, K)).astype(np.int8)
    got, op = H.run_int8(a, bn, "i32")
    assert "i8" in op.kernel_name
    ref = a.astype(np.int64) @ bn.astype(np.int64).T
    assert np.array_equal(got.astype(np.int64), ref)


@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (1, 16, 16), (300, 520, 144), (1000, 520, 2064), (2304, 4864, 256)])
def test_int8_cta_pair_kernel_bit_exact(M, N, K, monkeypatch):
    """The cta_group::2 kernel (256x256 tiles per CTA pair) on ragged and multi-tile shapes; the last shape takes it by
    default (>= 74 pair tiles), the others force it with CKB_SS_2CTA=1."""
    if (M, N) != (2304, 4864):
        monkeypatch.setenv("CKB_SS_2CTA", "1")
    rng = np.random.default_rng(M * 11 + N)
    a = rng.integers(-128, 128, (M, K)).astype(np.int8)
    bn = rng.integers(-128, 128, (N, K)).astype(np.int8)
    got, op = H.run_int8(a, bn, "i32")
    assert "2cta" in op.kernel_name, op.kernel_name
    ref = a.astype(np.int64) @ bn.astype(np.int64).T
    assert np.array_equal(got.astype(np.int64), ref)


def test_int8_extreme_values_full_size_k():
    # all operands at -128: |acc| = K * 2^14; K = 8192 -> 2^27, the cfg4 bound
    M, N, K = 128, 256, 8192
    a = np.full((M, K), -128, np.int8)
    bn = np.full((N, K), -128, np.int8)
    got, _ = H.run_int8(a, bn, "i32")
    assert (got == K * 16384).all()


def test_int8_cfg4_row_sample():
    """cfg4 shape class (large M=N=K) checked on sampled rows against the oracle: bit-exact."""
    o = orc()
    from composable_kernel_b200 import capi
    M = N = K = 2048
    rng = np.random.default_rng(4)
    a = rng.integers(-128, 128, (M, K)).astype(np.int8)
    bn = rng.integers(-128, 128, (N, K)).astype(np.int8)
    d0 = (rng.random(N) * 1e-3 + 5e-4).astype(np.float32)
    d1 = (rng.random(M) * 1e-3 + 5e-4).astype(np.float32)
    acc, _ = H.run_int8(a, bn, "i32")
    e, _ = H.run_int8(a, bn, "bf16", cde_op=capi.CDE_MULTIPLY_MULTIPLY, ds=(d0, d1), d_strides=((0, 1), (1, 0)))
    rows = (0, 2048)
    for r0 in (0, 1000, 1984):
        ref = o.gemm(a, o.I8, bn.T, o.I8, o.I32, cshuffle_dtype=o.I32, m_range=(r0, r0 + 64))
        assert np.array_equal(acc[r0:r0 + 64], ref[r0:r0 + 64])
        refe = o.gemm(a, o.I8, bn.T, o.I8, o.BF16, cshuffle_dtype=o.I32, m_range=(r0, r0 + 64),
                      ds=(np.broadcast_to(d0, (M, N)), np.broadcast_to(d1[:, None], (M, N))),
                      d_dtypes=(o.F32, o.F32), cde_op=o.CDE_MULTIPLY_MULTIPLY)
        assert np.array_equal(e[r0:r0 + 64], refe[r0:r0 + 64])
    assert rows


# ------------------------------------------------------------------------------------------ W4A16 decode
@pytest.mark.parametrize("M,N,K,zp", [(1, 128, 128, False), (16, 128, 128, True), (4, 128, 1024, False),
                                      (16, 4096, 4096, True), (7, 1000, 384, False), (48, 512, 1024, True),
                                      (1, 11008, 4096, False), (16, 4096, 11008, False)])
def test_w4a16_decode_vs_oracle(M, N, K, zp):
    o = orc()
    rng = np.random.default_rng(N + K + M)
    a = rand_bf16(rng, (M, K))
    q = rng.integers(0, 16, (K, N)).astype(np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N)) * 0.015 + 0.005).astype(np.float32))
    z = rng.integers(0, 16, (K // 128, N)).astype(np.uint8) if zp else None
    if zp and N % 16:
        pytest.skip("zero points need N % 16 == 0")
    got, op = H.run_w4a16(a, q, s, z, out_f32=True)
    assert "w4g128" in op.kernel_name
    ref = o.gemm(a, o.BF16, q, o.U8, o.F32, b_op=o.BOP_GROUP_F32, scale=s, s_dtype=o.BF16, zero=z, zero_const=8,
                 group_size=128)
    bad, err = H.f_close(got, ref, 1e-4, 2e-4 * np.sqrt(K / 128))  # fp32 accumulation-order noise only
    assert bad == 0, f"max err {err}"
    gotb, _ = H.run_w4a16(a, q, s, z)
    refb = o.bf16_trunc(ref)
    bad, err = H.bf16_close(gotb, refb, RTOL, ATOL)
    assert bad == 0, f"max err {err}"


@pytest.mark.parametrize("M,N,K,zp,cde", [(1, 4096, 4096, False, "add"), (4, 1000, 384, False, "add_gelu"), (16, 4096, 4096, True, "add"),
                                          (8, 11008, 4096, False, "gelu"), (40, 512, 1024, True, "add")])
def test_w4a16_decode_functor_epilogue_vs_oracle(M, N, K, zp, cde):
    """W4A16 decode rows with a CDE functor (bias Add / FastGelu / AddFastGelu, the epilogues of example 60 / client 30): the
    rotating-producer kernel serves them with its functor-epilogue instantiation (`_epi`), the plain epilogue with the lean one."""
    from composable_kernel_b200 import capi
    o = orc()
    rng = np.random.default_rng(N + K + 7 * M)
    a = rand_bf16(rng, (M, K))
    q = rng.integers(0, 16, (K, N)).astype(np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N)) * 0.015 + 0.005).astype(np.float32))
    z = rng.integers(0, 16, (K // 128, N)).astype(np.uint8) if zp else None
    bias = rand_bf16(rng, (N,))
    op_dev = {"add": capi.CDE_ADD, "gelu": capi.CDE_FASTGELU, "add_gelu": capi.CDE_ADD_FASTGELU}[cde]
    op_orc = {"add": o.CDE_ADD, "gelu": o.CDE_FASTGELU, "add_gelu": o.CDE_ADD_FASTGELU}[cde]
    has_d = "add" in cde
    got, op = H.run_w4a16(a, q, s, z, ds=(H.to_dev(bias, True),) if has_d else (), d_strides=((0, 1),) if has_d else (), cde_op=op_dev)
    ref = o.gemm(a, o.BF16, q, o.U8, o.BF16, b_op=o.BOP_GROUP_F32, scale=s, s_dtype=o.BF16, zero=z, zero_const=8, group_size=128,
                 ds=(np.broadcast_to(bias, (M, N)),) if has_d else (), d_dtypes=(o.BF16,) if has_d else (), cde_op=op_orc)
    bad, err = H.bf16_close(got, ref, RTOL, ATOL)
    assert bad == 0, f"{op.kernel_name}: max err {err}"
    if "wq_gemv2" in op.kernel_name:
        assert op.kernel_name.endswith("_epi"), op.kernel_name
        _, plain = H.run_w4a16(a, q, s, z)
        assert not plain.kernel_name.endswith("_epi"), plain.kernel_name


def test_w4a16_linearity_and_zero():
    """Size-independent properties at the BASELINE decode shape: E(0) = 0 and E(2a) = 2 E(a) exactly
    (power-of-two scaling commutes with every rounding on the path)."""
    o = orc()
    rng = np.random.default_rng(77)
    M, N, K = 16, 11008, 4096
    a = rand_bf16(rng, (M, K))
    q = rng.integers(0, 16, (K, N)).astype(np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N)) * 0.015 + 0.005).astype(np.float32))
    e1, _ = H.run_w4a16(a, q, s, out_f32=True)
    a2 = o.bf16_trunc(o.bf16_to_f32(a) * 2)
    e2, _ = H.run_w4a16(a2, q, s, out_f32=True)
    # no atomics anywhere on the path: the summation order is fixed, so power-of-two scaling is exact
    assert np.array_equal(e2, 2 * e1)
    e0, _ = H.run_w4a16(np.zeros_like(a), q, s, out_f32=True)
    assert not e0.any()


# ------------------------------------------------------------------------------------------ W8A16
@pytest.mark.parametrize("M,N,K,cde", [(8, 256, 512, "add"), (16, 128, 256, "pass"), (33, 384, 640, "add_gelu"),
                                       (200, 384, 512, "add"), (5, 256, 128, "gelu")])
def test_w8a16_bf16_multiply_vs_oracle(M, N, K, cde):
    o = orc()
    rng = np.random.default_rng(M + N + K)
    a = rand_bf16(rng, (M, K))
    qn = rng.integers(-128, 128, (N, K)).astype(np.int8)
    s = o.bf16_trunc((rng.random(N) * 0.02 + 0.001).astype(np.float32))
    bias = rand_bf16(rng, (N,))
    got, op = H.run_w8a16_bf16(a, qn, s, bias, cde)
    op_id = {"pass": o.CDE_PASSTHROUGH, "add": o.CDE_ADD, "gelu": o.CDE_FASTGELU, "add_gelu": o.CDE_ADD_FASTGELU}[cde]
    ds = (np.broadcast_to(bias, (M, N)),) if "add" in cde else ()
    ref = o.gemm(a, o.BF16, qn.T, o.I8, o.BF16, b_op=o.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)), s_dtype=o.BF16,
                 ds=ds, d_dtypes=(o.BF16,) * len(ds), cde_op=op_id)
    bad, err = H.bf16_close(got, ref, RTOL, ATOL)
    assert bad == 0, f"max err {err}"


@pytest.mark.parametrize("M,N,K", [(16, 256, 512), (1, 128, 128), (64, 384, 256)])
def test_fpaintb_fp16_vs_oracle(M, N, K):
    o = orc()
    rng = np.random.default_rng(M * N)
    a = rng.standard_normal((M, K)).astype(np.float16)
    qn = rng.integers(-128, 128, (N, K)).astype(np.int8)
    s = (rng.random(N) * 0.02 + 0.001).astype(np.float16)
    got, _ = H.run_fpaintb(a, qn, s)
    ref = o.gemm(a, o.F16, qn.T, o.I8, o.F16, b_op=o.BOP_FPAINTB, scale=np.broadcast_to(s, (K, N)), s_dtype=o.F16)
    bad, err = H.f_close(got, ref, RTOL, ATOL)
    assert bad == 0, f"max err {err}"


def test_fpaintb_example64_integer_inputs_exact():
    """example/64 init_method=1: A, B, scale in {-1,0,1} from FillUniformDistributionIntegerValue with a
    fresh mt19937(11939) per tensor (run_gemm_example.inc:33-50) -> every value is exact, so the device
    result must EQUAL the oracle (this is how the reference pins the path with its 1e-3 check)."""
    o = orc()
    M, N, K = 64, 256, 512
    a = o.fill_uniform(M * K, -1, 1, integer=True).reshape(M, K).astype(np.float16)
    qn = o.fill_uniform(N * K, -1, 1, integer=True).reshape(N, K).astype(np.int8)  # column-major B storage
    s = o.fill_uniform(N, -1, 1, integer=True).astype(np.float16)
    got, _ = H.run_fpaintb(a, qn, s)
    ref = o.gemm(a, o.F16, qn.T, o.I8, o.F16, b_op=o.BOP_FPAINTB, scale=np.broadcast_to(s, (K, N)), s_dtype=o.F16)
    assert np.array_equal(got.view(np.uint16), ref.view(np.uint16))


# ------------------------------------------------------------------------------------------ API behaviour
def test_empty_problem_and_unsupported():
    import torch
    from composable_kernel_b200 import capi, ops
    a = torch.zeros(0, 256, dtype=torch.bfloat16, device="cuda")
    b = torch.zeros(128, 256, dtype=torch.int8, device="cuda")
    e = torch.zeros(0, 128, dtype=torch.bfloat16, device="cuda")
    op = ops.GemmOp(a, b, e)
    assert op.supported
    op.run()  # no-op
    # K not a multiple of 16 bytes for int8 -> IsSupportedArgument == false, Run raises
    a = torch.zeros(4, 100, dtype=torch.int8, device="cuda")
    b = torch.zeros(8, 100, dtype=torch.int8, device="cuda")
    e = torch.zeros(4, 8, dtype=torch.int32, device="cuda")
    op = ops.GemmOp(a, b, e)
    assert not op.supported
    with pytest.raises(capi.CkbError):
        op.run()


def test_host_buffer_entry_point_matches_device_path():
    """ckb_gemm_run_host (pinned host A/E, device-resident weights) gives the same bits as the device path."""
    import ctypes as C
    import torch
    from composable_kernel_b200 import capi, ops
    o = orc()
    rng = np.random.default_rng(5)
    M, N, K = 4, 512, 1024
    a = rand_bf16(rng, (M, K))
    q = rng.integers(0, 16, (K, N)).astype(np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N)) * 0.015 + 0.005).astype(np.float32))
    ref, op = H.run_w4a16(a, q, s)
    p = capi.GemmProblem()
    C.memmove(C.byref(p), C.byref(op.problem), C.sizeof(p))
    ha = torch.from_numpy(a.view(np.int16)).pin_memory()
    he = torch.zeros(M, N, dtype=torch.int16).pin_memory()
    p.a, p.e = ha.data_ptr(), he.data_ptr()
    staging = torch.empty(int(capi.lib().ckb_gemm_host_staging_bytes(C.byref(p))), dtype=torch.uint8, device="cuda")
    ws = op.workspace.data_ptr() if op.workspace is not None else None
    capi.check(capi.lib().ckb_gemm_run_host(C.byref(p), staging.data_ptr(), ws,
                                            torch.cuda.current_stream().cuda_stream))
    assert np.array_equal(he.numpy().view(np.uint16), ref)
    # asynchronous form: three calls share one staging buffer (stream order), one synchronisation at the end
    hes = [torch.zeros(M, N, dtype=torch.int16).pin_memory() for _ in range(3)]
    for h in hes:
        p.e = h.data_ptr()
        capi.check(capi.lib().ckb_gemm_run_host_async(C.byref(p), staging.data_ptr(), ws,
                                                      torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    for h in hes:
        assert np.array_equal(h.numpy().view(np.uint16), ref)


# ------------------------------------------------------------------------------------------ prefill-shaped (M > 64)
@pytest.mark.parametrize("M,N,K", [(256, 512, 1024), (130, 384, 392), (1024, 256, 128)])
def test_prefill_w8a16_and_w4a16_vs_oracle(M, N, K):
    o = orc()
    rng = np.random.default_rng(M + N)
    a = rand_bf16(rng, (M, K))
    qn = rng.integers(-128, 128, (N, K + (-K) % 16)).astype(np.int8)[:, :K]   # ldb padded to 16
    s = o.bf16_trunc((rng.random(N) * 0.02 + 0.001).astype(np.float32))
    bias = rand_bf16(rng, (N,))
    import torch
    from composable_kernel_b200 import capi, ops
    e = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    bd = H.to_dev(np.ascontiguousarray(np.pad(qn, ((0, 0), (0, (-K) % 16)))))[:, :K]
    op = ops.GemmOp(H.to_dev(a, True), bd, e, b_op=capi.BOP_MULTIPLY, scale=H.to_dev(s, True), ds=(H.to_dev(bias, True),),
                    d_strides=((0, 1),), cde_op=capi.CDE_ADD)
    assert op.supported and "prefill" in op.kernel_name, (op.why, op.kernel_name)
    op.run()
    torch.cuda.synchronize()
    ref = o.gemm(a, o.BF16, qn.T, o.I8, o.BF16, b_op=o.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)), s_dtype=o.BF16,
                 ds=(np.broadcast_to(bias, (M, N)),), d_dtypes=(o.BF16,), cde_op=o.CDE_ADD)
    bad, err = H.bf16_close(H.from_dev_bf16(e), ref, RTOL, ATOL)
    assert bad == 0, f"w8 max err {err}"
    if K % 128 == 0:
        q = rng.integers(0, 16, (K, N)).astype(np.uint8)
        sg = o.bf16_trunc((rng.random((K // 128, N)) * 0.015 + 0.005).astype(np.float32))
        z = rng.integers(0, 16, (K // 128, N)).astype(np.uint8)
        got, op4 = H.run_w4a16(a, q, sg, z, rtn=True)
        assert "prefill" in op4.kernel_name
        ref4 = o.gemm(a, o.BF16, q, o.U8, o.BF16, b_op=o.BOP_GROUP_RTN, scale=sg, s_dtype=o.BF16, zero=z, group_size=128)
        bad, err = H.bf16_close(got, ref4, RTOL, ATOL)
        assert bad == 0, f"w4 max err {err}"


def test_prefill_fpaintb_cfg1_row_sample():
    """cfg1 class (example 64 default shape family) on a row sample against the oracle."""
    o = orc()
    rng = np.random.default_rng(64)
    M, N, K = 384, 512, 1024
    a = rng.standard_normal((M, K)).astype(np.float16)
    qn = rng.integers(-128, 128, (N, K)).astype(np.int8)
    s = (rng.random(N) * 0.02 + 0.001).astype(np.float16)
    got, op = H.run_fpaintb(a, qn, s)
    assert "prefill" in op.kernel_name
    ref = o.gemm(a, o.F16, qn.T, o.I8, o.F16, b_op=o.BOP_FPAINTB, scale=np.broadcast_to(s, (K, N)), s_dtype=o.F16)
    bad, err = H.f_close(got, ref, RTOL, ATOL)
    assert bad == 0, f"max err {err}"


# ------------------------------------------------------------------------------------------ grouped fixed-NK (client 31)
@pytest.mark.parametrize("layout,cde,N,K,Ms", [
    ("row", "add_gelu", 1024, 512, [32 + (7 * i * i) % 32 for i in range(16)]),   # client_example/31 shape
    ("row", "add", 256, 192, [1, 64, 65, 0, 130]),
    ("col", "pass", 192, 320, [5, 77]),
    ("col", "gelu", 128, 64, [64]),
    ("row", "pass", 80, 48, [3, 9, 200]),                                           # ragged N and K tiles
    ("col", "add", 80, 48, [17]),
])
def test_grouped_fixed_nk_vs_oracle(layout, cde, N, K, Ms):
    import torch
    from composable_kernel_b200 import capi, ops
    o = orc()
    rng = np.random.default_rng(N + K + len(Ms))
    cde_c = {"pass": capi.CDE_PASSTHROUGH, "add": capi.CDE_ADD, "gelu": capi.CDE_FASTGELU, "add_gelu": capi.CDE_ADD_FASTGELU}[cde]
    cde_o = {"pass": o.CDE_PASSTHROUGH, "add": o.CDE_ADD, "gelu": o.CDE_FASTGELU, "add_gelu": o.CDE_ADD_FASTGELU}[cde]
    hs, devs = [], ([], [], [], [], [])
    for M in Ms:
        a = rand_bf16(rng, (M, K))
        q_kn = rng.integers(-128, 128, (K, N)).astype(np.int8)
        s = o.bf16_trunc((rng.random(N) * 0.02 + 0.001).astype(np.float32))
        bias = rand_bf16(rng, (N,))
        hs.append((a, q_kn, s, bias))
        devs[0].append(H.to_dev(a, True).reshape(M, K))
        devs[1].append(H.to_dev(q_kn if layout == "row" else q_kn.T))
        devs[2].append(H.to_dev(s, True))
        devs[3].append(H.to_dev(bias, True))
        devs[4].append(torch.full((M, N), 7.0, dtype=torch.bfloat16, device="cuda"))
    op = ops.GroupedGemmOp(devs[0], devs[1], devs[2], devs[3] if "add" in cde else None, devs[4],
                           b_layout=capi.ROW if layout == "row" else capi.COL, cde_op=cde_c, d0_stride=0)
    assert op.supported, op.why
    op.run()
    torch.cuda.synchronize()
    for (a, q_kn, s, bias), e in zip(hs, devs[4]):
        M = a.shape[0]
        if M == 0:
            continue
        ds = (np.broadcast_to(bias, (M, N)),) if "add" in cde else ()
        ref = o.gemm(a, o.BF16, q_kn, o.I8, o.BF16, b_op=o.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)), s_dtype=o.BF16,
                     ds=ds, d_dtypes=(o.BF16,) * len(ds), cde_op=cde_o)
        bad, err = H.bf16_close(H.from_dev_bf16(e), ref, RTOL, ATOL)
        assert bad == 0, f"group M={M}: max err {err}"


def test_grouped_unsupported_and_null_args():
    import ctypes as C
    import torch
    from composable_kernel_b200 import capi, ops
    a = [torch.zeros(4, 40, dtype=torch.bfloat16, device="cuda")]
    b = [torch.zeros(40, 64, dtype=torch.int8, device="cuda")]
    e = [torch.zeros(4, 64, dtype=torch.bfloat16, device="cuda")]
    op = ops.GroupedGemmOp(a, b, None, None, e)     # K = 40 is not a multiple of 16
    assert not op.supported and "multiples of 16" in op.why
    with pytest.raises(capi.CkbError):
        op.run()
    a = [torch.zeros(4, 48, dtype=torch.bfloat16, device="cuda")]
    b = [torch.zeros(48, 64, dtype=torch.int8, device="cuda")]
    op = ops.GroupedGemmOp(a, b, None, None, e)
    assert op.supported
    # Run without SetDeviceKernelArgs: the reference throws (device_grouped_gemm_multi_abd_xdl_fixed_nk.hpp:597-600)
    rc = capi.lib().ckb_grouped_gemm_run(C.byref(op.problem), None, 1, 4, None)
    assert rc == capi.ERR_INVALID


# ------------------------------------------------------------------------------------------ layout tool, converter identities
@pytest.mark.parametrize("K,N", [(64, 16), (192, 100), (4096, 512)])
def test_device_prepack_equals_host_packer(K, N):
    """ckb_prepack_b_int4 (device) and composable_kernel_b200.packing.pack_int4 (numpy, offline tool) write the same bytes"""
    import torch
    from composable_kernel_b200 import ops, packing
    rng = np.random.default_rng(K * N)
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    dev = ops.prepack_int4(H.to_dev(q)).cpu().numpy()
    host = packing.pack_int4(q).reshape(-1).view(np.uint8)
    assert dev.size == host.size and np.array_equal(dev, host)


def test_gptq_checkpoint_import_runs_through_the_gemm():
    """AutoGPTQ qweight [K/8, N] int32 -> codes -> host packer -> device GEMM == oracle on the same codes"""
    import torch
    from composable_kernel_b200 import capi, ops, packing
    o = orc()
    rng = np.random.default_rng(8)
    M, N, K = 4, 256, 512
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    qweight = np.zeros((K // 8, N), np.uint32)
    for i in range(8):
        qweight |= q[i::8].astype(np.uint32) << np.uint32(4 * i)
    packed = packing.pack_int4(packing.codes_from_gptq(qweight.view(np.int32)))
    a = rand_bf16(rng, (M, K))
    s = o.bf16_trunc((rng.random((K // 128, N)) * 0.015 + 0.005).astype(np.float32))
    e = torch.zeros(M, N, dtype=torch.float32, device="cuda")
    op = ops.GemmOp(H.to_dev(a, True), torch.from_numpy(packed.reshape(-1).view(np.uint8)).cuda(), e, b_dtype=capi.I4, b_layout=capi.PREPACKED,
                    b_op=capi.BOP_GROUP, N=N, K=K, scale=H.to_dev(s, True), zero_const=8, group_size=128)
    assert op.supported, op.why
    op.run()
    torch.cuda.synchronize()
    ref = o.gemm(a, o.BF16, q, o.U8, o.F32, b_op=o.BOP_GROUP_F32, scale=s, s_dtype=o.BF16, zero_const=8, group_size=128)
    bad, err = H.f_close(e.cpu().numpy(), ref, 1e-4, 1e-3)
    assert bad == 0, err


@pytest.mark.parametrize("M", [1, 128])
def test_device_int8_converters_all_256_codes(M):
    """Reference-held identities on the DEVICE kernels (decode and prefill variants): with a one-hot activation row the GEMM
    returns the dequantised weight itself, so every int8 code must come back as fp16(q) * s (fpAintB: the 0x6400|u8 - 0x6480
    converter, unary_element_wise_operation.hpp:1061-1091) and as trunc_bf16(float(q) * float(s)) (Multiply<bhalf_t,int8_t,bhalf_t>,
    binary_element_wise_operation.hpp:184-190) -- bit exact, all 256 codes against a grid of scales."""
    o = orc()
    K, N = 128, 2560
    qs = np.arange(-128, 128, dtype=np.int32)
    code = np.tile(qs, N // 256).astype(np.int8)                      # column n carries code n % 256
    scales_f = np.array([1.0, 0.5, 0.0078125, 0.013, 0.0999, 1.7, 3.14159, 1e-3, 2.0, 0.25], np.float32)
    s_col = np.repeat(scales_f, 256)                                  # column block j uses scale j
    qn = np.zeros((N, K), np.int8)
    qn[:, 3] = code                                                   # only k = 3 is populated
    a16 = np.zeros((M, K), np.float16); a16[:, 3] = 1.0
    got, op = H.run_fpaintb(a16, qn, s_col.astype(np.float16))
    want = (code.astype(np.float16) * s_col.astype(np.float16))       # one fp16 rounding of q * s (reference_fpAintB_gemm.hpp:110)
    assert np.array_equal(got[0].view(np.uint16), want.view(np.uint16)), op.kernel_name
    abf = np.zeros((M, K), np.float32); abf[:, 3] = 1.0
    sb = o.bf16_trunc(s_col)
    gotb, opb = H.run_w8a16_bf16(o.bf16_trunc(abf), qn, sb, np.zeros(N, np.uint16), "pass")
    prod = code.astype(np.float32) * o.bf16_to_f32(sb)
    wantb = (prod.view(np.uint32) >> 16).astype(np.uint16)
    assert np.array_equal(gotb[0], wantb), opb.kernel_name


def test_torch_custom_ops_match_the_oracle():
    import torch
    from composable_kernel_b200 import ops, torch_ops  # noqa: F401
    o = orc()
    rng = np.random.default_rng(11)
    M, N, K = 5, 512, 1024
    a = rand_bf16(rng, (M, K))
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N)) * 0.015 + 0.005).astype(np.float32))
    y = torch.ops.ckb.w4a16_linear(H.to_dev(a, True), ops.prepack_int4(H.to_dev(q)), H.to_dev(s, True), None, N, K, False)
    ref = o.gemm(a, o.BF16, q, o.U8, o.BF16, b_op=o.BOP_GROUP_F32, scale=s, s_dtype=o.BF16, zero_const=8, group_size=128)
    bad, err = H.bf16_close(H.from_dev_bf16(y), ref, RTOL, ATOL)
    assert bad == 0, err
    ai = rng.integers(-128, 128, (64, 256), dtype=np.int8); bn = rng.integers(-128, 128, (128, 256), dtype=np.int8)
    d0 = (rng.random(128) * 1e-3).astype(np.float32); d1 = (rng.random(64) * 1e-3).astype(np.float32)
    e = torch.ops.ckb.int8_gemm_scaled(H.to_dev(ai), H.to_dev(bn), H.to_dev(d0), H.to_dev(d1))
    refe = o.gemm(ai, o.I8, bn.T, o.I8, o.BF16, cshuffle_dtype=o.I32, ds=(np.broadcast_to(d0, (64, 128)), np.broadcast_to(d1[:, None], (64, 128))),
                  d_dtypes=(o.F32, o.F32), cde_op=o.CDE_MULTIPLY_MULTIPLY)
    assert np.array_equal(H.from_dev_bf16(e), refe)


@pytest.mark.parametrize("M,N,K,cde", [(8, 256, 512, "add"), (16, 4096, 4096, "pass"), (33, 384, 640, "add_gelu"), (1, 11008, 4096, "add"),
                                       (16, 14336, 4096, "add"), (8, 14336, 1024, "add_gelu"), (3, 14336, 512, "pass")])  # cfg3's N: multi-pass slots
def test_w8a16_decode_both_kernels_vs_oracle(M, N, K, cde):
    """int8-weight decode rows on the N-split HMMA kernel (default) and on the stream-K tcgen05 kernel (CKB_FLAG_W8_STREAMK,
    the second factory instance): same oracle, same tolerance"""
    from composable_kernel_b200 import capi
    o = orc()
    rng = np.random.default_rng(M * 3 + N + K)
    a = rand_bf16(rng, (M, K))
    qn = rng.integers(-128, 128, (N, K), dtype=np.int8)
    s = o.bf16_trunc((rng.random(N) * 0.02 + 0.001).astype(np.float32))
    bias = rand_bf16(rng, (N,))
    op_id = {"pass": o.CDE_PASSTHROUGH, "add": o.CDE_ADD, "gelu": o.CDE_FASTGELU, "add_gelu": o.CDE_ADD_FASTGELU}[cde]
    ds = (np.broadcast_to(bias, (M, N)),) if "add" in cde else ()
    ref = o.gemm(a, o.BF16, qn.T, o.I8, o.BF16, b_op=o.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)), s_dtype=o.BF16,
                 ds=ds, d_dtypes=(o.BF16,) * len(ds), cde_op=op_id)
    for flags, tag in ((0, "wq_gemv2_w8mul"), (capi.FLAG_W8_STREAMK, "streamk")):
        got, op = H.run_w8a16_bf16(a, qn, s, bias, cde, flags=flags)
        assert tag in op.kernel_name, op.kernel_name
        bad, err = H.bf16_close(got, ref, RTOL, ATOL)
        assert bad == 0, f"{op.kernel_name}: max err {err}"


@pytest.mark.parametrize("M,N,K,f16", [(8, 256, 512, True), (16, 1000, 384, True), (4, 512, 1024, False), (40, 264, 256, False), (1, 16, 64, True)])
def test_int8_passthrough_weight_decode_vs_oracle(M, N, K, f16):
    """PassThrough<AType, int8_t> B functor at decode sizes (the split-K-reduce family's weight format) for both 16-bit A types"""
    import torch
    from composable_kernel_b200 import capi, ops
    o = orc()
    rng = np.random.default_rng(M + N + K)
    qn = rng.integers(-128, 128, (N, K + (-K) % 16), dtype=np.int8)[:, :K]
    if f16:
        a = (rng.standard_normal((M, K)) * 0.25).astype(np.float16)
        ad, e = H.to_dev(a), torch.zeros(M, N, dtype=torch.float16, device="cuda")
    else:
        a = rand_bf16(rng, (M, K), 0.25)
        ad, e = H.to_dev(a, True), torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    op = ops.GemmOp(ad, H.to_dev(np.ascontiguousarray(qn)), e, b_op=capi.BOP_PASSTHROUGH)
    assert op.supported, op.why
    op.run()
    torch.cuda.synchronize()
    dt = o.F16 if f16 else o.BF16
    ref = o.gemm(a, dt, qn.T, o.I8, dt, b_op=o.BOP_PASSTHROUGH)
    if f16:
        bad, err = H.f_close(e.cpu().numpy(), ref, RTOL, ATOL)
    else:
        bad, err = H.bf16_close(H.from_dev_bf16(e), ref, RTOL, ATOL)
    assert bad == 0, f"{op.kernel_name}: {err}"
