"""GPU parity at the FULL sizes of BASELINE.json's five configurations (pytest -m gpu).  The device computes the whole
problem through the C-ABI; the oracle evaluates a sample of rows (`m_range`), which costs seconds instead of minutes,
and the sampled rows are compared: bit-exact for int8, the 1e-2 check_err criterion for the 16-bit outputs (same
tolerances as tests/test_gpu_parity.py).  Plus the checks that pin the two W4A16 numerics contracts at every M
(no switch of definition at the decode / prefill boundary) and the shard-invariant K order on one device."""
import numpy as np
import pytest

from tests import helpers as H

pytestmark = pytest.mark.gpu
RTOL = ATOL = 1e-2


def orc():
    from oracle import ck_oracle
    return ck_oracle


def rand_bf16(rng, shape, scale=1.0):
    return orc().bf16_trunc((rng.standard_normal(shape, dtype=np.float32) * scale))


def row_blocks(M, rows=8):
    """first, middle and last `rows` rows of M (clipped, deduplicated)"""
    out = []
    for r0 in (0, (M // 2) // rows * rows, M - rows):
        r0 = max(0, min(r0, M - rows))
        blk = (r0, min(M, r0 + rows))
        if blk not in out:
            out.append(blk)
    return out


# ---------------------------------------------------------------- cfg1: example 64, fp16 A x int8 B, per-channel fp16 scale
def test_cfg1_fpaintb_3840x4096x4096():
    o = orc()
    rng = np.random.default_rng(641)
    M, N, K = 3840, 4096, 4096
    a = rng.standard_normal((M, K), dtype=np.float32).astype(np.float16)
    qn = rng.integers(-128, 128, (N, K), dtype=np.int8)
    s = (rng.random(N) * 0.02 + 0.001).astype(np.float16)
    got, op = H.run_fpaintb(a, qn, s)
    assert "fpaintb" in op.kernel_name
    for r0, r1 in row_blocks(M):
        ref = o.gemm(a, o.F16, qn.T, o.I8, o.F16, b_op=o.BOP_FPAINTB, scale=np.broadcast_to(s, (K, N)), s_dtype=o.F16, m_range=(r0, r1))
        bad, err = H.f_close(got[r0:r1], ref[r0:r1], RTOL, ATOL)
        assert bad == 0, f"rows {r0}:{r1} max err {err}"


# ---------------------------------------------------------------- cfg2: Llama-2-7B decode W4A16 g128, all nine (M, shape) cells
@pytest.mark.parametrize("N,K", [(4096, 4096), (11008, 4096), (4096, 11008)])
@pytest.mark.parametrize("M", [1, 4, 16])
def test_cfg2_w4a16_decode(M, N, K):
    o = orc()
    rng = np.random.default_rng(M * 7 + N + K)
    a = rand_bf16(rng, (M, K))
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N), dtype=np.float32) * 0.015 + 0.005))
    got, op = H.run_w4a16(a, q, s, out_f32=True)
    ref = o.gemm(a, o.BF16, q, o.U8, o.F32, b_op=o.BOP_GROUP_F32, scale=s, s_dtype=o.BF16, zero_const=8, group_size=128)
    bad, err = H.f_close(got, ref, 1e-4, 2e-4 * np.sqrt(K / 128))
    assert bad == 0, f"{op.kernel_name}: max err {err}"


# ---------------------------------------------------------------- cfg3: Llama-3-8B prefill W8A16 + bias, both B layouts
@pytest.mark.parametrize("b_layout", ["mk_nk_mn", "mk_kn_mn"])
def test_cfg3_w8a16_bias_4096x14336x4096(b_layout):
    import torch
    from composable_kernel_b200 import capi, ops
    o = orc()
    rng = np.random.default_rng(30)
    M, N, K = 4096, 14336, 4096
    a = rand_bf16(rng, (M, K))
    s = o.bf16_trunc((rng.random(N, dtype=np.float32) * 0.02 + 0.001))
    bias = rand_bf16(rng, (N,))
    e = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    if b_layout == "mk_nk_mn":
        qn = rng.integers(-128, 128, (N, K), dtype=np.int8)          # K-major: B[k,n] = qn[n,k]
        b_kn = qn.T
        bd = H.to_dev(qn)
    else:
        q_kn = rng.integers(-128, 128, (K, N), dtype=np.int8)        # row-major B (client 30): one-time K-major copy
        b_kn = q_kn
        src = H.to_dev(q_kn)
        bd = torch.empty(N, K, dtype=torch.int8, device="cuda")
        capi.check(capi.lib().ckb_prepack_b_rowmajor_i8(src.data_ptr(), N, K, N, bd.data_ptr(), None))
    op = ops.GemmOp(H.to_dev(a, True), bd, e, b_op=capi.BOP_MULTIPLY, scale=H.to_dev(s, True), ds=(H.to_dev(bias, True),),
                    d_strides=((0, 1),), cde_op=capi.CDE_ADD)
    assert op.supported, op.why
    op.run()
    torch.cuda.synchronize()
    got = H.from_dev_bf16(e)
    for r0, r1 in row_blocks(M):
        ref = o.gemm(a, o.BF16, b_kn, o.I8, o.BF16, b_op=o.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)), s_dtype=o.BF16,
                     ds=(np.broadcast_to(bias, (M, N)),), d_dtypes=(o.BF16,), cde_op=o.CDE_ADD, m_range=(r0, r1))
        bad, err = H.bf16_close(got[r0:r1], ref[r0:r1], RTOL, ATOL)
        assert bad == 0, f"{op.kernel_name} rows {r0}:{r1} max err {err}"


# ---------------------------------------------------------------- cfg4: int8 x int8, M = N = K = 8192, bit exact
def test_cfg4_int8_8192_cubed_bit_exact():
    from composable_kernel_b200 import capi
    o = orc()
    M = N = K = 8192
    rng = np.random.default_rng(65)
    a = rng.integers(-128, 128, (M, K), dtype=np.int8)
    bn = rng.integers(-128, 128, (N, K), dtype=np.int8)
    d0 = (rng.random(N, dtype=np.float32) * 1e-3 + 5e-4)
    d1 = (rng.random(M, dtype=np.float32) * 1e-3 + 5e-4)
    acc, op = H.run_int8(a, bn, "i32")
    e, _ = H.run_int8(a, bn, "bf16", cde_op=capi.CDE_MULTIPLY_MULTIPLY, ds=(d0, d1), d_strides=((0, 1), (1, 0)))
    for r0, r1 in row_blocks(M, 16):
        ref = o.gemm(a, o.I8, bn.T, o.I8, o.I32, cshuffle_dtype=o.I32, m_range=(r0, r1))
        assert np.array_equal(acc[r0:r1], ref[r0:r1]), f"{op.kernel_name}: int32 accumulators differ in rows {r0}:{r1}"
        refe = o.gemm(a, o.I8, bn.T, o.I8, o.BF16, cshuffle_dtype=o.I32, m_range=(r0, r1),
                      ds=(np.broadcast_to(d0, (M, N)), np.broadcast_to(d1[:, None], (M, N))), d_dtypes=(o.F32, o.F32),
                      cde_op=o.CDE_MULTIPLY_MULTIPLY)
        assert np.array_equal(e[r0:r1], refe[r0:r1]), f"MultiplyMultiply outputs differ in rows {r0}:{r1}"


# ---------------------------------------------------------------- cfg5: Llama-3-70B W4A16 MLP GEMM, N = 28672, K = 8192
@pytest.fixture(scope="module")
def cfg5_weights():
    o = orc()
    rng = np.random.default_rng(70)
    N, K = 28672, 8192
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N), dtype=np.float32) * 0.015 + 0.005))
    z = rng.integers(0, 16, (K // 128, N), dtype=np.uint8)
    return q, s, z


@pytest.mark.parametrize("M,zp", [(1, False), (1, True), (16, False), (16, True), (4096, False), (4096, True)])
def test_cfg5_w4a16_28672x8192(cfg5_weights, M, zp):
    o = orc()
    q, s, z = cfg5_weights
    K, N = q.shape
    rng = np.random.default_rng(M + int(zp))
    a = rand_bf16(rng, (M, K))
    zz = z if zp else None
    if M <= 64:
        # decode rows: the ab_scale contract (fp32 group scale), all rows against the oracle
        got, op = H.run_w4a16(a, q, s, zz, out_f32=True)
        ref = o.gemm(a, o.BF16, q, o.U8, o.F32, b_op=o.BOP_GROUP_F32, scale=s, s_dtype=o.BF16, zero=zz, zero_const=8, group_size=128)
        bad, err = H.f_close(got, ref, 1e-4, 2e-4 * np.sqrt(K / 128))
        assert bad == 0, f"{op.kernel_name}: max err {err}"
    else:
        # prefill rows: the one-rounding contract on the fused tcgen05 kernel, sampled rows
        got, op = H.run_w4a16(a, q, s, zz, rtn=True)
        assert "prefill_fused" in op.kernel_name
        for r0, r1 in row_blocks(M):
            ref = o.gemm(a, o.BF16, q, o.U8, o.BF16, b_op=o.BOP_GROUP_RTN, scale=s, s_dtype=o.BF16, zero=zz, zero_const=8, group_size=128,
                         m_range=(r0, r1))
            bad, err = H.bf16_close(got[r0:r1], ref[r0:r1], RTOL, ATOL)
            assert bad == 0, f"rows {r0}:{r1} max err {err}"


# ---------------------------------------------------------------- the two W4A16 contracts hold at EVERY M
@pytest.mark.parametrize("M", [1, 16, 33, 64, 65, 200])
@pytest.mark.parametrize("rtn", [False, True])
def test_w4a16_contract_is_independent_of_m(M, rtn):
    """CKB_BOP_GROUP (fp32 group scale on partial sums) and CKB_BOP_GROUP_RTN (one rounding into A's type) against their own
    oracle modes on both sides of the 64-row kernel boundary; fp32 outputs, so only accumulation-order noise is allowed."""
    o = orc()
    rng = np.random.default_rng(M + 1000 * int(rtn))
    N, K = 1024, 1024
    a = rand_bf16(rng, (M, K))
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N), dtype=np.float32) * 0.015 + 0.005))
    z = rng.integers(0, 16, (K // 128, N), dtype=np.uint8)
    got, op = H.run_w4a16(a, q, s, z, out_f32=True, rtn=rtn)
    if rtn:
        # oracle RTN mode with a float output = sum_k a_k * round_bf16((q - z) s) accumulated in fp32
        ref = o.gemm(a, o.BF16, q, o.U8, o.F32, b_op=o.BOP_GROUP_RTN, scale=s, s_dtype=o.BF16, zero=z, group_size=128)
        assert ("prefill_fused" in op.kernel_name) == (M > 64) and (M > 64 or "rtn" in op.kernel_name), op.kernel_name
    else:
        ref = o.gemm(a, o.BF16, q, o.U8, o.F32, b_op=o.BOP_GROUP_F32, scale=s, s_dtype=o.BF16, zero=z, group_size=128)
        assert "prefill" not in op.kernel_name, op.kernel_name
    bad, err = H.f_close(got, ref, 1e-4, 2e-4 * np.sqrt(K / 128))
    assert bad == 0, f"{op.kernel_name}: max err {err}"


@pytest.mark.parametrize("rtn", [False, True])
def test_w4a16_rows_do_not_change_across_the_64_row_boundary(rtn):
    """the same op must not change numerics between M = 64 and M = 65.  Rows 0..63 of an M = 65 problem against
    the M = 64 problem on the same inputs: identical definition, so fp32 outputs agree to accumulation-order noise (the
    two contracts differ from EACH OTHER by ~1e-3 relative, two orders of magnitude above the bar used here)."""
    o = orc()
    rng = np.random.default_rng(6465)
    N, K = 2048, 2048
    a = rand_bf16(rng, (65, K))
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N), dtype=np.float32) * 0.015 + 0.005))
    e64, op64 = H.run_w4a16(a[:64], q, s, out_f32=True, rtn=rtn)
    e65, op65 = H.run_w4a16(a, q, s, out_f32=True, rtn=rtn)
    scale = float(np.abs(e64).max())
    err = float(np.abs(e65[:64] - e64).max())
    assert err <= 2e-5 * scale + 1e-5, f"{op64.kernel_name} vs {op65.kernel_name}: {err} (scale {scale})"
    other, _ = H.run_w4a16(a[:64], q, s, out_f32=True, rtn=not rtn)
    assert float(np.abs(other - e64).max()) > 10 * err   # the test can tell the two contracts apart


# ---------------------------------------------------------------- shard-invariant K order (CKB_FLAG_FIXED_K_ORDER), one device
@pytest.mark.parametrize("M,N,K,parts", [(1, 28672, 8192, 8), (16, 28672, 8192, 4), (4, 4096, 4096, 2), (8, 11008, 4096, 2)])
def test_fixed_k_order_makes_column_shards_bit_identical(M, N, K, parts):
    """the concatenation of G column shards must equal the single-GPU E bit for bit.  With the flag the fp32
    summation order over K does not depend on N, so the slices computed as separate problems reproduce the full problem."""
    import torch
    from composable_kernel_b200 import capi, ops
    rng = np.random.default_rng(N + parts)
    a = H.to_dev(rand_bf16(rng, (M, K)), True)
    packed = torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device="cuda")
    sc = (torch.rand(K // 128, N, device="cuda") * 0.015 + 0.005).bfloat16()

    def run(p, s_, n):
        e = torch.zeros(M, n, dtype=torch.float32, device="cuda")
        op = ops.GemmOp(a, p, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=n, K=K, scale=s_, zero_const=8,
                        group_size=128, flags=capi.FLAG_FIXED_K_ORDER)
        assert op.supported and "korder" in op.kernel_name, (op.why, op.kernel_name)
        op.run()
        torch.cuda.synchronize()
        return e
    full = run(packed, sc, N)
    n_loc = N // parts
    if n_loc % 16:
        pytest.skip("shards must hold whole 16-channel groups")
    per_group = (K // 64) * 512
    for g in range(parts):
        p = packed[(g * n_loc // 16) * per_group:((g + 1) * n_loc // 16) * per_group].clone()
        e = run(p, sc[:, g * n_loc:(g + 1) * n_loc].contiguous(), n_loc)
        assert torch.equal(e.view(torch.int32), full[:, g * n_loc:(g + 1) * n_loc].contiguous().view(torch.int32)), f"shard {g} differs"
