"""Own bounds checks (compute-sanitizer is closed on this GPU pool, profiles/r2_sanitizer_unavailable.txt): every kernel family
writes its E into a window of a larger sentinel-filled buffer (guard rows above and below, guard columns to the right, lde > N)
with ragged M / N, and reads A from a window as well.  After the run the sentinels must be intact and the window must match the
oracle -- an out-of-bounds store or a store past N in a row shows up as a damaged sentinel."""
import numpy as np
import pytest

from tests import helpers as H

pytestmark = pytest.mark.gpu
G = 3            # guard rows
PADC = 24        # guard columns (keeps lde a multiple of 8)


def orc():
    from oracle import ck_oracle
    return ck_oracle


def window(M, N, dtype):
    import torch
    big = torch.full((M + 2 * G, N + PADC), -777 if dtype in (torch.int8, torch.int32) else -777.0, dtype=dtype, device="cuda")
    return big, big[G:G + M, :N]


def guards_intact(big, M, N):
    import torch
    sent = big.new_full((), -777)
    top, bot, right = big[:G], big[G + M:], big[G:G + M, N:]
    return bool((top == sent).all() and (bot == sent).all() and (right == sent).all())


@pytest.mark.parametrize("M,N,K,rtn", [(5, 1000, 384, False), (1, 136, 128, False), (16, 520, 1024, True), (32, 2056, 512, False),
                                       (40, 264, 256, False), (130, 520, 512, True), (257, 1000, 384, True)])
def test_w4a16_kernels_stay_inside_their_window(M, N, K, rtn):
    import torch
    from composable_kernel_b200 import capi, ops
    o = orc()
    rng = np.random.default_rng(M + N)
    a = o.bf16_trunc(rng.standard_normal((M, K)).astype(np.float32))
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    s = o.bf16_trunc((rng.random((K // 128, N)) * 0.015 + 0.005).astype(np.float32))
    big_a, a_win = window(M, K, torch.bfloat16)
    a_win.copy_(H.to_dev(a, True))
    big, e = window(M, N, torch.bfloat16)
    op = ops.GemmOp(a_win, ops.prepack_int4(H.to_dev(q)), e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP_RTN if rtn else capi.BOP_GROUP,
                    N=N, K=K, scale=H.to_dev(s, True), zero_const=8, group_size=128)
    assert op.supported, op.why
    op.run()
    torch.cuda.synchronize()
    assert guards_intact(big, M, N), f"{op.kernel_name} wrote outside its E window"
    ref = o.gemm(a, o.BF16, q, o.U8, o.BF16, b_op=o.BOP_GROUP_RTN if rtn else o.BOP_GROUP_F32, scale=s, s_dtype=o.BF16, zero_const=8, group_size=128)
    bad, err = H.bf16_close(H.from_dev_bf16(e.contiguous()), ref)
    assert bad == 0, f"{op.kernel_name}: {err}"


@pytest.mark.parametrize("M,N,K", [(7, 200, 256), (64, 328, 128), (100, 520, 384), (300, 1000, 256)])
def test_w8a16_kernels_stay_inside_their_window(M, N, K):
    import torch
    from composable_kernel_b200 import capi, ops
    o = orc()
    rng = np.random.default_rng(M * N)
    a = o.bf16_trunc(rng.standard_normal((M, K)).astype(np.float32))
    qn = rng.integers(-128, 128, (N, K), dtype=np.int8)
    s = o.bf16_trunc((rng.random(N) * 0.02 + 0.001).astype(np.float32))
    bias = o.bf16_trunc(rng.standard_normal(N).astype(np.float32))
    big, e = window(M, N, torch.bfloat16)
    op = ops.GemmOp(H.to_dev(a, True), H.to_dev(qn), e, b_op=capi.BOP_MULTIPLY, scale=H.to_dev(s, True), ds=(H.to_dev(bias, True),),
                    d_strides=((0, 1),), cde_op=capi.CDE_ADD)
    assert op.supported, op.why
    op.run()
    op.run()
    torch.cuda.synchronize()
    assert guards_intact(big, M, N), f"{op.kernel_name} wrote outside its E window"
    ref = o.gemm(a, o.BF16, qn.T, o.I8, o.BF16, b_op=o.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)), s_dtype=o.BF16,
                 ds=(np.broadcast_to(bias, (M, N)),), d_dtypes=(o.BF16,), cde_op=o.CDE_ADD)
    bad, err = H.bf16_close(H.from_dev_bf16(e.contiguous()), ref)
    assert bad == 0, f"{op.kernel_name}: {err}"


@pytest.mark.parametrize("M,N,K,pair", [(100, 200, 144, "0"), (640, 520, 256, "1"), (129, 257, 128, "0")])
def test_int8_kernels_stay_inside_their_window(M, N, K, pair, monkeypatch):
    import torch
    from composable_kernel_b200 import ops
    o = orc()
    monkeypatch.setenv("CKB_SS_2CTA", pair)
    rng = np.random.default_rng(M + K)
    a = rng.integers(-128, 128, (M, K), dtype=np.int8)
    bn = rng.integers(-128, 128, (N, K), dtype=np.int8)
    big, e = window(M, N, torch.int32)
    op = ops.GemmOp(H.to_dev(a), H.to_dev(bn), e)
    assert op.supported, op.why
    op.run()
    torch.cuda.synchronize()
    assert guards_intact(big, M, N), f"{op.kernel_name} wrote outside its E window"
    ref = o.gemm(a, o.I8, bn.T, o.I8, o.I32, cshuffle_dtype=o.I32)
    assert np.array_equal(e.contiguous().cpu().numpy(), ref), op.kernel_name
