"""Small end-to-end pass over every kernel family for compute-sanitizer (tests/scripts/sanitize.sh): one modest problem per kernel,
each checked against the oracle so that a sanitizer-clean run is also a correct run."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from composable_kernel_b200 import capi  # noqa: E402
from oracle import ck_oracle as o  # noqa: E402
from tests import helpers as H  # noqa: E402

rng = np.random.default_rng(1)
bf = lambda x: o.bf16_trunc(x.astype(np.float32))
only = set(sys.argv[1:])


def want(name):
    return not only or name in only


def w4(M, N, K, rtn, zp):
    a = bf(rng.standard_normal((M, K)))
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    s = bf(rng.random((K // 128, N)) * 0.015 + 0.005)
    z = rng.integers(0, 16, (K // 128, N), dtype=np.uint8) if zp else None
    got, op = H.run_w4a16(a, q, s, z, out_f32=True, rtn=rtn)
    ref = o.gemm(a, o.BF16, q, o.U8, o.F32, b_op=o.BOP_GROUP_RTN if rtn else o.BOP_GROUP_F32, scale=s, s_dtype=o.BF16, zero=z, zero_const=8, group_size=128)
    bad, err = H.f_close(got, ref, 1e-4, 1e-3)
    print(f"w4 M{M} N{N} K{K} rtn={rtn} zp={zp}: {op.kernel_name} bad={bad} err={err:.2e}", flush=True)
    return bad


bad = 0
if want("gemv2"):
    bad += w4(4, 512, 1024, False, True)       # wq_gemv2 (rotating producer), zero points
    bad += w4(16, 1024, 512, True, False)      # wq_gemv2 RTN instantiation, two token tiles
if want("gemv"):
    os.environ["CKB_GEMV"] = "1"
    bad += w4(8, 512, 1024, False, False)      # single-producer kernel
    del os.environ["CKB_GEMV"]
if want("fused"):
    bad += w4(256, 512, 512, True, True)       # fused TMA -> dequant -> tcgen05 pair kernel
if want("w8"):
    M, N, K = 8, 256, 512                      # stream-K TMEM-A decode kernel (W8A16 Multiply)
    a = bf(rng.standard_normal((M, K))); qn = rng.integers(-128, 128, (N, K), dtype=np.int8); s = bf(rng.random(N) * 0.02 + 0.001)
    bias = bf(rng.standard_normal(N))
    got, op = H.run_w8a16_bf16(a, qn, s, bias, "add")
    ref = o.gemm(a, o.BF16, qn.T, o.I8, o.BF16, b_op=o.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)), s_dtype=o.BF16,
                 ds=(np.broadcast_to(bias, (M, N)),), d_dtypes=(o.BF16,), cde_op=o.CDE_ADD)
    b_, err = H.bf16_close(got, ref)
    print(f"w8 decode: {op.kernel_name} bad={b_} err={err:.2e}", flush=True)
    bad += b_
    M = 192                                    # fused prefill, W8 Multiply + bias
    a = bf(rng.standard_normal((M, K)))
    got, op = H.run_w8a16_bf16(a, qn, s, bias, "add")
    ref = o.gemm(a, o.BF16, qn.T, o.I8, o.BF16, b_op=o.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)), s_dtype=o.BF16,
                 ds=(np.broadcast_to(bias, (M, N)),), d_dtypes=(o.BF16,), cde_op=o.CDE_ADD)
    b_, err = H.bf16_close(got, ref)
    print(f"w8 prefill: {op.kernel_name} bad={b_} err={err:.2e}", flush=True)
    bad += b_
if want("i8"):
    M, N, K = 256, 512, 256                    # gemm_ss (1-CTA) and the CTA-pair kernel
    a = rng.integers(-128, 128, (M, K), dtype=np.int8); bn = rng.integers(-128, 128, (N, K), dtype=np.int8)
    ref = o.gemm(a, o.I8, bn.T, o.I8, o.I32, cshuffle_dtype=o.I32)
    for pair in ("0", "1"):
        os.environ["CKB_SS_2CTA"] = pair
        got, op = H.run_int8(a, bn, "i32")
        ok = np.array_equal(got, ref)
        print(f"int8 2cta={pair}: {op.kernel_name} exact={ok}", flush=True)
        bad += 0 if ok else 1
    del os.environ["CKB_SS_2CTA"]
torch.cuda.synchronize()
print("sanitize driver:", "OK" if bad == 0 else f"{bad} mismatches")
sys.exit(1 if bad else 0)
