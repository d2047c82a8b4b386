#!/usr/bin/env python
"""Generate the committed golden vectors under tests/golden/ from the CPU oracle.

The reference keeps no stored vectors for this path and its host code cannot be built
here, so these fixtures are produced by the restated oracle (oracle/ck_oracle.cpp), which is itself
pinned by tests/test_oracle.py (reference KAT constants + an independent numpy evaluation).
Each .npz holds the inputs, the parameters and the expected E of one small GEMM.

    python oracle/make_golden.py          # rewrites tests/golden/*.npz deterministically
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ck_oracle as orc  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def bf16(x):
    return orc.bf16_trunc(np.asarray(x, dtype=np.float32))


def case_w4a16(name, M, N, K, seed, with_zero, f16=False):
    rng = np.random.default_rng(seed)
    G = 128
    q = rng.integers(0, 16, (K, N)).astype(np.uint8)
    if f16:
        a = rng.standard_normal((M, K)).astype(np.float16)
        s = (rng.random((K // G, N)) * 0.015 + 0.005).astype(np.float16)
        dt = orc.F16
    else:
        a = bf16(rng.standard_normal((M, K)))
        s = bf16(rng.random((K // G, N)) * 0.015 + 0.005)
        dt = orc.BF16
    z = rng.integers(0, 16, (K // G, N)).astype(np.uint8) if with_zero else np.zeros((0, 0), np.uint8)
    kw = dict(scale=s, s_dtype=dt, zero=z if with_zero else None, zero_const=8, group_size=G)
    # e / e_f32: the decode definition (exact (q - zp), group scale applied in fp32 -- CK's ab_scale host pattern);
    # e_rtn: (q - zp) * s rounded once into A's type first, what the prefill path feeds the dense tensor-core GEMM
    e     = orc.gemm(a, dt, q, orc.U8, dt, b_op=orc.BOP_GROUP_F32, **kw)
    e_f32 = orc.gemm(a, dt, q, orc.U8, orc.F32, b_op=orc.BOP_GROUP_F32, **kw)
    e_rtn = orc.gemm(a, dt, q, orc.U8, dt, b_op=orc.BOP_GROUP_RTN, **kw)
    np.savez_compressed(os.path.join(OUT, name), kind="w4a16", a=a, q=q, s=s, z=z, f16=f16, e=e, e_f32=e_f32, e_rtn=e_rtn)


def case_w8a16_bf16(name, M, N, K, seed, cde):
    rng = np.random.default_rng(seed)
    a = bf16(rng.standard_normal((M, K)))
    qn = rng.integers(-128, 128, (N, K)).astype(np.int8)  # [N,K] storage = column-major B (mk_nk_mn)
    s = bf16(rng.random(N) * 0.02 + 0.001)
    bias = bf16(rng.standard_normal(N))
    op = {"pass": orc.CDE_PASSTHROUGH, "add": orc.CDE_ADD, "gelu": orc.CDE_FASTGELU, "add_gelu": orc.CDE_ADD_FASTGELU}[cde]
    ds = (np.broadcast_to(bias, (M, N)),) if "add" in cde else ()
    e = orc.gemm(a, orc.BF16, qn.T, orc.I8, orc.BF16, b_op=orc.BOP_MULTIPLY, scale=np.broadcast_to(s, (K, N)),
                 s_dtype=orc.BF16, ds=ds, d_dtypes=(orc.BF16,) * len(ds), cde_op=op)
    np.savez_compressed(os.path.join(OUT, name), kind="w8a16_bf16", a=a, qn=qn, s=s, bias=bias, cde=cde, e=e)


def case_fpaintb(name, M, N, K, seed):
    # example/64_fpAintB_gemm: fp16 A row-major, int8 B column-major (+128 biased on device), fp16 scale[N]
    rng = np.random.default_rng(seed)
    a = rng.standard_normal((M, K)).astype(np.float16)
    qn = rng.integers(-128, 128, (N, K)).astype(np.int8)
    s = (rng.random(N) * 0.02 + 0.001).astype(np.float16)
    e = orc.gemm(a, orc.F16, qn.T, orc.I8, orc.F16, b_op=orc.BOP_FPAINTB, scale=np.broadcast_to(s, (K, N)), s_dtype=orc.F16)
    np.savez_compressed(os.path.join(OUT, name), kind="fpaintb", a=a, qn=qn, s=s, e=e)


def case_int8(name, M, N, K, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(-128, 128, (M, K)).astype(np.int8)
    bn = rng.integers(-128, 128, (N, K)).astype(np.int8)
    bias = rng.integers(-4000, 4000, N).astype(np.int32)
    d0 = (rng.random(N) * 1e-3 + 5e-4).astype(np.float32)
    d1 = (rng.random(M) * 1e-3 + 5e-4).astype(np.float32)
    acc = orc.gemm(a, orc.I8, bn.T, orc.I8, orc.I32, cshuffle_dtype=orc.I32)
    e_q = orc.gemm(a, orc.I8, bn.T, orc.I8, orc.I8, cshuffle_dtype=orc.I32, cde_op=orc.CDE_ACT_MUL_CLAMP, cde_scale=0.03)
    e_bq = orc.gemm(a, orc.I8, bn.T, orc.I8, orc.I8, cshuffle_dtype=orc.I32, ds=(np.broadcast_to(bias, (M, N)),),
                    d_dtypes=(orc.I32,), cde_op=orc.CDE_ADD_ACT_MUL_CLAMP, cde_scale=0.01, act=orc.ACT_RELU)
    e_mm = orc.gemm(a, orc.I8, bn.T, orc.I8, orc.BF16, cshuffle_dtype=orc.I32,
                    ds=(np.broadcast_to(d0, (M, N)), np.broadcast_to(d1[:, None], (M, N))),
                    d_dtypes=(orc.F32, orc.F32), cde_op=orc.CDE_MULTIPLY_MULTIPLY)
    np.savez_compressed(os.path.join(OUT, name), kind="int8", a=a, bn=bn, bias=bias, d0=d0, d1=d1, acc=acc, e_q=e_q,
                        e_bq=e_bq, e_mm=e_mm)


def main():
    os.makedirs(OUT, exist_ok=True)
    case_w4a16("w4a16_bf16_m4_n256_k256.npz", 4, 256, 256, 101, with_zero=False)
    case_w4a16("w4a16_bf16_m16_n384_k512_zp.npz", 16, 384, 512, 102, with_zero=True)
    case_w4a16("w4a16_f16_m1_n128_k384.npz", 1, 128, 384, 103, with_zero=True, f16=True)
    case_w8a16_bf16("w8a16_bf16_m8_n256_k256_add.npz", 8, 256, 256, 201, "add")
    case_w8a16_bf16("w8a16_bf16_m33_n128_k384_addgelu.npz", 33, 128, 384, 202, "add_gelu")
    case_fpaintb("fpaintb_f16_m16_n256_k256.npz", 16, 256, 256, 301)
    case_int8("int8_m96_n272_k208.npz", 96, 272, 208, 401)
    total = sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT))
    print("wrote", sorted(os.listdir(OUT)), total, "bytes")


if __name__ == "__main__":
    main()
