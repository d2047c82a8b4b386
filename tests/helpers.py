"""Shared helpers for the GPU parity tests: numpy inputs -> device -> C-ABI GEMM -> numpy output.
Everything on the device side goes through composable_kernel_b200.ops.GemmOp, i.e. ckb_gemm_query /
ckb_gemm_run; the oracle (oracle/ck_oracle.py) is only ever the checker."""
import numpy as np


def to_dev(x, bf16=False):
    import torch
    if bf16:
        return torch.from_numpy(np.ascontiguousarray(x).view(np.int16)).cuda().view(torch.bfloat16)
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def from_dev_bf16(t):
    import torch
    return t.view(torch.int16).cpu().numpy().view(np.uint16)


def run_w4a16(a, q_kn, s, z=None, zero_const=8, f16=False, out_f32=False, ds=(), d_strides=(), cde_op=0, rtn=False, flags=0):
    """a [M,K] (uint16 bf16 bits or float16), q_kn uint8 [K,N] codes, s [K/128,N], z uint8 [K/128,N] or None."""
    import torch
    from composable_kernel_b200 import capi, ops
    M, K = a.shape
    N = q_kn.shape[1]
    ad, sd = to_dev(a, not f16), to_dev(s, not f16)
    packed = ops.prepack_int4(to_dev(q_kn))
    zd = to_dev(z) if z is not None else None
    e = torch.zeros(M, N, dtype=torch.float32 if out_f32 else (torch.float16 if f16 else torch.bfloat16), device="cuda")
    # rtn selects the numerics contract (capi.BOP_GROUP_RTN: one rounding into A's type; default capi.BOP_GROUP: fp32 group scale)
    op = ops.GemmOp(ad, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP_RTN if rtn else capi.BOP_GROUP, N=N, K=K,
                    scale=sd, zero=zd, zero_const=zero_const, group_size=128, ds=ds, d_strides=d_strides, cde_op=cde_op, flags=flags)
    assert op.supported, op.why
    op.run()
    op.run()  # second run: the self-cleaning stream-K workspace must have been restored
    torch.cuda.synchronize()
    if out_f32:
        return e.cpu().numpy(), op
    return (e.cpu().numpy() if f16 else from_dev_bf16(e)), op


def run_w8a16_bf16(a, qn, s, bias, cde, b_op="mul", flags=0):
    """a bf16 bits [M,K]; qn int8 [N,K] (column-major B); s, bias bf16 bits [N]."""
    import torch
    from composable_kernel_b200 import capi, ops
    M, K = a.shape
    N = qn.shape[0]
    e = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    cde_op = {"pass": capi.CDE_PASSTHROUGH, "add": capi.CDE_ADD, "gelu": capi.CDE_FASTGELU,
              "add_gelu": capi.CDE_ADD_FASTGELU}[cde]
    ds = (to_dev(bias, True),) if "add" in cde else ()
    op = ops.GemmOp(to_dev(a, True), to_dev(qn), e, b_op=capi.BOP_MULTIPLY if b_op == "mul" else capi.BOP_PASSTHROUGH,
                    scale=to_dev(s, True) if b_op == "mul" else None, ds=ds, d_strides=((0, 1),) * len(ds), cde_op=cde_op, flags=flags)
    assert op.supported, op.why
    op.run()
    torch.cuda.synchronize()
    return from_dev_bf16(e), op


def run_fpaintb(a, qn, s):
    import torch
    from composable_kernel_b200 import capi, ops
    M, K = a.shape
    N = qn.shape[0]
    e = torch.zeros(M, N, dtype=torch.float16, device="cuda")
    bu8 = ops.prepack_u8(to_dev(qn))  # UnsignedWeightPreprocessor
    op = ops.GemmOp(to_dev(a), bu8, e, b_op=capi.BOP_FPAINTB, scale=to_dev(s))
    assert op.supported, op.why
    op.run()
    torch.cuda.synchronize()
    return e.cpu().numpy(), op


def run_int8(a, bn, e_dtype, cde_op=0, ds=(), d_strides=(), cde_scale=1.0, act=0):
    import torch
    from composable_kernel_b200 import ops
    M, N = a.shape[0], bn.shape[0]
    tdt = {"i32": torch.int32, "i8": torch.int8, "bf16": torch.bfloat16}[e_dtype]
    e = torch.zeros(M, N, dtype=tdt, device="cuda")
    op = ops.GemmOp(to_dev(a), to_dev(bn), e, ds=tuple(to_dev(d) for d in ds), d_strides=d_strides, cde_op=cde_op,
                    cde_scale=cde_scale, act=act)
    assert op.supported, op.why
    op.run()
    torch.cuda.synchronize()
    return (from_dev_bf16(e) if e_dtype == "bf16" else e.cpu().numpy()), op


def bf16_close(got_bits, ref_bits, rtol=1e-2, atol=1e-2):
    """check_err criterion (library/include/ck/library/utility/check_err.hpp:54) on bf16 bit patterns."""
    from oracle import ck_oracle as orc
    g, r = orc.bf16_to_f32(got_bits).astype(np.float64), orc.bf16_to_f32(ref_bits).astype(np.float64)
    err = np.abs(g - r)
    bad = ~(err <= atol + rtol * np.abs(r)) | ~np.isfinite(g)
    return int(bad.sum()), float(err.max())


def f_close(got, ref, rtol=1e-2, atol=1e-2):
    g, r = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    err = np.abs(g - r)
    bad = ~(err <= atol + rtol * np.abs(r)) | ~np.isfinite(g)
    return int(bad.sum()), float(err.max())
