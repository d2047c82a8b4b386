"""torch.library custom ops over the C-ABI (the role of python/ck4inductor/universal_gemm/op.py in the reference:
make the CK GEMMs callable from PyTorch graphs).  Thin plumbing: each op allocates E, builds (or re-uses) the ckb plan for the
exact pointers and launches on torch's current stream.  Fake (meta) implementations give torch.compile / FakeTensor tracing the
output shape; there is no CPU implementation -- on a machine without an sm_100 GPU the ops raise.

    torch.ops.ckb.w4a16_linear(x, packed, scales, zeros, N, K, rtn)   x [M, K] bf16|fp16 -> [M, N]
    torch.ops.ckb.w8a16_linear(x, weight_nk, scales, bias)            bf16 x int8 [N, K], per-channel bf16 scale, optional bias
    torch.ops.ckb.int8_gemm_scaled(a, b_nk, col_scale, row_scale)     int8 x int8, int32 accumulate, MultiplyMultiply -> bf16
"""
from __future__ import annotations

from typing import Optional

import torch

from . import capi, ops

_PLANS: dict = {}


def _cached(key, make):
    op = _PLANS.get(key)
    if op is None:
        if len(_PLANS) > 4096:
            _PLANS.clear()
        op = _PLANS[key] = make()
        if not op.supported:
            del _PLANS[key]
            raise RuntimeError(f"ckb: unsupported problem: {op.why}")
    return op


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


@torch.library.custom_op("ckb::w4a16_linear", mutates_args=())
def w4a16_linear(x: torch.Tensor, packed: torch.Tensor, scales: torch.Tensor, zeros: Optional[torch.Tensor], N: int, K: int,
                 rtn: bool) -> torch.Tensor:
    x2 = x.reshape(-1, K).contiguous()
    e = torch.empty(x2.shape[0], N, dtype=x.dtype, device=x.device)
    if x2.shape[0] == 0:
        return e.reshape(*x.shape[:-1], N)
    key = ("w4", x2.data_ptr(), packed.data_ptr(), scales.data_ptr(), zeros.data_ptr() if zeros is not None else 0, e.data_ptr(),
           x2.shape[0], N, K, rtn, x.dtype)
    op = _cached(key, lambda: ops.GemmOp(x2, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED,
                                         b_op=capi.BOP_GROUP_RTN if rtn else capi.BOP_GROUP, N=N, K=K, scale=scales, zero=zeros,
                                         zero_const=8, group_size=128))
    op.run(_stream(x))
    return e.reshape(*x.shape[:-1], N)


@w4a16_linear.register_fake
def _(x, packed, scales, zeros, N, K, rtn):
    return x.new_empty(*x.shape[:-1], N)


@torch.library.custom_op("ckb::w8a16_linear", mutates_args=())
def w8a16_linear(x: torch.Tensor, weight_nk: torch.Tensor, scales: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    N, K = weight_nk.shape
    x2 = x.reshape(-1, K).contiguous()
    e = torch.empty(x2.shape[0], N, dtype=x.dtype, device=x.device)
    if x2.shape[0] == 0:
        return e.reshape(*x.shape[:-1], N)
    key = ("w8", x2.data_ptr(), weight_nk.data_ptr(), scales.data_ptr(), bias.data_ptr() if bias is not None else 0, e.data_ptr(), x2.shape[0], N, K)
    op = _cached(key, lambda: ops.GemmOp(x2, weight_nk, e, b_op=capi.BOP_MULTIPLY, scale=scales, ds=(bias,) if bias is not None else (),
                                         d_strides=((0, 1),) if bias is not None else (),
                                         cde_op=capi.CDE_ADD if bias is not None else capi.CDE_PASSTHROUGH))
    op.run(_stream(x))
    return e.reshape(*x.shape[:-1], N)


@w8a16_linear.register_fake
def _(x, weight_nk, scales, bias):
    return x.new_empty(*x.shape[:-1], weight_nk.shape[0])


@torch.library.custom_op("ckb::int8_gemm_scaled", mutates_args=())
def int8_gemm_scaled(a: torch.Tensor, b_nk: torch.Tensor, col_scale: torch.Tensor, row_scale: torch.Tensor) -> torch.Tensor:
    M, K = a.shape
    N = b_nk.shape[0]
    e = torch.empty(M, N, dtype=torch.bfloat16, device=a.device)
    if M == 0:
        return e
    key = ("i8", a.data_ptr(), b_nk.data_ptr(), col_scale.data_ptr(), row_scale.data_ptr(), e.data_ptr(), M, N, K)
    op = _cached(key, lambda: ops.GemmOp(a, b_nk, e, ds=(col_scale, row_scale), d_strides=((0, 1), (1, 0)), cde_op=capi.CDE_MULTIPLY_MULTIPLY))
    op.run(_stream(a))
    return e


@int8_gemm_scaled.register_fake
def _(a, b_nk, col_scale, row_scale):
    return a.new_empty(a.shape[0], b_nk.shape[0], dtype=torch.bfloat16)
