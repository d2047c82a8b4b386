"""Thin torch-tensor front end over the C-ABI (plumbing for tests and bench.py).

`GemmOp` mirrors the CK device-operator life cycle on one problem:
    MakeArgument  -> GemmOp(...)            (ckb_gemm_query: validates, picks the kernel, encodes TMA maps)
    IsSupportedArgument -> op.supported
    GetWorkSpaceSize / SetWorkSpacePointer -> op.workspace (zeroed once)
    Invoker::Run  -> op.run(stream)         (ckb_gemm_run)
The compute always happens in libckb.so; nothing here falls back to torch.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import capi

_TORCH_DT = {
    torch.float32: capi.F32, torch.float16: capi.F16, torch.bfloat16: capi.BF16,
    torch.int8: capi.I8, torch.uint8: capi.U8, torch.int32: capi.I32,
}


def _dt(t: torch.Tensor) -> int:
    return _TORCH_DT[t.dtype]


class GemmOp:
    def __init__(self, a, b, e, *, b_dtype=None, b_layout=capi.COL, b_op=capi.BOP_PASSTHROUGH,
                 N=None, K=None, ldb=None, scale=None, zero=None, zero_const=8, group_size=0,
                 ds=(), d_strides=(), cde_op=capi.CDE_PASSTHROUGH, cde_scale=1.0,
                 act=capi.ACT_PASSTHROUGH, flags=0, kbatch=0):
        """a: [M,K] row-major (stride(1)==1).  b: weight storage tensor; for COL layout a [N,K] tensor
        (K contiguous), for PREPACKED a flat uint8 tensor from prepack_int4 (then N,K are required).
        e: [M,N] row-major output.  ds: D tensors with explicit (stride_m, stride_n) element strides."""
        lib = capi.lib()
        p = capi.GemmProblem()
        M = a.shape[0]
        if K is None:
            K = a.shape[1]
        if N is None:
            N = b.shape[0] if b_layout == capi.COL else b.shape[1]
        p.M, p.N, p.K = M, N, K
        p.a_dtype, p.a_layout, p.a, p.lda = _dt(a), capi.ROW, a.data_ptr(), a.stride(0) if M > 0 else K
        p.b_dtype = b_dtype if b_dtype is not None else _dt(b)
        p.b_layout, p.b = b_layout, b.data_ptr()
        p.ldb = ldb if ldb is not None else (b.stride(0) if b.dim() == 2 else 0)
        p.b_op = b_op
        if scale is not None:
            p.s_dtype, p.scale = _dt(scale), scale.data_ptr()
        if zero is not None:
            assert zero.dtype == torch.uint8
            p.zero = zero.data_ptr()
        p.zero_const, p.group_size = zero_const, group_size
        p.num_d = len(ds)
        for i, d in enumerate(ds):
            p.d_dtype[i], p.d[i] = _dt(d), d.data_ptr()
            p.d_sm[i], p.d_sn[i] = d_strides[i]
        p.cde_op, p.cde_scale, p.act = cde_op, cde_scale, act
        p.e_dtype, p.e, p.lde = _dt(e), e.data_ptr(), e.stride(0) if M > 0 else N
        p.kbatch, p.flags = kbatch, flags
        self.problem = p
        self.plan = capi.GemmPlan()
        self._keep = (a, b, e, scale, zero, ds)
        capi.check(lib.ckb_gemm_query(C.byref(p), C.byref(self.plan)))
        self.supported = bool(self.plan.supported)
        self.why = capi.last_error() if not self.supported else ""
        self.workspace = None
        if self.supported and self.plan.workspace_bytes:
            self.workspace = torch.zeros(int(self.plan.workspace_bytes), dtype=torch.uint8, device=a.device)

    @property
    def kernel_name(self) -> str:
        return self.plan.kernel_name.decode()

    def run(self, stream=None):
        if not self.supported:
            raise capi.CkbError(capi.ERR_UNSUPPORTED, self.why)
        s = stream if stream is not None else torch.cuda.current_stream().cuda_stream
        ws = self.workspace.data_ptr() if self.workspace is not None else None
        capi.check(capi.lib().ckb_gemm_run(C.byref(self.problem), C.byref(self.plan), ws, s))


def prepack_int4(codes_kn: torch.Tensor) -> torch.Tensor:
    """codes_kn: uint8 [K,N] (any strides) of 4-bit codes -> tile-major packed uint8 buffer."""
    K, N = codes_kn.shape
    nbytes = capi.lib().ckb_prepack_b_int4_bytes(N, K)
    out = torch.empty(int(nbytes), dtype=torch.uint8, device=codes_kn.device)
    capi.check(capi.lib().ckb_prepack_b_int4(codes_kn.data_ptr(), codes_kn.stride(0), codes_kn.stride(1), N, K,
                                             out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return out


def prepack_u8(w_i8: torch.Tensor) -> torch.Tensor:
    """int8 -> biased uint8 (UnsignedWeightPreprocessor); contiguous storage is converted as-is."""
    out = torch.empty_like(w_i8, dtype=torch.uint8)
    assert w_i8.is_contiguous() or w_i8.t().is_contiguous()
    capi.check(capi.lib().ckb_prepack_b_u8(w_i8.data_ptr(), out.data_ptr(), w_i8.numel(),
                                           torch.cuda.current_stream().cuda_stream))
    return out
