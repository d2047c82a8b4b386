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
                 act=capi.ACT_PASSTHROUGH, flags=0, kbatch=0, e_peers=None, flag_peers=None, peer_rank=0):
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
        if e_peers:
            # fused all-gather epilogue: raw peer-mapped pointers (this rank's slot in every rank's buffer, every rank's counter)
            p.num_peers, p.peer_rank = len(e_peers), peer_rank
            for g, (ep_, fp_) in enumerate(zip(e_peers, flag_peers)):
                p.e_peer[g], p.flag_peer[g] = int(ep_), int(fp_)
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


class GroupedGemmOp:
    """DeviceGroupedGemmMultiABDFixedNK life cycle (client_example/31): G GEMMs sharing N, K and the functors, the
    per-group pointers / M / strides in a DEVICE array of GroupedGemmMultiABDKernelArgument<1,NumB,NumD> structs
    (SetDeviceKernelArgs).  as_[i]: bf16 [M_i,K]; b0s[i]: int8, [K,N] (b_layout ROW) or [N,K] (COL); b1s[i]: bf16 [N]
    or None (PassThrough); d0s[i]: bf16 tensor + its row stride (0 = bias broadcast along M) or None; es[i]: bf16 [M_i,N]."""

    def __init__(self, as_, b0s, b1s, d0s, es, *, b_layout=capi.ROW, cde_op=capi.CDE_PASSTHROUGH, d0_stride=0, flags=0):
        import struct
        lib = capi.lib()
        G = len(as_)
        num_b = 2 if b1s is not None else 1
        num_d = 1 if d0s is not None else 0
        p = capi.GemmProblem()
        K = as_[0].shape[1]
        N = es[0].shape[1]
        p.M, p.N, p.K = sum(a.shape[0] for a in as_), N, K
        p.a_dtype, p.a_layout, p.lda = _dt(as_[0]), capi.ROW, as_[0].stride(0)
        p.b_dtype, p.b_layout, p.ldb = _dt(b0s[0]), b_layout, b0s[0].stride(0)
        p.b_op = capi.BOP_MULTIPLY if num_b == 2 else capi.BOP_PASSTHROUGH
        if num_b == 2:
            p.s_dtype = _dt(b1s[0])
        p.num_d = num_d
        if num_d:
            p.d_dtype[0] = _dt(d0s[0])
        p.cde_op, p.e_dtype, p.lde, p.flags = cde_op, _dt(es[0]), es[0].stride(0), flags
        self.problem, self.G, self.sum_m = p, G, p.M
        rc = lib.ckb_grouped_gemm_query(C.byref(p), G)
        self.supported = rc == 0
        self.why = capi.last_error() if rc else ""
        self._keep = (as_, b0s, b1s, d0s, es)
        # GroupedGemmMultiABDKernelArgument<1, num_b, num_d> as the C++ compiler lays it out: a zero-length std::array member
        # still occupies storage (one byte + padding), so NumD = 0 has the offsets of NumD = 1 with the D slots unused
        stride = int(lib.ckb_grouped_gemm_arg_bytes(num_b, num_d))
        blob = bytearray()
        for i in range(G):
            ptrs = [as_[i].data_ptr(), b0s[i].data_ptr()] + ([b1s[i].data_ptr()] if num_b == 2 else []) + \
                   [d0s[i].data_ptr() if num_d else 0] + [es[i].data_ptr()]
            ints = [as_[i].shape[0], N, K, as_[i].stride(0), b0s[i].stride(0)] + ([0] if num_b == 2 else []) + \
                   [d0_stride if num_d else 0] + [es[i].stride(0)]
            rec = struct.pack(f"<{len(ptrs)}Q{len(ints)}i", *ptrs, *ints)
            blob += rec + b"\0" * (stride - len(rec))
        self.dev_args = torch.frombuffer(blob, dtype=torch.uint8).cuda() if G else None

    def run(self, stream=None):
        if not self.supported:
            raise capi.CkbError(capi.ERR_UNSUPPORTED, self.why)
        s = stream if stream is not None else torch.cuda.current_stream().cuda_stream
        capi.check(capi.lib().ckb_grouped_gemm_run(C.byref(self.problem), self.dev_args.data_ptr(), self.G, self.sum_m, s))
