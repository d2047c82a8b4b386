"""Probe: host-side cost per call of ckb_gemm_query / ckb_gemm_run / ckb_gemm_run_host_async (us)."""
import ctypes as C
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from composable_kernel_b200 import capi, ops  # noqa: E402

M, N, K = 4, 4096, 4096
a = torch.randn(M, K, device="cuda").bfloat16()
q = torch.randint(0, 16, (K, N), dtype=torch.uint8, device="cuda")
sc = torch.rand(K // 128, N, device="cuda").bfloat16()
e = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
op = ops.GemmOp(a, ops.prepack_int4(q), e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=N, K=K, scale=sc,
                zero_const=8, group_size=128)
lib = capi.lib()
s = torch.cuda.current_stream().cuda_stream
n = 500


def t(f):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    host = (time.perf_counter() - t0) / n * 1e6
    torch.cuda.synchronize()
    return host, (time.perf_counter() - t0) / n * 1e6


ha = a.cpu().pin_memory()
he = torch.zeros(M, N, dtype=torch.bfloat16).pin_memory()
p0 = capi.GemmProblem()
C.memmove(C.byref(p0), C.byref(op.problem), C.sizeof(p0))
p0.a, p0.e = ha.data_ptr(), he.data_ptr()
st0 = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
plan = capi.GemmPlan()
print("query            host %.1f us" % t(lambda: lib.ckb_gemm_query(C.byref(op.problem), C.byref(plan)))[0])
for pdl in ("1", "0"):
    os.environ["CKB_GEMV_PDL"] = pdl
    print("run (PDL=%s)      host %.1f us, with drain %.1f us" % ((pdl,) + t(lambda: lib.ckb_gemm_run(C.byref(op.problem), C.byref(op.plan), None, s))))
    print("run_host_async (PDL=%s) host %.1f us, with drain %.1f us" % ((pdl,) + t(lambda: lib.ckb_gemm_run_host_async(C.byref(p0), st0.data_ptr(), None, s))))
ha = a.cpu().pin_memory()
he = torch.zeros(M, N, dtype=torch.bfloat16).pin_memory()
p = capi.GemmProblem()
C.memmove(C.byref(p), C.byref(op.problem), C.sizeof(p))
p.a, p.e = ha.data_ptr(), he.data_ptr()
st = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
print("run_host_async   host %.1f us, with drain %.1f us" % t(lambda: lib.ckb_gemm_run_host_async(C.byref(p), st.data_ptr(), None, s)))
print("h2d 32 KB        host %.1f us, with drain %.1f us" % t(lambda: a.copy_(ha, non_blocking=True)))
print("d2h 32 KB        host %.1f us, with drain %.1f us" % t(lambda: he.copy_(e, non_blocking=True)))
