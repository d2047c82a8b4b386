"""ckb_gemm_run_static_scratch (include/ckb.h): the allocation-free variant of ckb_gemm_run for the stream-K decode kernel, whose
scratch lives in a static device array inside libckb.so.  Called through the C-ABI and compared with the existing entry point
(ckb_gemm_run with a caller workspace) on two calls in a row -- the second call finds the first one's leftovers in the shared array.
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from tests import helpers as H

RTOL = ATOL = 1e-2   # tests/test_gpu_parity.py: check_err's bf16 tolerance (the stream-K partial sums meet in arrival order)


def _problem(M, N, K, seed, cde="add"):
    import torch
    from composable_kernel_b200 import capi, ops
    from oracle import ck_oracle as o
    rng = np.random.default_rng(seed)
    a = o.bf16_trunc((rng.standard_normal((M, K)) * 0.5).astype(np.float32))
    qn = rng.integers(-128, 128, (N, K), dtype=np.int8)
    s = o.bf16_trunc((rng.random(N) * 0.02 + 0.001).astype(np.float32))
    bias = o.bf16_trunc(rng.standard_normal(N).astype(np.float32))
    e = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    ds = (H.to_dev(bias, True),) if cde == "add" else ()
    op = ops.GemmOp(H.to_dev(a, True), H.to_dev(qn), e, b_op=capi.BOP_MULTIPLY, scale=H.to_dev(s, True), ds=ds, d_strides=((0, 1),) * len(ds),
                    cde_op=capi.CDE_ADD if cde == "add" else capi.CDE_PASSTHROUGH, flags=capi.FLAG_W8_STREAMK)
    assert op.supported, op.why
    assert "wq_decode" in op.kernel_name and op.plan.workspace_bytes > 0, op.kernel_name   # the kernel family that uses scratch memory
    return op, e


def _run_static(op, stream=None):
    import torch
    from composable_kernel_b200 import capi
    s = stream if stream is not None else torch.cuda.current_stream().cuda_stream
    return capi.lib().ckb_gemm_run_static_scratch(C.byref(op.problem), C.byref(op.plan), s)


@pytest.mark.gpu
def test_static_scratch_matches_ckb_gemm_run_twice_in_a_row():
    import torch
    from composable_kernel_b200 import capi
    assert int(capi.lib().ckb_static_scratch_bytes()) == 8 << 20
    # two differently shaped problems back to back, each twice: every call shares the one array and finds the previous call's contents
    for (M, N, K, cde, seed) in ((16, 4096, 4096, "add", 1), (5, 1000, 384, "pass", 2), (64, 11008, 1024, "add", 3)):
        op, e = _problem(M, N, K, seed, cde)
        assert op.plan.workspace_bytes <= capi.lib().ckb_static_scratch_bytes()
        op.run()                                    # existing entry point: ckb_gemm_run with a zeroed caller workspace
        torch.cuda.synchronize()
        want = H.from_dev_bf16(e)
        for call in range(2):
            e.fill_(float("nan"))
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            capi.check(_run_static(op))
            torch.cuda.synchronize()
            free_after = torch.cuda.mem_get_info()[0]
            got = H.from_dev_bf16(e)
            bad, err = H.bf16_close(got, want, RTOL, ATOL)
            assert bad == 0, f"{op.kernel_name} M{M} N{N} K{K} call {call}: {err}"
            if call == 1:   # (the first call may load the kernel's module lazily)
                assert free_after == free_before, "the allocation-free entry point changed the device's free memory"


@pytest.mark.gpu
def test_static_scratch_rejects_what_it_cannot_serve():
    import torch
    from composable_kernel_b200 import capi, ops
    # scratch beyond the built-in array: N = 40000 at M = 64 needs 313 tiles x 64 x 128 x 4 B > 8 MiB
    op, _ = _problem(64, 40000, 256, 4, "pass")
    assert op.plan.workspace_bytes > capi.lib().ckb_static_scratch_bytes()
    assert _run_static(op) == capi.ERR_WORKSPACE and "8388608" in capi.last_error()
    op.run()                                        # the existing entry point still serves it
    torch.cuda.synchronize()
    # a plan of a kernel that uses no scratch memory
    a = torch.zeros(4, 512, dtype=torch.bfloat16, device="cuda")
    b = torch.zeros(256, 512, dtype=torch.int8, device="cuda")
    s = torch.ones(256, dtype=torch.bfloat16, device="cuda")
    e = torch.zeros(4, 256, dtype=torch.bfloat16, device="cuda")
    plain = ops.GemmOp(a, b, e, b_op=capi.BOP_MULTIPLY, scale=s)
    assert plain.supported and plain.plan.workspace_bytes == 0
    assert _run_static(plain) == capi.ERR_UNSUPPORTED and "no scratch" in capi.last_error()
