"""Stress probe of the stream-K decode kernel (CKB_FLAG_W8_STREAMK): repeats the launch sequence of the parity test -- scratch
cleared by torch.zeros (a memset) immediately in front of the launch -- and counts runs whose output differs from the first
HMMA-kernel result by more than the parity tolerance.  PAD_MB > 0 puts the scratch at the END of a larger zeros() tensor so that
the clear takes longer (widens the window of a launch that starts before the clear has finished).
usage: python tests/streamk_stress_probe.py [iters] [pad_mb]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ctypes as C  # noqa: E402

import torch  # noqa: E402

from composable_kernel_b200 import capi, ops  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
pad_mb = int(sys.argv[2]) if len(sys.argv) > 2 else 0
torch.manual_seed(0)
for (M, N, K) in [(1, 11008, 4096), (16, 14336, 4096), (16, 4096, 4096)]:
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randint(-128, 128, (N, K), dtype=torch.int8, device="cuda")
    sc = (torch.rand(N, device="cuda") * 0.02 + 0.001).bfloat16()
    bias = torch.randn(N, device="cuda").bfloat16()
    e0 = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    ref_op = ops.GemmOp(a, b, e0, b_op=capi.BOP_MULTIPLY, scale=sc, ds=(bias,), d_strides=((0, 1),), cde_op=capi.CDE_ADD)
    ref_op.run()
    torch.cuda.synchronize()
    ref = e0.float()
    e = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    op = ops.GemmOp(a, b, e, b_op=capi.BOP_MULTIPLY, scale=sc, ds=(bias,), d_strides=((0, 1),), cde_op=capi.CDE_ADD,
                    flags=capi.FLAG_W8_STREAMK)
    assert "streamk" in op.kernel_name, op.kernel_name
    wsb = int(op.plan.workspace_bytes)
    bad_runs, bad_tiles, dirty = 0, set(), 0
    s = torch.cuda.current_stream().cuda_stream
    for it in range(iters):
        e.fill_(0)
        big = torch.zeros(pad_mb * (1 << 20) + wsb, dtype=torch.uint8, device="cuda")   # the clear, right in front of the launch
        ws = big[pad_mb * (1 << 20):]
        capi.check(capi.lib().ckb_gemm_run(C.byref(op.problem), C.byref(op.plan), ws.data_ptr(), s))
        torch.cuda.synchronize()
        err = (e.float() - ref).abs()
        bad = err > (1e-2 + 1e-2 * ref.abs())
        if bad.any():
            bad_runs += 1
            dirty += int(ws.view(torch.int32).abs().max()) != 0   # a partial sum that arrived after the fix-up read
            tiles = sorted(set((bad.any(dim=0).nonzero().flatten() // 128).tolist()))
            bad_tiles |= set(tiles)
            if bad_runs <= 2:
                # which stream-K segment of the tile is missing (or doubled)?  partial products of every CTA's k range, in fp32
                grid, ku = int(op.plan.grid), K // 128
                for t in tiles[:2]:
                    cols = slice(128 * t, min(128 * t + 128, N))
                    diff = (e.float() - ref)[:, cols]                       # got - ref
                    rows = bad[:, cols].any(dim=0).nonzero().flatten().tolist()
                    toks = bad[:, cols].any(dim=1).nonzero().flatten().tolist()
                    r0 = rows[0]
                    line = (f"   run {it} tile {t}: bad elements {int(bad[:, cols].sum())}, channel rows {rows[0]}..{rows[-1]} ({len(rows)}), tokens {toks}, "
                            f"row {r0}: got {e.float()[:, cols][:, r0].tolist()[:4]} ref {ref[:, cols][:, r0].tolist()[:4]}\n      ")
                    for cta in range(grid):
                        b0, e0_ = C.c_int64(), C.c_int64()
                        capi.lib().ckb_streamk_range(N // 128 * ku if N % 128 == 0 else ((N + 127) // 128) * ku, cta, grid, C.byref(b0), C.byref(e0_))
                        lo, hi = max(b0.value, t * ku), min(e0_.value, (t + 1) * ku)
                        if lo >= hi:
                            continue
                        k0, k1 = (lo - t * ku) * 128, (hi - t * ku) * 128
                        part = (a[:, k0:k1].float() @ b[cols, k0:k1].float().t()) * sc[cols].float()
                        rr = torch.tensor(rows, device="cuda")
                        rel = float((diff + part)[:, rr].abs().max() / part[:, rr].abs().max())       # ~0 when exactly this segment is missing (bad rows only)
                        rel2 = float((diff - part)[:, rr].abs().max() / part[:, rr].abs().max())      # ~0 when it was added twice
                        line += f"[cta {cta} k {k0}-{k1}: missing? {rel:.3f} doubled? {rel2:.3f}] "
                    print(line)
            del big, ws
            continue
        # the scratch must be zero again on exit
        assert int(ws.view(torch.int32).abs().max()) == 0, "scratch not zero on exit"
    print(f"M{M} N{N} K{K} {op.kernel_name}: {bad_runs} bad runs of {iters} (pad {pad_mb} MB; scratch left non-zero in {dirty}), tiles {sorted(bad_tiles)[:12]}")
