"""Probe: raster band height of K3 (CKB_SS_BAND = M-blocks of 128 rows per band) vs time.  usage: python tests/ss_band_probe.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from composable_kernel_b200 import ops  # noqa: E402
from tests.ss_pair_probe import timed  # noqa: E402

for (M, N, K) in [(4096, 14336, 4096), (8192, 8192, 8192), (2048, 4096, 4096)]:
    for dt in ("bf16", "i8"):
        torch.manual_seed(1)
        if dt == "i8":
            a = torch.randint(-128, 127, (M, K), dtype=torch.int8, device="cuda")
            b = torch.randint(-128, 127, (N, K), dtype=torch.int8, device="cuda")
            e = torch.empty(M, N, dtype=torch.int32, device="cuda")
        else:
            a = torch.randn(M, K, device="cuda").bfloat16()
            b = torch.randn(N, K, device="cuda").bfloat16()
            e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        out = []
        for band in (4, 8, 16, 32, 64):
            os.environ["CKB_SS_BAND"] = str(band)
            op = ops.GemmOp(a, b, e)
            ms = timed(op)
            out.append(f"band {band}: {ms*1e3:.1f} us ({2*M*N*K/ms/1e9:.0f})")
        print(dt, f"{M}x{N}x{K}", op.kernel_name, " | ".join(out), flush=True)
