"""Probe: K3 (tcgen05 SS GEMM) 1-CTA vs CTA-pair kernel, timed with CUDA events.  usage: python tests/ss_pair_probe.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from composable_kernel_b200 import capi, ops  # noqa: E402


def timed(op, iters=20):
    for _ in range(3):
        op.run()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(iters):
        op.run()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / iters


def main():
    shapes = [(8192, 8192, 8192), (4096, 14336, 4096), (2048, 4096, 4096)]
    if len(sys.argv) > 1:
        shapes = [tuple(int(x) for x in sys.argv[1].split('x'))]
    for (M, N, K) in shapes:
        for dt in ("i8", "bf16"):
            res = {}
            torch.manual_seed(0)
            for pair in ("0", "1"):
                os.environ["CKB_SS_2CTA"] = pair
                torch.manual_seed(M + N + K)
                if dt == "i8":
                    a = torch.randint(-128, 127, (M, K), dtype=torch.int8, device="cuda")
                    b = torch.randint(-128, 127, (N, K), dtype=torch.int8, device="cuda")
                    e = torch.empty(M, N, dtype=torch.int32, device="cuda")
                else:
                    a = torch.randn(M, K, device="cuda").bfloat16()
                    b = torch.randn(N, K, device="cuda").bfloat16()
                    e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
                op = ops.GemmOp(a, b, e)
                assert op.supported, op.why
                ms = timed(op)
                res[pair] = (ms, e.clone())
                print(f"{dt} {M}x{N}x{K} pair={pair} {op.kernel_name}: {ms*1e3:.1f} us  {2*M*N*K/ms/1e9:.0f} T(FL)OP/s", flush=True)
            same = torch.equal(res["0"][1], res["1"][1])
            print(f"   outputs identical: {same}")


if __name__ == "__main__":
    main()
