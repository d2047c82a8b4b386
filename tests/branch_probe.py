"""Graph-topology probe for the W4A16 decode kernel: the same list of GEMMs captured as one linear chain or as B parallel
branches (independent GEMMs of a layer -- q|k|v, gate|up -- on forked streams).  usage: python tests/branch_probe.py M N K [branches...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from composable_kernel_b200 import capi, ops  # noqa: E402


def build(M, N, K, n):
    a = torch.randn(M, K, device="cuda").bfloat16()
    lst = []
    for _ in range(n):
        packed = torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device="cuda")
        sc = (torch.rand(K // 128, N, device="cuda") * 0.01).bfloat16()
        e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        lst.append(ops.GemmOp(a, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=N, K=K,
                              scale=sc, zero_const=8, group_size=128))
    return lst


def time_graph(lst, branches, iters=5):
    main = torch.cuda.Stream()
    side = [torch.cuda.Stream() for _ in range(branches - 1)]
    def body():
        # groups of `branches` independent GEMMs: fork, one GEMM per stream, join
        for i in range(0, len(lst), branches):
            grp = lst[i:i + branches]
            if len(grp) > 1:
                ev = torch.cuda.Event()
                ev.record(main)
                for s in side[:len(grp) - 1]:
                    s.wait_event(ev)
            grp[0].run(main.cuda_stream)
            for op, s in zip(grp[1:], side):
                op.run(s.cuda_stream)
                e2 = torch.cuda.Event()
                e2.record(s)
                main.wait_event(e2)
    with torch.cuda.stream(main):
        body()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=main):
            body()
        g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(main)
        for _ in range(iters):
            g.replay()
        b.record(main)
        torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / (iters * len(lst))


if __name__ == "__main__":
    M, N, K = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    brs = [int(x) for x in sys.argv[4:]] or [1, 2, 3]
    n = int(400e6 // (N * K // 2)) + 1
    n = n // 6 * 6
    lst = build(M, N, K, n)
    by = 2 * M * K + N * K // 2 + (K // 128) * N * 2 + 2 * M * N
    for b in brs:
        us = time_graph(lst, b)
        print(f"w4 M{M} N{N} K{K} branches={b}: {us:.2f} us per GEMM  {by/us/1e3:.0f} GB/s  ({lst[0].kernel_name})")
