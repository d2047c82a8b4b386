"""Quick device-time measurements of single GEMM configurations (CUDA events, rotating operands > L2).
usage: python tests/perf_probe.py [i8 M N K | w4 M N K | w8 M N K] ..."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from composable_kernel_b200 import capi, ops  # noqa: E402


def timeit(ops_list, iters=20):
    if os.environ.get("PERF_ONCE"):
        # three eager launches of the first operand set (for an ncu capture: the third one is warm)
        s = torch.cuda.current_stream().cuda_stream
        for _ in range(3):
            ops_list[0].run(s)
        torch.cuda.synchronize()
        return float("nan")
    if os.environ.get("PERF_GRAPH"):
        # CUDA-graph replay of the whole rotation (launch overhead of the host API taken out)
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            for op in ops_list:
                op.run(st.cuda_stream)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                for op in ops_list:
                    op.run(st.cuda_stream)
            g.replay()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            for _ in range(iters):
                g.replay()
            b.record(st)
            torch.cuda.synchronize()
        return a.elapsed_time(b) * 1e3 / (iters * len(ops_list))
    s = torch.cuda.current_stream().cuda_stream
    for op in ops_list:
        op.run(s)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        for op in ops_list:
            op.run(s)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / (iters * len(ops_list))   # us per launch


def case_i8(M, N, K):
    nrot = max(1, int(300e6 // (M * K + N * K)) + 1)
    lst = []
    for _ in range(min(nrot, 4)):
        a = torch.randint(-128, 128, (M, K), dtype=torch.int8, device="cuda")
        b = torch.randint(-128, 128, (N, K), dtype=torch.int8, device="cuda")
        d0 = torch.rand(N, device="cuda") * 1e-3
        d1 = torch.rand(M, device="cuda") * 1e-3
        e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        lst.append(ops.GemmOp(a, b, e, ds=(d0, d1), d_strides=((0, 1), (1, 0)), cde_op=capi.CDE_MULTIPLY_MULTIPLY))
    us = timeit(lst, 10)
    print(f"i8 mulmul M{M} N{N} K{K}: {us:.1f} us  {2*M*N*K/us/1e6:.1f} TOP/s  kernel={lst[0].kernel_name}")


def BOP_W4(M):
    """numerics contract of the W4A16 runs: PERF_RTN=0/1 forces one, default = the fast one of the regime"""
    r = os.environ.get("PERF_RTN")
    rtn = (M > 64) if r is None else bool(int(r))
    return capi.BOP_GROUP_RTN if rtn else capi.BOP_GROUP


def case_w4(M, N, K):
    nrot = int(400e6 // (N * K // 2)) + 1
    nrot = int(os.environ.get("PERF_NROT", nrot))   # small values keep the weights L2-resident (an upper bound for prefetching)
    lst = []
    a = torch.randn(M, K, device="cuda").bfloat16()
    for _ in range(nrot):
        packed = torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device="cuda")
        sc = (torch.rand(K // 128, N, device="cuda") * 0.01).bfloat16()
        e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        # PERF_NODEP=n: every GEMM whose index is not a multiple of n declares its inputs independent of its predecessor
        nd = int(os.environ.get("PERF_NODEP", 0))
        fl = capi.FLAG_INPUTS_READY if nd and (len(lst) % nd) else 0
        extra = {}
        if os.environ.get("PERF_W4_BIAS"):   # functor epilogue: E = acc + bias[n]
            extra = dict(ds=(torch.randn(N, device="cuda").bfloat16(),), d_strides=((0, 1),), cde_op=capi.CDE_ADD)
        lst.append(ops.GemmOp(a, packed, e, flags=fl, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=BOP_W4(M), N=N, K=K,
                              scale=sc, zero_const=8, group_size=128, **extra))
    lst = lst * int(os.environ.get("PERF_REPEAT", 1))   # longer graphs for small rotations
    us = timeit(lst, 5)
    by = 2 * M * K + N * K // 2 + (K // 128) * N * 2 + 2 * M * N
    print(f"w4 M{M} N{N} K{K}: {us:.2f} us  {by/us/1e3:.0f} GB/s  {2*M*N*K/us/1e6:.2f} TFLOP/s  grid={lst[0].plan.grid} kernel={lst[0].kernel_name}")


def case_w8(M, N, K):
    nrot = int(300e6 // (N * K)) + 1
    lst = []
    a = torch.randn(M, K, device="cuda").bfloat16()
    for _ in range(min(nrot, 8)):
        b = torch.randint(-128, 128, (N, K), dtype=torch.int8, device="cuda")
        sc = (torch.rand(N, device="cuda") * 0.01).bfloat16()
        bias = torch.randn(N, device="cuda").bfloat16()
        e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        lst.append(ops.GemmOp(a, b, e, b_op=capi.BOP_MULTIPLY, scale=sc, ds=(bias,), d_strides=((0, 1),), cde_op=capi.CDE_ADD,
                              flags=int(os.environ.get("PERF_W8_FLAGS", 0))))   # 128 = CKB_FLAG_W8_STREAMK
    us = timeit(lst, 3)
    by = 2 * M * K + N * K + 2 * N * 2 + 2 * M * N
    print(f"w8 mul+bias M{M} N{N} K{K}: {us:.1f} us  {by/us/1e3:.0f} GB/s  {2*M*N*K/us/1e6:.1f} TFLOP/s  kernel={lst[0].kernel_name}")


if __name__ == "__main__":
    argv = sys.argv[1:]
    i = 0
    while i < len(argv):
        kind, M, N, K = argv[i], int(argv[i + 1]), int(argv[i + 2]), int(argv[i + 3])
        {"i8": case_i8, "w4": case_w4, "w8": case_w8}[kind](M, N, K)
        i += 4
