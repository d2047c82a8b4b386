"""Per-warp timeline of the W4 decode kernel over consecutive launches of a CUDA graph (steady state under programmatic
dependent launch).  Needs the trace build: `make -C composable_kernel_b200 trace`.
usage: CKB_LIB=composable_kernel_b200/lib/libckb_trace.so CKB_G2_TRACE=/tmp/g2.txt python tests/g2trace_probe.py M N K [nlaunch]
then:  python tests/g2trace_probe.py --report /tmp/g2.txt"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

NAMES = ["entry", "prefill_issued", "pdl_wait_done", "first_full", "kloop_done", "exit"]


def report(path):
    rows = np.loadtxt(path, dtype=np.int64)
    launches = sorted(set(rows[:, 0]))
    launches = launches[len(launches) // 2:]   # first half = eager warm-up launches, second half = the replayed graph
    rows = rows[rows[:, 0] >= launches[0]]
    t0 = rows[rows[:, 3] > 0][:, 3].min()
    print(f"{len(launches)} launches; times in ns relative to the first entry of launch 0; min / median / max over all warps")
    prev_exit = None
    for l in launches:
        r = rows[rows[:, 0] == l]
        r = r[r[:, 3] > 0]
        if not len(r):
            continue
        line = [f"L{l:02d}"]
        for k, nm in enumerate(NAMES):
            col = r[:, 3 + k] - t0
            line.append(f"{nm}={col.min():.0f}/{np.median(col):.0f}/{col.max():.0f}")
        wait, issue, loop, tot = r[:, 9], r[:, 10], r[:, 11], r[:, 12]
        ghz = np.median(tot / np.maximum(1, r[:, 8] - r[:, 3]))
        line.append(f"wait_clk={np.median(wait):.0f}/{wait.max():.0f} issue_clk={np.median(issue):.0f}/{issue.max():.0f} "
                    f"loop_clk={np.median(loop):.0f}/{loop.max():.0f} sm_ghz={ghz:.2f}")
        ex = (r[:, 8] - t0).max()
        if prev_exit is not None:
            line.append(f"period={ex - prev_exit:.0f}")
        prev_exit = ex
        print("  ".join(line))


def main():
    if sys.argv[1] == "--report":
        return report(sys.argv[2])
    import torch
    from composable_kernel_b200 import capi, ops
    M, N, K = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    nl = int(sys.argv[4]) if len(sys.argv) > 4 else 12
    a = torch.randn(M, K, device="cuda").bfloat16()
    lst = []
    for _ in range(nl):
        packed = torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device="cuda")
        sc = (torch.rand(K // 128, N, device="cuda") * 0.01).bfloat16()
        e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        lst.append(ops.GemmOp(a, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=N, K=K,
                              scale=sc, zero_const=8, group_size=128))
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for op in lst:
            op.run(st.cuda_stream)           # warm-up (eager; also allocates the trace buffer outside the capture)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for op in lst:
                op.run(st.cuda_stream)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
    print(f"traced {nl} launches of {lst[0].kernel_name}")


if __name__ == "__main__":
    main()
