"""Spans of back-to-back launches of the W4 decode kernel (trace build): how much consecutive kernels overlap under PDL.
usage: CKB_LIB=composable_kernel_b200/lib/libckb_trace.so CKB_GSPAN=/tmp/span.txt python tests/gspan_probe.py M N K [graph]"""
import os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
if os.environ.get("GSPAN_CHILD") != "1":
    env = dict(os.environ, GSPAN_CHILD="1", CKB_GSPAN="/tmp/span.txt", CKB_GTRACE_AT="70")
    subprocess.check_call([sys.executable] + sys.argv, env=env)
    import numpy as np
    r = np.loadtxt("/tmp/span.txt", dtype=np.float64)
    r = r[-40:]
    t0 = r[0, 1]
    print("launch: entry  first_full  pdl_done  exit   (ns, relative)   | period")
    for i in range(len(r)):
        print(f"{int(r[i,0]):5d}: {r[i,1]-t0:9.0f} {r[i,3]-t0:9.0f} {r[i,4]-t0:9.0f} {r[i,2]-t0:9.0f} | {r[i,1]-r[i-1,1] if i else 0:6.0f}  len {r[i,2]-r[i,1]:6.0f}")
    if os.path.exists("/tmp/span.txt.cta"):
        NAMES = ["entry", "prologue_done", "prefill_issued", "first_full", "-", "-", "kloop_done", "exit"]
        lines = open("/tmp/span.txt.cta").read().splitlines()
        rows = np.array([[int(x) for x in l.split()[1:]] for l in lines if l and not l.startswith("#")], dtype=np.float64)
        for l in lines:
            if l.startswith("#"): print(l)
        if len(rows):
            for k, nm in enumerate(NAMES):
                col = rows[:, k]; col = col[col >= 0]
                if len(col): print(f"  {nm:16s} {col.min():8.0f} {np.median(col):8.0f} {col.max():8.0f}")
    sys.exit(0)
import torch
from composable_kernel_b200 import capi, ops
M, N, K = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
use_graph = len(sys.argv) > 4
a = torch.randn(M, K, device="cuda").bfloat16()
lst = []
for _ in range(40):
    packed = torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device="cuda")
    sc = (torch.rand(K // 128, N, device="cuda") * 0.01).bfloat16()
    e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    lst.append(ops.GemmOp(a, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=N, K=K, scale=sc, zero_const=8, group_size=128))
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    for op in lst: op.run(st.cuda_stream)
    torch.cuda.synchronize()
    if use_graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for op in lst: op.run(st.cuda_stream)
        g.replay(); torch.cuda.synchronize()
    else:
        for op in lst: op.run(st.cuda_stream)
        torch.cuda.synchronize()
