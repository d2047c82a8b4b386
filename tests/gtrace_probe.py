"""Per-CTA globaltimer timeline of the decode kernel (needs lib/libckb_trace.so: `make -C composable_kernel_b200 trace`).
usage: CKB_LIB=composable_kernel_b200/lib/libckb_trace.so python tests/gtrace_probe.py M N K [M N K ...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import torch
from composable_kernel_b200 import capi, ops

NAMES = ["entry", "prologue_done", "prod_prefill_issued", "first_full", "-", "-", "kloop_done", "exit"]


def probe(M, N, K, out):
    a = torch.randn(M, K, device="cuda").bfloat16()
    lst = []
    for _ in range(3):
        packed = torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device="cuda")
        sc = (torch.rand(K // 128, N, device="cuda") * 0.01).bfloat16()
        e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        lst.append(ops.GemmOp(a, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=N, K=K,
                              scale=sc, zero_const=8, group_size=128))
    for op in lst:
        op.run()
    torch.cuda.synchronize()
    if os.path.exists(out):
        os.remove(out)
    os.environ["CKB_GTRACE"] = out
    lst[0].run()
    torch.cuda.synchronize()
    del os.environ["CKB_GTRACE"]
    rows = np.array([[int(x) for x in l.split()[1:]] for l in open(out) if not l.startswith("#")], dtype=np.float64)
    for l in open(out):
        if l.startswith("# chunks"):
            print(l.rstrip())
    print(f"--- M{M} N{N} K{K} grid={len(rows)} kernel={lst[0].kernel_name} (ns after first CTA entry: min / median / max)")
    for k, nm in enumerate(NAMES):
        col = rows[:, k]
        col = col[col >= 0]
        if len(col):
            print(f"  {nm:16s} {col.min():8.0f} {np.median(col):8.0f} {col.max():8.0f}")


if __name__ == "__main__":
    argv = [int(x) for x in sys.argv[1:]]
    for i in range(0, len(argv), 3):
        probe(argv[i], argv[i + 1], argv[i + 2], "/tmp/gtrace.txt")
