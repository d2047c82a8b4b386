import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from composable_kernel_b200 import capi, ops
M, N, K = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
a = torch.randn(M, K, device="cuda").bfloat16()
lst = []
for _ in range(3):
    packed = torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device="cuda")
    sc = (torch.rand(K // 128, N, device="cuda") * 0.01).bfloat16()
    e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    lst.append(ops.GemmOp(a, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=N, K=K, scale=sc, zero_const=8, group_size=128))
for op in lst: op.run()
torch.cuda.synchronize()
os.environ["CKB_TRACE"] = sys.argv[4]
lst[0].run()
torch.cuda.synchronize()
print(open(sys.argv[4]).read())
