"""Which kernel serves the fp32-group-scale contract at M > 32 / > 64, and does every (M, N, K) run (eager launches)?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from composable_kernel_b200 import capi, ops

import itertools
cases = [(M, N, K) for (N, K), M in itertools.product([(4096, 4096), (11008, 4096), (4096, 11008), (28672, 8192), (1024, 1024)],
                                                      [8, 9, 17, 24, 32, 33, 64, 65, 130])]
for (M, N, K) in cases:
    a = torch.randn(M, K, device="cuda").bfloat16()
    packed = torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device="cuda")
    sc = (torch.rand(K // 128, N, device="cuda") * 0.01).bfloat16()
    e = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    op = ops.GemmOp(a, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=N, K=K, scale=sc, zero_const=8, group_size=128)
    print(M, N, K, op.supported, op.kernel_name, "grid", op.plan.grid, "smem", op.plan.smem_bytes, flush=True)
    op.run()
    torch.cuda.synchronize()
    print("   ok", flush=True)
