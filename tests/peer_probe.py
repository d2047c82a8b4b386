"""Cost decomposition of the fused all-gather (torchrun worker): GEMM alone / + peer stores / + flags / + wait, graph replay.
CKB_GEMV_DBG bits (debug only): 32 = data stores to the own buffer only, 64 = arrival counts to the own counter only;
bits 8.. = own rank."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from composable_kernel_b200 import capi, ops, shard

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
dev = torch.device("cuda")
N, K = 28672, 8192
n_loc = N // world
for M in (1, 16):
    a = torch.randn(M, K, device=dev).bfloat16()
    nrot = 8
    sh = []
    for _ in range(nrot):
        sh.append((torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(n_loc, K)),), dtype=torch.uint8, device=dev),
                   (torch.rand(K // 128, n_loc, device=dev) * 0.01).bfloat16()))
    e = torch.empty(M, n_loc, dtype=torch.bfloat16, device=dev)
    pg = shard.PeerGather(M, n_loc, world, rank, torch.bfloat16, slots=2)
    plain = [ops.GemmOp(a, p, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=n_loc, K=K, scale=s, zero_const=8, group_size=128) for p, s in sh]
    fused = [ops.GemmOp(a, p, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=n_loc, K=K, scale=s, zero_const=8, group_size=128,
                        e_peers=pg.e_peers(i % 2), flag_peers=pg.flag_peers(), peer_rank=rank) for i, (p, s) in enumerate(sh)]
    st = torch.cuda.Stream()

    def timed(body, label):
        with torch.cuda.stream(st):
            body(); torch.cuda.synchronize(); dist.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                body()
            g.replay(); torch.cuda.synchronize(); dist.barrier()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record(st)
            for _ in range(5):
                g.replay()
            t1.record(st); torch.cuda.synchronize(); dist.barrier()
        t = torch.tensor([t0.elapsed_time(t1) * 1e3 / (5 * nrot)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"M{M} world {world} {label}: {t.item():.2f} us", flush=True)

    def run_plain():
        for op in plain: op.run(st.cuda_stream)
    def run_fused_nowait():
        for op in fused: op.run(st.cuda_stream)
    def run_fused():
        for op in fused:
            op.run(st.cuda_stream); pg.wait(op.plan.grid, st.cuda_stream)
    def run_plain_wait():   # plain GEMM + a wait that is satisfied by nobody -> skip
        pass
    timed(run_plain, "gemm only (" + plain[0].kernel_name + ")")
    # the fused code path on ORDINARY device memory (one "peer" = a cudaMalloc buffer): separates the kernel's own cost (fence,
    # counter, generic epilogue) from the cost of writing peer-mapped symmetric memory
    e2 = torch.empty(M, n_loc, dtype=torch.bfloat16, device=dev)
    fl = torch.zeros(64, dtype=torch.int32, device=dev)
    loc = [ops.GemmOp(a, p, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=n_loc, K=K, scale=s, zero_const=8, group_size=128,
                      e_peers=[e2.data_ptr()], flag_peers=[fl.data_ptr()], peer_rank=0) for p, s in sh]
    def run_loc():
        for op in loc: op.run(st.cuda_stream)
    timed(run_loc, "fused code path, one peer = ordinary device memory, no wait (" + loc[0].kernel_name + ")")
    sy = [ops.GemmOp(a, p, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=n_loc, K=K, scale=s, zero_const=8, group_size=128,
                     e_peers=[pg.e_peers(0)[rank]], flag_peers=[pg.flag_peers()[rank]], peer_rank=0) for p, s in sh]
    def run_sy():
        for op in sy: op.run(st.cuda_stream)
    timed(run_sy, "fused code path, one peer = own symmetric buffer, no wait")
    os.environ["CKB_GEMV_DBG"] = "128"
    timed(run_loc, "   same on ordinary memory, signalling block skipped")
    os.environ["CKB_GEMV_DBG"] = "256"
    timed(run_loc, "   same on ordinary memory, no explicit fence (release add only)")
    os.environ["CKB_GEMV_DBG"] = "0"
    os.environ["CKB_GEMV_DBG"] = str(32 | 64 | 0)
    timed(run_fused_nowait, "fused kernel, local stores + local flags, no wait")
    os.environ["CKB_GEMV_DBG"] = str(64 | 0)
    timed(run_fused_nowait, "fused kernel, peer stores + local flags, no wait")
    os.environ["CKB_GEMV_DBG"] = str(32 | 0)
    timed(run_fused_nowait, "fused kernel, local stores + peer flags, no wait")
    os.environ["CKB_GEMV_DBG"] = "0"
    timed(run_fused_nowait, "fused kernel, peer stores + peer flags, no wait")
    pg.buf[:256].zero_(); torch.cuda.synchronize(); dist.barrier()
    timed(run_fused, "fused kernel + wait (full)")
    del pg
dist.destroy_process_group()
