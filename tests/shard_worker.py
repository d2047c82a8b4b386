"""torchrun worker of tests/test_gpu_shard.py: one process per GPU (NCCL).  Every rank builds the SAME global W4A16 problem,
runs the C-ABI GEMM on its column shard (composable_kernel_b200.shard.column_shards), all-gathers the E slices over NVLink and
compares the re-assembled E[M][N] with the single-GPU result of the full problem computed on the same device.
Prints one line per case and exits non-zero on a mismatch."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from composable_kernel_b200 import capi, ops, shard  # noqa: E402


def w4_op(a, packed, sc, z, e, N, K):
    # decode rows: fp32-group-scale contract with the shard-invariant K order; prefill rows: one-rounding contract (the tcgen05
    # kernel's K order never depends on N)
    M = a.shape[0]
    return ops.GemmOp(a, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP_RTN if M > 64 else capi.BOP_GROUP,
                      N=N, K=K, scale=sc, zero=z, zero_const=8, group_size=128, flags=capi.FLAG_FIXED_K_ORDER)


def fused_gather_cases(rank, world, dev):
    """Fused all-gather (epilogue stores into every rank's buffer over NVLink + ckb_peer_wait) against the NCCL gather of the
    same slices, eager and under CUDA-graph replay with rotating buffers."""
    bad = 0
    for M, N, K in [(1, 28672, 8192), (16, 28672, 8192), (4, 4096, 4096)]:
        if not shard.equal_shards(N, world, 16):
            continue
        n_loc = N // world
        g = torch.Generator(device=dev)
        g.manual_seed(99 + M)
        a = torch.randn(M, K, device=dev, generator=g).bfloat16()
        packed = torch.randint(0, 256, (int(capi.lib().ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device=dev, generator=g)
        sc = (torch.rand(K // 128, N, device=dev, generator=g) * 0.015 + 0.005).bfloat16()
        per_group = (K // 64) * 512
        p_loc = packed[(rank * n_loc // 16) * per_group:((rank + 1) * n_loc // 16) * per_group].clone()
        s_loc = sc[:, rank * n_loc:(rank + 1) * n_loc].contiguous()
        e_loc = torch.zeros(M, n_loc, dtype=torch.bfloat16, device=dev)
        ref_op = ops.GemmOp(a, p_loc, e_loc, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=n_loc, K=K, scale=s_loc,
                            zero_const=8, group_size=128, flags=capi.FLAG_FIXED_K_ORDER)
        ref_op.run()
        torch.cuda.synchronize()
        want = shard.all_gather_columns(e_loc, world, row_major=False)        # NCCL reference: E_all[G][M][n_loc]
        pg = shard.PeerGather(M, n_loc, world, rank, torch.bfloat16, slots=2)
        opsl = [ops.GemmOp(a, p_loc, e_loc, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=n_loc, K=K, scale=s_loc,
                           zero_const=8, group_size=128, flags=capi.FLAG_FIXED_K_ORDER, e_peers=pg.e_peers(s), flag_peers=pg.flag_peers(), peer_rank=rank)
                for s in range(2)]
        assert all(o.supported for o in opsl), opsl[0].why
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            for s in range(2):                                               # eager: one step per slot
                opsl[s].run(st.cuda_stream)
                pg.wait(opsl[s].plan.grid, st.cuda_stream)
            torch.cuda.synchronize()
            ok_eager = all(torch.equal(pg.gathered(s).view(torch.int16), want.view(torch.int16)) for s in range(2))
            dist.barrier()
            pg.buf[shard.PeerGather.HEADER:].zero_()
            torch.cuda.synchronize()
            dist.barrier()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                for rep in range(4):
                    for s in range(2):
                        opsl[s].run(st.cuda_stream)
                        pg.wait(opsl[s].plan.grid, st.cuda_stream)
            for _ in range(3):
                gr.replay()
            torch.cuda.synchronize()
            ok_graph = all(torch.equal(pg.gathered(s).view(torch.int16), want.view(torch.int16)) for s in range(2))
            dist.barrier()
        print(f"rank {rank}/{world} fused-gather M{M} N{N} K{K}: eager {ok_eager} graph-replay {ok_graph} kernel {opsl[0].kernel_name}", flush=True)
        bad += 0 if (ok_eager and ok_graph) else 1
        del pg, gr
    return bad


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    dev = torch.device("cuda")
    cases = [(1, 28672, 8192, False), (16, 28672, 8192, True), (4096, 28672, 8192, False), (4, 4096, 4096, False),
             (16, 11008, 4096, False), (256, 4096, 1024, True)]
    if os.environ.get("SHARD_SMALL"):
        cases = [(4, 4096, 1024, True), (256, 2048, 1024, False)]
    bad = 0
    for M, N, K, use_zp in cases:
        if not shard.equal_shards(N, world, 16):
            continue
        g = torch.Generator(device=dev)
        g.manual_seed(1234 + M)          # identical global problem on every rank
        a = torch.randn(M, K, device=dev, generator=g).bfloat16()
        q = torch.randint(0, 16, (K, N), dtype=torch.uint8, device=dev, generator=g)
        sc = (torch.rand(K // 128, N, device=dev, generator=g) * 0.015 + 0.005).bfloat16()
        z = torch.randint(0, 16, (K // 128, N), dtype=torch.uint8, device=dev, generator=g) if use_zp else None
        packed = ops.prepack_int4(q)
        del q
        # single-GPU result of the full problem
        e_full = torch.zeros(M, N, dtype=torch.bfloat16, device=dev)
        op = w4_op(a, packed, sc, z, e_full, N, K)
        assert op.supported, op.why
        op.run()
        # this rank's shard: contiguous groups of the fragment-major packed weight, columns of scales / zero points
        n0, n1 = shard.column_shards(N, world, 16)[rank]
        per_group = (K // 64) * 512
        p_loc = packed[(n0 // 16) * per_group:(n1 // 16) * per_group].clone()
        s_loc = sc[:, n0:n1].contiguous()
        z_loc = z[:, n0:n1].contiguous() if use_zp else None
        e_loc = torch.zeros(M, n1 - n0, dtype=torch.bfloat16, device=dev)
        op_l = w4_op(a, p_loc, s_loc, z_loc, e_loc, n1 - n0, K)
        assert op_l.supported, op_l.why
        op_l.run()
        torch.cuda.synchronize()
        e_all = shard.all_gather_columns(e_loc, world)          # row-major E[M][N]
        blocked = shard.all_gather_columns(e_loc, world, row_major=False)
        torch.cuda.synchronize()
        same = torch.equal(e_all.view(torch.int16), e_full.view(torch.int16))
        own = torch.equal(blocked[rank].view(torch.int16), e_loc.view(torch.int16))
        nd = int((e_all.view(torch.int16) != e_full.view(torch.int16)).sum().item())
        md = float((e_all.float() - e_full.float()).abs().max().item())
        print(f"rank {rank}/{world} M{M} N{N} K{K} zp={int(use_zp)}: gathered==single {same} (differing elements {nd}, max |diff| {md:.3e}) "
              f"own-slice {own}  kernels {op.kernel_name} | {op_l.kernel_name}", flush=True)
        if not (same and own):
            bad += 1
        del packed, p_loc, e_full, e_all, blocked
        torch.cuda.empty_cache()
    bad += fused_gather_cases(rank, world, dev)
    t = torch.tensor([bad], device=dev)
    dist.all_reduce(t)
    dist.destroy_process_group()
    sys.exit(1 if int(t.item()) else 0)


if __name__ == "__main__":
    main()
