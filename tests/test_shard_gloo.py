"""Host-side logic of the N-sharded (column-parallel) path on CPU: shard arithmetic, and a world_size-2
gloo run in which every rank evaluates ITS column slice (with the oracle standing in for the device
GEMM) and the all-gather reassembles exactly the single-process result (per-column
arithmetic is unchanged by sharding, so the concatenation must be bit-identical)."""
import os
import socket

import numpy as np
import pytest

from composable_kernel_b200 import shard


def test_column_shards_cover_and_align():
    for N, G in [(28672, 8), (28672, 4), (4096, 2), (11008, 8), (1000, 3), (128, 4), (0, 2)]:
        s = shard.column_shards(N, G)
        assert len(s) == G and s[0][0] == 0 and s[-1][1] == N
        for (a, b), (c, d) in zip(s, s[1:]):
            assert b == c and a <= b
        for a, b in s[:-1]:
            assert (b - a) % 128 == 0
    assert shard.equal_shards(28672, 8) and shard.column_shards(28672, 8)[0] == (0, 3584)
    assert not shard.equal_shards(11008, 8)
    with pytest.raises(ValueError):
        shard.column_shards(16, 0)


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from oracle import ck_oracle as orc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(0)  # identical inputs on every rank (A replicated, B sharded by column)
    M, N, K, G = 4, 512, 256, 128
    a = orc.bf16_trunc(rng.standard_normal((M, K)).astype(np.float32))
    qk = rng.integers(0, 16, (K, N)).astype(np.uint8)
    s = orc.bf16_trunc((rng.random((K // G, N)) * 0.01 + 0.005).astype(np.float32))
    n0, n1 = shard.column_shards(N, world)[rank]
    e_loc = orc.gemm(a, orc.BF16, qk[:, n0:n1], orc.U8, orc.BF16, b_op=orc.BOP_GROUP_RTN, scale=s[:, n0:n1],
                     s_dtype=orc.BF16, zero_const=8, group_size=G)
    e_t = torch.from_numpy(e_loc.astype(np.int32))  # gloo has no 16-bit integer transport; NCCL moves bf16 directly
    e_all = shard.all_gather_columns(e_t, world)
    blocked = shard.all_gather_columns(e_t, world, row_major=False)
    full = orc.gemm(a, orc.BF16, qk, orc.U8, orc.BF16, b_op=orc.BOP_GROUP_RTN, scale=s, s_dtype=orc.BF16,
                    zero_const=8, group_size=G)
    ok = np.array_equal(e_all.numpy().astype(np.uint16), full) and tuple(blocked.shape) == (world, M, N // world) \
        and np.array_equal(blocked[rank].numpy().astype(np.uint16), e_loc)
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_two_rank_gloo_all_gather_matches_single_process():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]
