"""Column-parallel (N-sharded) placement of the weight matrix across the GPUs of one node.

[EXT] -- the reference has no multi-GPU code.  Rank g of G owns the output channels
[n0, n1) of every GEMM: its slice of the (packed) weights, scales, zero points and bias.  A is replicated,
no exchange happens during the GEMM; the only collective is one all-gather of the E slices over NVLink
(torch.distributed / NCCL), which lands them in the blocked layout E_all[G][M][N/G] and, on request,
re-interleaves them into row-major E[M][N].
"""
from __future__ import annotations

from typing import List, Tuple


def column_shards(N: int, world: int, align: int = 128) -> List[Tuple[int, int]]:
    """Split N output channels into `world` contiguous ranges whose sizes are multiples of `align`
    (the kernels' 128-channel tile) except possibly the last; earlier ranks get the larger shares."""
    if world <= 0 or N < 0:
        raise ValueError("bad shard request")
    tiles = (N + align - 1) // align
    base, extra = divmod(tiles, world)
    out, n0 = [], 0
    for g in range(world):
        n1 = min(N, n0 + (base + (1 if g < extra else 0)) * align)
        out.append((n0, n1))
        n0 = n1
    return out


def equal_shards(N: int, world: int, align: int = 128) -> bool:
    s = column_shards(N, world, align)
    return len({b - a for a, b in s}) == 1


def all_gather_columns(e_local, world: int, group=None, row_major: bool = True):
    """All-gather the per-rank E slices [M, N/G] (equal sizes).  Returns E_all[G, M, N/G] or, with
    row_major=True, the re-interleaved E[M, N] (one extra permute-copy on the device)."""
    import torch
    import torch.distributed as dist
    M, n_loc = e_local.shape
    flat = torch.empty((world * M, n_loc), dtype=e_local.dtype, device=e_local.device)
    dist.all_gather_into_tensor(flat, e_local.contiguous(), group=group)
    out = flat.view(world, M, n_loc)
    if not row_major:
        return out
    return out.permute(1, 0, 2).reshape(M, world * n_loc)


class PeerGather:
    """Fused all-gather of column-sharded outputs over NVLink peer memory (no collective call on the data path).

    Every rank owns one symmetric allocation (torch.distributed._symmetric_memory: the same buffer mapped into every
    process of the node) holding `slots` gathered outputs E_all[G][M][n_loc] plus an arrival counter and an epoch word.
    The GEMM epilogue stores this rank's slice straight into slot [rank] of EVERY rank's buffer and bumps every rank's
    counter (ckb_gemm_problem.e_peer / flag_peer); `wait()` enqueues ckb_peer_wait, after which the local E_all is
    complete.  Buffers rotate over `slots` >= 2 so that a rank running one step ahead never overwrites a buffer a slower
    rank still reads (a rank can only enter step i + 2 after every rank signalled step i + 1, i.e. finished reading i)."""
    HEADER = 256  # bytes: uint32 arrival counter at 0, the kernels' scratch word at 4, uint32 epoch at 128

    def __init__(self, M: int, n_loc: int, world: int, rank: int, dtype, slots: int = 2, group=None):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.M, self.n_loc, self.world, self.rank, self.slots = M, n_loc, world, rank, slots
        self.esize = torch.empty((), dtype=dtype).element_size()
        self.slot_bytes = (world * M * n_loc * self.esize + 255) // 256 * 256
        total = self.HEADER + slots * self.slot_bytes
        self.buf = symm.empty(total, dtype=torch.uint8, device="cuda")
        self.buf.zero_()
        torch.cuda.synchronize()
        grp = group if group is not None else dist.group.WORLD
        self.hdl = symm.rendezvous(self.buf, grp.group_name)
        self.hdl.barrier()
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.dtype = dtype

    def e_peers(self, slot: int):
        """this rank's [M][n_loc] slice inside slot `slot` of every rank's buffer"""
        off = self.HEADER + slot * self.slot_bytes + self.rank * self.M * self.n_loc * self.esize
        return [p + off for p in self.ptrs]

    def flag_peers(self):
        return list(self.ptrs)

    def gathered(self, slot: int):
        """local view E_all[G][M][n_loc] of slot `slot`"""
        o = self.HEADER + slot * self.slot_bytes
        n = self.world * self.M * self.n_loc * self.esize
        return self.buf[o:o + n].view(self.dtype).view(self.world, self.M, self.n_loc)

    def wait(self, grid: int, stream) -> None:
        """enqueue the consumer-side wait for one step (one arrival per rank; `grid` is kept for callers' bookkeeping)"""
        from . import capi
        base = self.buf.data_ptr()
        capi.check(capi.lib().ckb_peer_wait(base, base + 128, self.world, stream))
