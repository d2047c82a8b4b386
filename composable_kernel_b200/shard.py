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
