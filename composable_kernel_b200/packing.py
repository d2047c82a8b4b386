"""Weight layout tool for the W4A16 path (host side, numpy): the fragment-major packed-int4 layout of ckb_prepack_b_int4 and the
importers that bring common 4-bit checkpoint formats to it.  One-time weight preparation -- never on the GEMM path (the device
path packs with ckb_prepack_b_int4; this module produces bit-identical buffers offline, e.g. when converting a checkpoint on a
machine without a GPU).  SURVEY.md F3; the reference's only weight pre-processing is UnsignedWeightPreprocessor
(example/64_fpAintB_gemm/common.hpp:56-83) and it has no packed-int4 format.

Packed layout (composable_kernel_b200/csrc/prepack.cu): uint32 words [n16 group][k64 block][lane = 4 g + t][s = 0..3]; word
(ng, kb, lane, s) holds rows r_lo = 16 ng + g and r_hi = r_lo + 8 at k_i = 64 kb + 16 t + 4 s + i:
    nibble 0 = (r_lo,k0)  1 = (r_hi,k0)  2 = (r_lo,k2)  3 = (r_hi,k2)  4 = (r_lo,k1)  5 = (r_hi,k1)  6 = (r_lo,k3)  7 = (r_hi,k3)
so that (w >> 4p) & 0x000F000F is register p of the mma.sync m16n8k16 A fragment.  Codes outside N x K are 0.

CLI:  python -m composable_kernel_b200.packing pack   codes.npy packed.bin      (codes: uint8 [K, N], values 0..15)
      python -m composable_kernel_b200.packing gptq   qweight.npy packed.bin    (int32 [K/8, N], AutoGPTQ order)
      python -m composable_kernel_b200.packing awq    qweight.npy packed.bin    (int32 [K, N/8], AutoAWQ order)
"""
from __future__ import annotations

import sys

import numpy as np

# (row half, k offset) of nibble j of a word
_NIBBLE = [(0, 0), (1, 0), (0, 2), (1, 2), (0, 1), (1, 1), (0, 3), (1, 3)]
AWQ_ORDER = (0, 2, 4, 6, 1, 3, 5, 7)   # AutoAWQ stores the 8 channels of an int32 in this interleaved order


def packed_bytes(N: int, K: int) -> int:
    return ((N + 15) // 16) * ((K + 63) // 64) * 512


def _index_maps(N: int, K: int):
    ng, kb = (N + 15) // 16, (K + 63) // 64
    g = np.arange(8)[:, None].repeat(4, 1).reshape(-1)        # lane -> g
    t = np.tile(np.arange(4), 8)                               # lane -> t
    s = np.arange(4)
    rows = np.empty((ng, 32, 8), np.int64)                     # [group, lane, nibble]
    ks = np.empty((kb, 32, 4, 8), np.int64)                    # [block, lane, s, nibble]
    for j, (half, dk) in enumerate(_NIBBLE):
        rows[:, :, j] = 16 * np.arange(ng)[:, None] + g[None, :] + 8 * half
        ks[:, :, :, j] = 64 * np.arange(kb)[:, None, None] + 16 * t[None, :, None] + 4 * s[None, None, :] + dk
    return rows, ks


def pack_int4(codes_kn: np.ndarray) -> np.ndarray:
    """uint8 codes [K, N] (0..15) -> uint32 words [n16 groups, k64 blocks, 32, 4] (the buffer ckb_prepack_b_int4 writes)."""
    q = np.asarray(codes_kn)
    if q.ndim != 2 or q.dtype != np.uint8 or (q > 15).any():
        raise ValueError("codes must be a uint8 [K, N] array of 4-bit values")
    K, N = q.shape
    ng, kb = (N + 15) // 16, (K + 63) // 64
    padded = np.zeros((kb * 64, ng * 16), np.uint32)
    padded[:K, :N] = q
    rows, ks = _index_maps(N, K)
    words = np.zeros((ng, kb, 32, 4), np.uint32)
    for j in range(8):
        r = rows[:, None, :, None, j]                          # [ng, 1, 32, 1]
        k = ks[None, :, :, :, j]                               # [1, kb, 32, 4]
        words |= padded[k, r] << np.uint32(4 * j)
    return words


def unpack_int4(words: np.ndarray, N: int, K: int) -> np.ndarray:
    """inverse of pack_int4: uint8 codes [K, N]"""
    ng, kb = (N + 15) // 16, (K + 63) // 64
    w = np.asarray(words, np.uint32).reshape(ng, kb, 32, 4)
    rows, ks = _index_maps(N, K)
    out = np.zeros((kb * 64, ng * 16), np.uint8)
    for j in range(8):
        r = np.broadcast_to(rows[:, None, :, None, j], w.shape)
        k = np.broadcast_to(ks[None, :, :, :, j], w.shape)
        out[k, r] = ((w >> np.uint32(4 * j)) & np.uint32(15)).astype(np.uint8)
    return out[:K, :N]


def codes_from_gptq(qweight: np.ndarray) -> np.ndarray:
    """AutoGPTQ 4-bit qweight: int32 [K/8, N], nibble i of word (r, n) = code of k = 8 r + i  ->  uint8 codes [K, N]"""
    qw = np.asarray(qweight).view(np.uint32)
    out = np.empty((qw.shape[0] * 8, qw.shape[1]), np.uint8)
    for i in range(8):
        out[i::8] = (qw >> np.uint32(4 * i)) & np.uint32(15)
    return out


def codes_from_awq(qweight: np.ndarray) -> np.ndarray:
    """AutoAWQ 4-bit qweight: int32 [K, N/8], nibble i of word (k, c) = code of channel 8 c + AWQ_ORDER[i]  ->  codes [K, N]"""
    qw = np.asarray(qweight).view(np.uint32)
    out = np.empty((qw.shape[0], qw.shape[1] * 8), np.uint8)
    for i, ch in enumerate(AWQ_ORDER):
        out[:, ch::8] = (qw >> np.uint32(4 * i)) & np.uint32(15)
    return out


def main(argv) -> int:
    if len(argv) != 4 or argv[1] not in ("pack", "gptq", "awq"):
        print(__doc__)
        return 2
    src = np.load(argv[2])
    codes = src if argv[1] == "pack" else (codes_from_gptq(src) if argv[1] == "gptq" else codes_from_awq(src))
    words = pack_int4(np.ascontiguousarray(codes, dtype=np.uint8))
    words.tofile(argv[3])
    print(f"{argv[1]}: K = {codes.shape[0]}, N = {codes.shape[1]} -> {words.nbytes} bytes ({argv[3]})")
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
