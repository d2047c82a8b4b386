"""Host-only unit tests (no GPU): the scheduler arithmetic the kernels use (exported through the C-ABI test hooks, in the style
of the reference's test/block_to_ctile_map/test_block_to_ctile_map.cpp), the int4 pack / unpack round trip and checkpoint
importers (test/data_type/test_int4.cpp pins the 4-bit value range), and the reference-held conversion identities."""
import ctypes as C

import numpy as np
import pytest

from composable_kernel_b200 import capi, packing


# ------------------------------------------------------------------ output-tile raster (K3 / fused prefill kernels)
@pytest.mark.parametrize("mb,nb,group", [(32, 56, 8), (30, 16, 8), (1, 112, 8), (7, 3, 4), (64, 64, 16), (5, 1, 8)])
def test_tile_raster_visits_every_tile_once_and_keeps_bands(mb, nb, group):
    lib = capi.lib()
    seen = np.zeros((mb, nb), np.int32)
    m, n = C.c_int32(), C.c_int32()
    prev_band = -1
    for t in range(mb * nb):
        assert lib.ckb_tile_coord(t, mb, nb, group, C.byref(m), C.byref(n)) == 0
        assert 0 <= m.value < mb and 0 <= n.value < nb
        seen[m.value, n.value] += 1
        band = m.value // group
        assert band >= prev_band          # bands are finished one after the other (a band of A stays L2 resident)
        prev_band = band
    assert (seen == 1).all()
    # inside a band consecutive tiles walk down M first: concurrently running CTAs share the B panel (same n block)
    assert lib.ckb_tile_coord(0, mb, nb, group, C.byref(m), C.byref(n)) == 0 and (m.value, n.value) == (0, 0)
    if mb > 1:
        lib.ckb_tile_coord(1, mb, nb, group, C.byref(m), C.byref(n))
        assert (m.value, n.value) == (1, 0)
    assert lib.ckb_tile_coord(mb * nb, mb, nb, group, C.byref(m), C.byref(n)) != 0    # out of range is rejected


# ------------------------------------------------------------------ stream-K ranges (K2b)
@pytest.mark.parametrize("units,grid", [(1000, 148), (148, 148), (147, 148), (1, 148), (0, 4), (123457, 296), (2 ** 33 + 5, 1184)])
def test_streamk_ranges_partition_the_unit_space(units, grid):
    lib = capi.lib()
    b, e = C.c_int64(), C.c_int64()
    end_prev, sizes = 0, []
    for cta in range(grid):
        assert lib.ckb_streamk_range(units, cta, grid, C.byref(b), C.byref(e)) == 0
        assert b.value == end_prev and e.value >= b.value       # contiguous, no gap, no overlap
        end_prev = e.value
        sizes.append(e.value - b.value)
    assert end_prev == units
    assert max(sizes) - min(sizes) <= 1                          # equal shares up to the remainder


# ------------------------------------------------------------------ W4A16 decode schedule (K2)
NAMES = ["nt", "rb", "n_groups", "kb_total", "grid", "passes", "n_slots", "slot_q", "slot_r", "g_hi", "wn_log2", "gw", "kbu", "upw",
         "n_chunks", "ring", "stage_bytes", "off_w", "off_s", "off_z", "red_off", "bar_off", "smem_bytes", "warps"]


def sched(M, N, K, zp=0, sms=148, mode=0):
    out = (C.c_int32 * 24)()
    rc = capi.lib().ckb_gemv_schedule(M, N, K, zp, sms, mode, out)
    return rc, dict(zip(NAMES, list(out)))


@pytest.mark.parametrize("M", [1, 4, 8, 16, 32])
@pytest.mark.parametrize("N,K", [(4096, 4096), (11008, 4096), (4096, 11008), (28672, 8192), (3584, 8192), (128, 128), (1000, 384), (16, 64)])
@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
def test_decode_schedule_covers_the_problem_and_fits_the_sm(M, N, K, mode):
    rc, s = sched(M, N, K, zp=1, mode=mode)
    assert rc == 0
    assert s["n_groups"] == (N + 15) // 16 and s["kb_total"] == (K + 63) // 64
    assert 1 <= s["grid"] <= 148 and s["n_slots"] == s["grid"] * s["passes"]
    # every 16-channel group belongs to exactly one slot
    assert s["slot_q"] * s["n_slots"] + s["slot_r"] == s["n_groups"] and 0 <= s["slot_r"] < s["n_slots"]
    assert s["g_hi"] == s["slot_q"] + (1 if s["slot_r"] else 0)
    wn = 1 << s["wn_log2"]
    wk = s["warps"] // wn
    assert s["gw"] * wn >= s["g_hi"]                              # the warp grid holds the largest slot
    chk = wk * s["upw"] * s["kbu"]
    assert s["n_chunks"] * chk >= s["kb_total"] > (s["n_chunks"] - 1) * chk   # the stages cover K with no empty stage
    assert s["ring"] >= 1 and s["ring"] * s["stage_bytes"] <= s["bar_off"]
    assert s["smem_bytes"] <= 227 * 1024
    assert s["off_w"] < s["off_s"] <= s["off_z"] <= s["stage_bytes"]
    if mode & 1:   # CKB_FLAG_FIXED_K_ORDER: the K order may not depend on the shape
        assert (s["wn_log2"], s["upw"], s["kbu"]) == ((1 if s["nt"] == 1 else 0), 1, 2) and s["gw"] >= 2
    if mode & 6:   # RTN numerics (2), peer-store or functor epilogue (4): those instantiations exist for 128-k units only
        assert s["kbu"] == 2


@pytest.mark.parametrize("M", [1, 8, 16, 32])
@pytest.mark.parametrize("N,K", [(4096, 4096), (11008, 4096), (14336, 4096), (28672, 8192), (4096, 14336), (1024, 512)])
def test_decode_schedule_int8_weights_shrinks_slots_until_a_stage_fits(M, N, K):
    """mode 8 (int8 weights, 1 KB per group and 64-k block): cfg3's N = 14336 gives 7 groups per CTA, whose ring stage does not fit
    in shared memory -- the schedule then takes more passes with smaller slots instead of handing the shape to the stream-K kernel.
    A schedule is either refused (the dispatcher falls back to stream-K) or has no empty slot."""
    rc, s = sched(M, N, K, mode=8)
    if rc != 0:
        # 32-row activation stages leave room for one group per slot only (more slots than groups)
        assert M == 32
        return
    assert s["slot_q"] >= 1 and s["slot_q"] * s["n_slots"] + s["slot_r"] == s["n_groups"] and s["n_slots"] == s["grid"] * s["passes"]
    wn = 1 << s["wn_log2"]
    assert s["gw"] * wn >= s["g_hi"] and s["kbu"] == 2
    assert s["ring"] >= (2 if s["passes"] * s["n_chunks"] > 1 else 1)
    assert s["ring"] * s["stage_bytes"] <= s["bar_off"] and s["smem_bytes"] <= 227 * 1024


def test_fixed_k_order_schedule_is_shard_invariant():
    for M in (1, 16):
        ref = None
        for n in (28672, 14336, 7168, 3584, 4096, 11008):
            rc, s = sched(M, n, 8192, mode=1)
            assert rc == 0
            key = (s["nt"], s["wn_log2"], s["upw"], s["kbu"])
            ref = ref or key
            assert key == ref


def test_decode_schedule_rejects_more_than_32_rows():
    rc, _ = sched(33, 4096, 4096)
    assert rc == capi.ERR_UNSUPPORTED if hasattr(capi, "ERR_UNSUPPORTED") else rc != 0


# ------------------------------------------------------------------ int4 pack / unpack, importers
@pytest.mark.parametrize("K,N", [(64, 16), (128, 48), (192, 100), (100, 24), (4096, 256)])
def test_int4_pack_unpack_round_trip(K, N):
    rng = np.random.default_rng(K + N)
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    words = packing.pack_int4(q)
    assert words.nbytes == packing.packed_bytes(N, K) == int(capi.lib().ckb_prepack_b_int4_bytes(N, K))
    assert np.array_equal(packing.unpack_int4(words, N, K), q)
    # the documented fragment property: (w >> 4p) & 0x000F000F = (r_lo | r_hi alternating, k pairs) -- register p of the MMA fragment
    w = int(words[0, 0, 5, 2])                 # group 0, block 0, lane 5 (g = 1, t = 1), s = 2: rows 1 and 9, k = 16 + 8 + {0..3}
    k0 = 16 * 1 + 4 * 2
    def code(r, k):
        return int(q[k, r]) if r < N and k < K else 0
    assert (w >> 0) & 0x000F000F == code(1, k0) | code(1, k0 + 1) << 16
    assert (w >> 4) & 0x000F000F == code(9, k0) | code(9, k0 + 1) << 16
    assert (w >> 8) & 0x000F000F == code(1, k0 + 2) | code(1, k0 + 3) << 16
    assert (w >> 12) & 0x000F000F == code(9, k0 + 2) | code(9, k0 + 3) << 16


def test_int4_value_range_matches_reference_limits():
    """test/data_type/test_int4.cpp:39-44: NumericLimits<int4_t> = [-8, 7]; codes 0..15 with the symmetric zero point 8 span it"""
    codes = np.arange(16, dtype=np.int32)
    assert (codes - 8).min() == -8 and (codes - 8).max() == 7
    with pytest.raises(ValueError):
        packing.pack_int4(np.full((64, 16), 16, np.uint8))


def test_gptq_and_awq_importers():
    rng = np.random.default_rng(5)
    K, N = 128, 64
    q = rng.integers(0, 16, (K, N), dtype=np.uint8)
    gptq = np.zeros((K // 8, N), np.uint32)
    for i in range(8):
        gptq |= q[i::8].astype(np.uint32) << np.uint32(4 * i)
    assert np.array_equal(packing.codes_from_gptq(gptq.view(np.int32)), q)
    awq = np.zeros((K, N // 8), np.uint32)
    for i, ch in enumerate(packing.AWQ_ORDER):
        awq |= q[:, ch::8].astype(np.uint32) << np.uint32(4 * i)
    assert np.array_equal(packing.codes_from_awq(awq.view(np.int32)), q)


# ------------------------------------------------------------------ reference-held conversion identities vs the oracle
def test_fast_u8_to_f16_converter_identity_all_codes():
    """FastNumericArrayConverter<uint8_t, half_t> (element/unary_element_wise_operation.hpp:1061-1091): (0x6400 | u8) read as fp16
    minus 0x6480 (= 1152) is the int8 value whose biased code is u8 -- for ALL 256 codes; and the oracle's fpAintB dequant (q * s with
    s = 1) returns exactly that value."""
    from oracle import ck_oracle as orc
    u = np.arange(256, dtype=np.uint16)
    h = (np.uint16(0x6400) | u).view(np.float16)
    magic = np.array([0x6480], np.uint16).view(np.float16)[0]
    got = (h - magic).astype(np.float32)
    want = (u.astype(np.int32) - 128).astype(np.float32)
    assert np.array_equal(got, want)
    q = (u.astype(np.int32) - 128).astype(np.int8)[None, :]            # B[k = 0, n] = every int8 value
    a = np.ones((1, 1), np.float16)
    e = orc.gemm(a, orc.F16, q, orc.I8, orc.F16, b_op=orc.BOP_FPAINTB, scale=np.ones((1, 256), np.float16), s_dtype=orc.F16)
    assert np.array_equal(e.astype(np.float32)[0], want)


def test_multiply_bf16_int8_functor_all_codes_times_scale_grid():
    """Multiply<bhalf_t, int8_t, bhalf_t> (element/binary_element_wise_operation.hpp:184-190): y = bf16(float(x0) * float(x1)) with CK's
    truncating float -> bf16 conversion; every int8 value against a grid of bf16 scales, oracle GEMM with K = 1 and A = 1."""
    from oracle import ck_oracle as orc
    scales = orc.bf16_trunc(np.array([1.0, 0.5, 0.0078125, 0.013, 0.0999, 1.7, 3.14159, 1e-3, 127.0, 2.0 ** -20], np.float32))
    qs = np.arange(-128, 128, dtype=np.int32)
    q = np.repeat(qs, len(scales)).astype(np.int8)[None, :]
    s = np.tile(scales, len(qs))[None, :]
    a = orc.bf16_trunc(np.ones((1, 1), np.float32))
    e = orc.gemm(a, orc.BF16, q, orc.I8, orc.BF16, b_op=orc.BOP_MULTIPLY, scale=s, s_dtype=orc.BF16)
    prod = q[0].astype(np.float32) * orc.bf16_to_f32(s[0])
    want = (prod.view(np.uint32) >> 16).astype(np.uint16)               # truncation = the upper 16 bits
    assert np.array_equal(e[0], want)


# ------------------------------------------------------------------ torch custom ops: registration + fake (meta) shapes
def test_torch_custom_ops_trace_with_fake_tensors():
    import torch
    from torch._subclasses.fake_tensor import FakeTensorMode
    from composable_kernel_b200 import torch_ops  # noqa: F401  (registers torch.ops.ckb.*)
    with FakeTensorMode():
        x = torch.empty(2, 3, 4096, dtype=torch.bfloat16)
        packed = torch.empty(packing.packed_bytes(11008, 4096), dtype=torch.uint8)
        sc = torch.empty(32, 11008, dtype=torch.bfloat16)
        y = torch.ops.ckb.w4a16_linear(x, packed, sc, None, 11008, 4096, False)
        assert tuple(y.shape) == (2, 3, 11008) and y.dtype == torch.bfloat16
        w = torch.empty(14336, 4096, dtype=torch.int8)
        y8 = torch.ops.ckb.w8a16_linear(x, w, torch.empty(14336, dtype=torch.bfloat16), torch.empty(14336, dtype=torch.bfloat16))
        assert tuple(y8.shape) == (2, 3, 14336)
        a = torch.empty(64, 256, dtype=torch.int8)
        e = torch.ops.ckb.int8_gemm_scaled(a, torch.empty(128, 256, dtype=torch.int8), torch.empty(128), torch.empty(64))
        assert tuple(e.shape) == (64, 128) and e.dtype == torch.bfloat16


# ------------------------------------------------------------------ ASan + UBSan over the oracle and the header-only host layer
def test_oracle_and_host_layer_under_asan_ubsan():
    """`make -C oracle sanitize` builds oracle/_san/sanitize_main with -fsanitize=address,undefined (the reference's sanitizer build
    switch, CMakeLists.txt:111-118): every oracle weight format / epilogue on ragged shapes, strided and broadcast operands, row
    sampling, threads; Tensor / functors / conversions of include/ck.  Any report aborts the binary (non-zero exit)."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run(["make", "-C", os.path.join(root, "oracle"), "sanitize", "-s"], capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("no sanitizer runtime for the host compiler here: " + r.stderr[-200:])
    r = subprocess.run([os.path.join(root, "oracle", "_san", "sanitize_main")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "sanitize_main: OK" in r.stdout, (r.stdout[-1500:], r.stderr[-3000:])


def test_int8_scale_fma_identity():
    """The device conversion of the int8 `Multiply` weight (csrc/dequant.cuh, s8_biased_byte_times_scale): one FFMA per code,
    fma(2^15 + q + 128, s, -(2^15 + 128) s), must equal float(q) * float(s) bit for bit (binary_element_wise_operation.hpp:184-190
    multiplies in fp32) for every code and every finite bf16 scale below 2^112; zeros may differ in sign only."""
    bits = np.arange(65536, dtype=np.uint32)
    s = (bits << 16).view(np.float32)
    s = s[np.isfinite(s) & (np.abs(s) < 2.0 ** 112)]
    q = np.arange(-128, 128, dtype=np.int32)
    xf = (np.uint32(0x47000000) | ((q + 128).astype(np.uint32) << 8)).view(np.float32)
    assert np.array_equal(xf, (32768 + q + 128).astype(np.float32))
    with np.errstate(all="ignore"):
        nb = (s * np.float32(-32896.0)).astype(np.float32)
        assert np.array_equal(nb.astype(np.float64), s.astype(np.float64) * -32896.0)  # the addend is exact
        ref = q.astype(np.float32)[None, :] * s[:, None]
        # float64 holds the product (24 + 8 bits) and the sum exactly: one rounding to fp32 = the FFMA
        new = (xf.astype(np.float64)[None, :] * s.astype(np.float64)[:, None] + nb.astype(np.float64)[:, None]).astype(np.float32)
    same = ref.view(np.uint32) == new.view(np.uint32)
    both_zero = (ref == 0) & (new == 0)
    assert np.all(same | both_zero)
