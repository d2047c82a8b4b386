"""ctypes binding of the CPU oracle (TEST INFRASTRUCTURE -- never imported by the product).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use this.
See oracle/ck_oracle.h for what it restates (reference file:line) and the parity-pin status.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libck_oracle.so")

F32, F16, BF16, I8, U8, I32 = 0, 1, 2, 3, 4, 5
BOP_PASSTHROUGH, BOP_MULTIPLY, BOP_FPAINTB, BOP_GROUP_F32, BOP_GROUP_RTN = 0, 1, 2, 3, 4
(CDE_PASSTHROUGH, CDE_ADD, CDE_FASTGELU, CDE_ADD_FASTGELU, CDE_MULTIPLY, CDE_MULTIPLY_ADD,
 CDE_MULTIPLY_FASTGELU, CDE_MULTIPLY_ADD_FASTGELU, CDE_MULTIPLY_MULTIPLY, CDE_ACT_MUL_CLAMP,
 CDE_ADD_ACT_MUL_CLAMP, CDE_ACT_MUL2_CLAMP, CDE_ADD_ACT_MUL2_CLAMP) = range(13)
ACT_PASSTHROUGH, ACT_RELU = 0, 1

_NP_OF = {F32: np.float32, F16: np.float16, BF16: np.uint16, I8: np.int8, U8: np.uint8, I32: np.int32}


class OrcProblem(C.Structure):
    _fields_ = [
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("a_dtype", C.c_int32), ("a", C.c_void_p), ("a_sm", C.c_int64), ("a_sk", C.c_int64),
        ("b_dtype", C.c_int32), ("b", C.c_void_p), ("b_sk", C.c_int64), ("b_sn", C.c_int64),
        ("b_op", C.c_int32),
        ("s_dtype", C.c_int32), ("scale", C.c_void_p), ("s_sk", C.c_int64), ("s_sn", C.c_int64),
        ("zero", C.c_void_p), ("z_sk", C.c_int64), ("z_sn", C.c_int64),
        ("zero_const", C.c_int32), ("group_size", C.c_int32),
        ("cshuffle_dtype", C.c_int32),
        ("num_d", C.c_int32), ("d_dtype", C.c_int32 * 2), ("d", C.c_void_p * 2),
        ("d_sm", C.c_int64 * 2), ("d_sn", C.c_int64 * 2),
        ("cde_op", C.c_int32), ("cde_scale", C.c_float), ("act", C.c_int32),
        ("e_dtype", C.c_int32), ("e", C.c_void_p), ("e_sm", C.c_int64), ("e_sn", C.c_int64),
        ("m_begin", C.c_int32), ("m_end", C.c_int32), ("num_threads", C.c_int32),
    ]


def build(force: bool = False) -> str:
    """Compile oracle/libck_oracle.so with the committed Makefile (g++ only)."""
    src = os.path.join(_HERE, "ck_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_gemm.argtypes = [C.POINTER(OrcProblem)]
        _lib.orc_gemm.restype = C.c_int
        _lib.orc_last_error.restype = C.c_char_p
        for name, res, args in [
            ("orc_f32_to_bf16_trunc", C.c_uint16, [C.c_float]),
            ("orc_f32_to_bf16_rtn", C.c_uint16, [C.c_float]),
            ("orc_bf16_to_f32", C.c_float, [C.c_uint16]),
            ("orc_f32_to_f16", C.c_uint16, [C.c_float]),
            ("orc_f16_to_f32", C.c_float, [C.c_uint16]),
            ("orc_i8_to_bf16", C.c_uint16, [C.c_int8]),
            ("orc_bf16_to_i8", C.c_int8, [C.c_uint16]),
            ("orc_fast_gelu", C.c_float, [C.c_float]),
            ("orc_hardware_threads", C.c_int, []),
        ]:
            f = getattr(_lib, name)
            f.restype, f.argtypes = res, args
        _lib.orc_fill_uniform_f32.argtypes = [C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_int]
        _lib.orc_int8_to_biased_u8.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    return _lib


# ---- bf16 helpers on numpy uint16 bit patterns (ck::bhalf_t = ushort) ---------------------------

def bf16_trunc(x: np.ndarray) -> np.ndarray:
    """fp32 -> bf16 bit pattern by truncation (type_convert.hpp:53-64)."""
    return (np.ascontiguousarray(x, dtype=np.float32).view(np.uint32) >> 16).astype(np.uint16)


def bf16_rtn(x: np.ndarray) -> np.ndarray:
    """fp32 -> bf16 bit pattern, round-to-nearest-even (type_convert.hpp:508-549), finite inputs."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = u + 0x7FFF + ((u >> 16) & 1)
    return (u >> 16).astype(np.uint16)


def bf16_to_f32(x: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(x, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)


def _strides_elems(a: np.ndarray):
    return tuple(s // a.itemsize for s in a.strides)


def fill_uniform(n: int, a: float, b: float, integer: bool = False) -> np.ndarray:
    """FillUniformDistribution[IntegerValue]{a,b} of fill.hpp (fresh mt19937(11939) each call)."""
    out = np.empty(n, dtype=np.float32)
    lib().orc_fill_uniform_f32(out.ctypes.data, n, a, b, int(integer))
    return out


def gemm(a, a_dtype, b, b_dtype, e_dtype, *, b_op=BOP_PASSTHROUGH, scale=None, s_dtype=F32,
         zero=None, zero_const=0, group_size=0, cshuffle_dtype=F32, ds=(), d_dtypes=(),
         cde_op=CDE_PASSTHROUGH, cde_scale=1.0, act=ACT_PASSTHROUGH, m_range=None, num_threads=0,
         e_out=None):
    """Run the oracle.  `a` is a 2-D numpy array indexed [m,k]; `b` is indexed [k,n]; `scale`,
    `zero` are indexed [k or k//group, n]; each d is indexed [m,n].  Arrays may be arbitrary strided
    views (broadcast_to gives stride-0 tensors like CK's StrideB1 = 0).  bf16 data are uint16 bit
    patterns.  Returns E as a numpy array [M,N] of the dtype's storage type."""
    M, K = a.shape
    K2, N = b.shape
    assert K == K2
    p = OrcProblem()
    p.M, p.N, p.K = M, N, K
    p.a_dtype, p.a = a_dtype, a.ctypes.data
    p.a_sm, p.a_sk = _strides_elems(a)
    p.b_dtype, p.b = b_dtype, b.ctypes.data
    p.b_sk, p.b_sn = _strides_elems(b)
    p.b_op = b_op
    keep = [a, b]
    if scale is not None:
        p.s_dtype, p.scale = s_dtype, scale.ctypes.data
        p.s_sk, p.s_sn = _strides_elems(scale)
        keep.append(scale)
    if zero is not None:
        assert zero.dtype == np.uint8
        p.zero = zero.ctypes.data
        p.z_sk, p.z_sn = _strides_elems(zero)
        keep.append(zero)
    p.zero_const, p.group_size = zero_const, group_size
    p.cshuffle_dtype = cshuffle_dtype
    p.num_d = len(ds)
    for i, d in enumerate(ds):
        p.d_dtype[i] = d_dtypes[i]
        p.d[i] = d.ctypes.data
        p.d_sm[i], p.d_sn[i] = _strides_elems(d)
        keep.append(d)
    p.cde_op, p.cde_scale, p.act = cde_op, cde_scale, act
    e = e_out if e_out is not None else np.zeros((M, N), dtype=_NP_OF[e_dtype])
    p.e_dtype, p.e = e_dtype, e.ctypes.data
    p.e_sm, p.e_sn = _strides_elems(e)
    if m_range is not None:
        p.m_begin, p.m_end = m_range
    p.num_threads = num_threads
    rc = lib().orc_gemm(C.byref(p))
    if rc != 0:
        raise ValueError("oracle: " + lib().orc_last_error().decode())
    return e
