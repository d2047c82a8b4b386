"""ctypes view of include/ckb.h (the C-ABI of libckb.so).

This is plumbing for the Python tests / bench: device memory comes from torch, every GEMM goes
through the same `ckb_gemm_query` / `ckb_gemm_run` entry points the C++ CK-style classes in
include/ck/ use.  There is no fallback: if the library is missing or the device is not sm_100 the
calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CKB_LIB selects a debug build of the same library (e.g. lib/libckb_trace.so); never a different implementation
LIB_PATH = os.environ.get("CKB_LIB") or os.path.join(_HERE, "lib", "libckb.so")

# enums (include/ckb.h)
F32, F16, BF16, I8, U8, I32, I4 = range(7)
ROW, COL, PREPACKED = 0, 1, 2
BOP_PASSTHROUGH, BOP_MULTIPLY, BOP_FPAINTB, BOP_GROUP, BOP_GROUP_RTN = range(5)
FLAG_BF16_RNE, FLAG_FIXED_K_ORDER = 1, 2
FLAG_SS_1CTA, FLAG_SS_2CTA, FLAG_PREFILL_TWO_PASS, FLAG_GEMV_SINGLE_PRODUCER, FLAG_INPUTS_READY, FLAG_W8_STREAMK = 4, 8, 16, 32, 64, 128
(CDE_PASSTHROUGH, CDE_ADD, CDE_FASTGELU, CDE_ADD_FASTGELU, CDE_MULTIPLY, CDE_MULTIPLY_ADD,
 CDE_MULTIPLY_FASTGELU, CDE_MULTIPLY_ADD_FASTGELU, CDE_MULTIPLY_MULTIPLY, CDE_ACT_MUL_CLAMP,
 CDE_ADD_ACT_MUL_CLAMP, CDE_ACT_MUL2_CLAMP, CDE_ADD_ACT_MUL2_CLAMP) = range(13)
ACT_PASSTHROUGH, ACT_RELU = 0, 1
FLAG_BF16_RNE = 1
OK, ERR_UNSUPPORTED, ERR_INVALID, ERR_NO_DEVICE, ERR_CUDA, ERR_WORKSPACE = 0, -1, -2, -3, -4, -5
MAX_D = 2
MAX_PEERS = 8


class GemmProblem(C.Structure):
    _fields_ = [
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("a_dtype", C.c_int32), ("a_layout", C.c_int32), ("a", C.c_void_p), ("lda", C.c_int64),
        ("b_dtype", C.c_int32), ("b_layout", C.c_int32), ("b", C.c_void_p), ("ldb", C.c_int64),
        ("b_op", C.c_int32),
        ("s_dtype", C.c_int32), ("scale", C.c_void_p),
        ("zero", C.c_void_p), ("zero_const", C.c_int32), ("group_size", C.c_int32),
        ("num_d", C.c_int32), ("d_dtype", C.c_int32 * MAX_D), ("d", C.c_void_p * MAX_D),
        ("d_sm", C.c_int64 * MAX_D), ("d_sn", C.c_int64 * MAX_D),
        ("cde_op", C.c_int32), ("cde_scale", C.c_float), ("act", C.c_int32),
        ("e_dtype", C.c_int32), ("e", C.c_void_p), ("lde", C.c_int64),
        ("kbatch", C.c_int32), ("flags", C.c_uint32),
        ("num_peers", C.c_int32), ("peer_rank", C.c_int32), ("e_peer", C.c_void_p * MAX_PEERS), ("flag_peer", C.c_void_p * MAX_PEERS),
    ]


class GemmPlan(C.Structure):
    _fields_ = [
        ("supported", C.c_int32), ("kernel_id", C.c_int32), ("kernel_name", C.c_char * 96),
        ("workspace_bytes", C.c_uint64),
        ("grid", C.c_int32), ("block", C.c_int32), ("smem_bytes", C.c_int32), ("split_units", C.c_int32),
        ("algo_bytes", C.c_uint64), ("flops", C.c_uint64),
        ("_pad", C.c_ubyte * 48),            # opaque is 64-byte aligned in C
        ("opaque", C.c_ubyte * 1280),
    ]


_lib = None


class CkbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"ckb error {code}: {msg}")
        self.code = code


def lib():
    """Load libckb.so; fail loudly when it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "or `make -C composable_kernel_b200`")
        l = C.CDLL(LIB_PATH)
        l.ckb_last_error.restype = C.c_char_p
        l.ckb_gemm_query.argtypes = [C.POINTER(GemmProblem), C.POINTER(GemmPlan)]
        l.ckb_gemm_run.argtypes = [C.POINTER(GemmProblem), C.POINTER(GemmPlan), C.c_void_p, C.c_void_p]
        l.ckb_gemm_host_staging_bytes.argtypes = [C.POINTER(GemmProblem)]
        l.ckb_gemm_host_staging_bytes.restype = C.c_uint64
        l.ckb_gemm_run_host.argtypes = [C.POINTER(GemmProblem), C.c_void_p, C.c_void_p, C.c_void_p]
        l.ckb_gemm_run_host_async.argtypes = [C.POINTER(GemmProblem), C.c_void_p, C.c_void_p, C.c_void_p]
        l.ckb_prepack_b_u8.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        l.ckb_prepack_b_rowmajor_i8.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
        l.ckb_prepack_b_int4_bytes.argtypes = [C.c_int32, C.c_int32]
        l.ckb_prepack_b_int4_bytes.restype = C.c_uint64
        l.ckb_prepack_b_int4.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                         C.c_void_p]
        if os.environ.get("CKB_LIB") and not hasattr(l, "ckb_tile_coord"):
            pass  # an older debug build selected for an A/B measurement: the newer entry points are simply absent
        else:
            l.ckb_tile_coord.argtypes = [C.c_int32] * 4 + [C.POINTER(C.c_int32)] * 2
            l.ckb_streamk_range.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
            l.ckb_gemv_schedule.argtypes = [C.c_int32] * 6 + [C.POINTER(C.c_int32)]
            l.ckb_transpose_a_colmajor.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
            l.ckb_peer_wait.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        if hasattr(l, "ckb_gemm_run_static_scratch"):   # absent from older experiment builds selected with CKB_LIB
            l.ckb_static_scratch_bytes.argtypes = []
            l.ckb_static_scratch_bytes.restype = C.c_uint64
            l.ckb_gemm_run_static_scratch.argtypes = [C.POINTER(GemmProblem), C.POINTER(GemmPlan), C.c_void_p]
        l.ckb_grouped_gemm_arg_bytes.argtypes = [C.c_int32, C.c_int32]
        l.ckb_grouped_gemm_arg_bytes.restype = C.c_uint64
        l.ckb_grouped_gemm_query.argtypes = [C.POINTER(GemmProblem), C.c_int32]
        l.ckb_grouped_gemm_run.argtypes = [C.POINTER(GemmProblem), C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
        _lib = l
    return _lib


EXPORTED = [
    "ckb_gemm_query", "ckb_gemm_run", "ckb_gemm_host_staging_bytes", "ckb_gemm_run_host", "ckb_gemm_run_host_async",
    "ckb_prepack_b_u8", "ckb_prepack_b_rowmajor_i8", "ckb_prepack_b_int4_bytes", "ckb_prepack_b_int4", "ckb_grouped_gemm_run", "ckb_grouped_gemm_query", "ckb_grouped_gemm_arg_bytes",
    "ckb_last_error", "ckb_version", "ckb_device_sm_count", "ckb_peer_wait",
    "ckb_tile_coord", "ckb_streamk_range", "ckb_gemv_schedule", "ckb_transpose_a_colmajor",
    "ckb_static_scratch_bytes", "ckb_gemm_run_static_scratch",
]


def check(rc):
    if rc != 0:
        raise CkbError(rc, lib().ckb_last_error().decode())


def last_error() -> str:
    return lib().ckb_last_error().decode()
