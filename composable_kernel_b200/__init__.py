"""composable_kernel_b200 -- B200 (sm_100a) mixed-precision GEMM behind Composable Kernel's operator API.

The product is the C-ABI library `lib/libckb.so` (include/ckb.h) plus the CK-style C++ headers in
include/ck/.  This Python package only loads that library for tests and benchmarks.
"""
from . import capi  # noqa: F401

__all__ = ["capi"]
