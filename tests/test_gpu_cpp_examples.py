"""GPU acceptance tests: the C++ ports of the reference's self-checking examples (tests/cpp), built against the CK-style
headers in include/ck and linked to libckb.so.  Each binary verifies itself against the CK-style host reference operators
(oracle/include) with ck::utils::check_err and exits non-zero on mismatch -- the same way the reference pins this path."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
BIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "bin")

CASES = [
    ("example_fp16int8_gemm", ["1", "1", "0", "256", "512", "384"]),            # example 64, integer init: exact
    ("example_fp16int8_gemm", ["1", "2", "0", "64", "256", "1024"]),            # real-valued init
    ("client_gemm_bias_bf16_i8", ["64", "512", "256", "1", "1", "add"]),        # client 30, mk_nk_mn
    ("client_gemm_bias_bf16_i8", ["48", "256", "384", "0", "1", "add_gelu"]),   # mk_kn_mn (row-major B) + AddFastGelu
    ("example_gemm_quantization_int8", ["0", "256", "512", "384"]),
    ("example_gemm_quantization_int8", ["1", "256", "512", "384"]),
    ("example_gemm_quantization_int8", ["2", "384", "256", "512"]),
    ("example_gemm_quantization_int8", ["3", "1024", "1024", "1024"]),
    ("example_w4a16_group_gemm", ["16", "512", "1024", "1", "0"]),
    ("example_w4a16_group_gemm", ["1", "1024", "512", "0", "0"]),
]


@pytest.mark.parametrize("exe,args", CASES, ids=[f"{e}-{'_'.join(a)}" for e, a in CASES])
def test_cpp_example(exe, args):
    path = os.path.join(BIN, exe)
    assert os.path.exists(path), f"{path} not built (run __graft_entry__.build())"
    r = subprocess.run([path] + args, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert "PASSED" in r.stdout
