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
    ("client_gemm_bias_bf16_i8", ["200", "512", "256", "2", "1", "add"]),       # km_kn_mn: column-major A (prefill rows)
    ("client_gemm_bias_bf16_i8", ["16", "256", "512", "2", "1", "add_gelu"]),   # km_kn_mn, decode rows
    ("example_gemm_quantization_int8", ["0", "256", "512", "384"]),
    ("example_gemm_quantization_int8", ["1", "256", "512", "384"]),
    ("example_gemm_quantization_int8", ["2", "384", "256", "512"]),
    ("example_gemm_quantization_int8", ["3", "1024", "1024", "1024"]),
    ("example_w4a16_group_gemm", ["16", "512", "1024", "1", "0"]),
    ("example_w4a16_group_gemm", ["1", "1024", "512", "0", "0"]),
    ("client_grouped_gemm_bias_fastgelu_bf16_i8", []),                           # client 31: 16 groups, N=1024, K=512, mk_kn_mn
    ("client_grouped_gemm_bias_fastgelu_bf16_i8", ["5", "256", "192", "1", "1"]),  # mk_nk_mn
    ("client_grouped_gemm_bias_fastgelu_bf16_i8", ["6", "256", "256", "0", "1", "64"]),  # km_kn_mn, pitch 64: vector loads along m
    ("client_grouped_gemm_bias_fastgelu_bf16_i8", ["5", "128", "192", "0", "1", "67"]),  # km_kn_mn, odd pitch: element loads
    ("client_grouped_gemm_fastgelu_bf16_i8", ["0"]),                             # client 31 FastGelu: KernelArgument<1,2,0> (NumD = 0)
    ("client_grouped_gemm_fastgelu_bf16_i8", ["1"]),                             # client 31 plain: PassThrough, NumD = 0
    ("client_grouped_gemm_fastgelu_bf16_i8", ["0", "5", "256", "192", "1"]),     # NumD = 0, mk_nk_mn
    ("example_w4a16_group_gemm", ["16", "512", "1024", "1", "0", "1"]),          # DequantizeGroupRound contract, decode rows
    ("example_w4a16_group_gemm", ["200", "512", "1024", "0", "0", "1"]),         # DequantizeGroupRound, prefill rows (tcgen05)
    ("example_w4a16_group_gemm", ["200", "512", "1024", "1", "0", "0"]),         # DequantizeGroup above 64 rows (row slabs)
    ("interfaces_plain", []),                                                    # DeviceGemm / SplitK / StreamK / V3 / ABScale
]

PROFILER = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiler", "bin", "ckProfiler")
# ckProfiler <op> ...: every factory instance verified against the host reference, then timed (reference test strategy:
# the op tests are thin wrappers around the profiler's profile_*_impl functions)
PROFILER_CASES = [
    ["gemm_multiply_multiply", "3", "1", "1", "2", "0", "1", "512", "768", "640", "-1", "-1", "0", "0", "-1"],
    ["gemm_multi_abd", "0", "0", "3", "1", "2", "0", "1", "128", "512", "256", "-1", "-1", "-1"],
    ["gemm_multi_abd", "0", "1", "1", "1", "2", "0", "1", "16", "512", "1024", "-1", "-1", "-1"],
    ["gemm_multi_abd", "0", "1", "0", "1", "1", "0", "0", "300", "256", "128", "-1", "-1", "-1"],
    ["gemm_quantization", "0", "1", "2", "0", "1", "256", "512", "384", "0.03"],
    ["gemm_quantization", "1", "1", "2", "0", "1", "384", "256", "512", "0.01"],
    ["fpAintB_gemm", "1", "1", "0", "1", "128", "256", "512"],
    ["fpAintB_gemm", "1", "2", "0", "1", "8", "512", "1024"],
    ["gemm_w4a16", "1", "2", "0", "1", "4", "1024", "1024", "1"],
    ["gemm_w4a16", "1", "2", "0", "1", "256", "512", "1024", "0"],
    ["grouped_gemm_fixed_nk", "0", "0", "1", "2", "0", "1", "40,63,1,130", "512,512,512,512", "256,256,256,256", "-1", "-1", "-1"],
    ["gemm_universal_reduce", "0", "0", "1", "2", "0", "1", "16", "512", "1024", "-1", "-1", "-1", "4"],     # decode-shaped, KBatch 4
    ["gemm_universal_reduce", "0", "0", "1", "1", "0", "1", "300", "256", "384", "-1", "-1", "-1", "1"],     # dense path, integer init
    ["grouped_gemm_fixed_nk", "0", "1", "1", "1", "0", "1", "17,64", "128,128", "320,320", "-1", "-1", "-1", "1", "2", "5"],
]


@pytest.mark.parametrize("exe,args", CASES, ids=[f"{e}-{'_'.join(a)}" for e, a in CASES])
def test_cpp_example(exe, args):
    path = os.path.join(BIN, exe)
    assert os.path.exists(path), f"{path} not built (run __graft_entry__.build())"
    r = subprocess.run([path] + args, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert "PASSED" in r.stdout


@pytest.mark.parametrize("args", PROFILER_CASES, ids=["_".join(a[:4]) + f"_{i}" for i, a in enumerate(PROFILER_CASES)])
def test_ck_profiler(args):
    assert os.path.exists(PROFILER), f"{PROFILER} not built (run __graft_entry__.build())"
    r = subprocess.run([PROFILER] + args, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert "verification PASSED" in r.stdout
    if args[0] != "gemm_multi_abd" or args[7] == "1":
        assert "Best Perf:" in r.stdout


def test_verbatim_reference_instances_run():
    """the reference's own instance declarations (full template argument lists) compile against include/ck and run"""
    r = subprocess.run([os.path.join(BIN, "verbatim_instances"), "run"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    for cls in ("DeviceFpAintBGemm_Wmma_CShuffle<sm_100a", "DeviceGemmMultipleD_Xdl_CShuffle<sm_100a", "DeviceGemmMultipleABD_Xdl_CShuffle<sm_100a",
                "DeviceGemmMultiD_Xdl_CShuffle_V3<sm_100a"):
        assert cls in r.stdout
    assert "gemm_ss" in r.stdout          # GetTypeString names the chosen kernel once an argument has been made


def test_ck_profiler_rotating_buffers_report_the_hbm_number():
    """with rotating buffers + L2 flush (flush_cache.hpp semantics) a decode-shaped GEMM whose weights fit the
    126 MB L2 must time SLOWER than the warm back-to-back loop, i.e. the profiler reports the HBM number, not the L2 number."""
    import json
    args = ["gemm_w4a16", "0", "2", "0", "1", "1", "11008", "4096", "0", "3", "30"]

    def perf(env):
        r = subprocess.run([PROFILER] + args, capture_output=True, text=True, timeout=300, env={**os.environ, **env})
        assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
        rows = [json.loads(l.split("PerfJSON: ", 1)[1]) for l in r.stdout.splitlines() if l.startswith("PerfJSON: ")]
        assert rows and all(k in rows[0] for k in ("instance", "ms", "gbs", "hbm_frac", "timing"))
        return rows[0]
    warm = perf({"CKP_ROTATING_MB": "0"})
    cold = perf({"CKP_ROTATING_MB": "512"})
    assert "warm" in warm["timing"] and "cold" in cold["timing"]
    assert cold["ms"] > 1.1 * warm["ms"], (warm, cold)


def test_ck_profiler_unknown_operation_and_help():
    r = subprocess.run([PROFILER, "no_such_op"], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "cannot find operation" in r.stderr
    r = subprocess.run([PROFILER], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "gemm_multiply_multiply" in r.stdout and "grouped_gemm_fixed_nk" in r.stdout
