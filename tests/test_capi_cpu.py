"""CPU-side checks of the drop-in boundary (no GPU, no compute calls): the C-ABI library loads, exports
every symbol include/ckb.h declares, its POD layouts match the ctypes mirror, and argument validation
(CK's IsSupportedArgument) behaves like the reference's error conventions."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

from composable_kernel_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "ckb.h")


def declared_functions():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ckb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = capi.lib()
    names = declared_functions()
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), f"libckb.so does not export {n}"
    assert set(names) == set(capi.EXPORTED)
    assert lib.ckb_version() == 100


def test_struct_layouts_match_header(tmp_path):
    prog = tmp_path / "layout.c"
    prog.write_text('#include "ckb.h"\n#include <stdio.h>\n#include <stddef.h>\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                    "sizeof(ckb_gemm_problem),offsetof(ckb_gemm_problem,scale),offsetof(ckb_gemm_problem,d),"
                    "offsetof(ckb_gemm_problem,e),sizeof(ckb_gemm_plan),offsetof(ckb_gemm_plan,algo_bytes),"
                    "offsetof(ckb_gemm_plan,opaque));return 0;}\n")
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])  # plain C
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    P, L = capi.GemmProblem, capi.GemmPlan
    assert got == [C.sizeof(P), P.scale.offset, P.d.offset, P.e.offset, C.sizeof(L), L.algo_bytes.offset, L.opaque.offset]


def _problem(**kw):
    p = capi.GemmProblem()
    p.M, p.N, p.K = 16, 256, 256
    p.a_dtype, p.a_layout, p.a, p.lda = capi.BF16, capi.ROW, 0x1000, 256
    p.b_dtype, p.b_layout, p.b, p.ldb = capi.I8, capi.COL, 0x2000, 256
    p.b_op, p.s_dtype, p.scale = capi.BOP_MULTIPLY, capi.BF16, 0x3000
    p.e_dtype, p.e, p.lde = capi.BF16, 0x4000, 256
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def _query(p):
    plan = capi.GemmPlan()
    rc = capi.lib().ckb_gemm_query(C.byref(p), C.byref(plan))
    return rc, plan, capi.last_error()


@pytest.mark.parametrize("kw,needle", [
    (dict(K=0), "K must be positive"),
    (dict(a_layout=capi.COL), "row-major"),
    (dict(b_layout=capi.ROW), "K-major"),
    (dict(a_dtype=capi.F16), "Multiply B-op"),
    (dict(lda=100), "lda"),
    (dict(cde_op=capi.CDE_ADD), "D tensors"),
    (dict(cde_op=capi.CDE_ACT_MUL_CLAMP), "int8 inputs"),
    (dict(b_op=capi.BOP_GROUP), "pre-packed int4"),
    (dict(e_dtype=capi.I8), "E must be"),
])
def test_unsupported_arguments_are_reported_not_raised(kw, needle):
    rc, plan, msg = _query(_problem(**kw))
    assert rc == capi.OK and plan.supported == 0, (rc, msg)
    assert needle in msg
    # Run on an unsupported plan = the reference's `throw std::runtime_error` (device_fpAintB_gemm_wmma.hpp:426-433)
    assert capi.lib().ckb_gemm_run(C.byref(_problem(**kw)), C.byref(plan), None, None) == capi.ERR_UNSUPPORTED


def test_invalid_arguments():
    lib = capi.lib()
    assert lib.ckb_gemm_query(None, None) == capi.ERR_INVALID
    rc, _, _ = _query(_problem(M=-1))
    assert rc == capi.ERR_INVALID


def test_empty_problem_is_supported_without_a_device():
    rc, plan, _ = _query(_problem(M=0))
    assert rc == capi.OK and plan.supported == 1 and plan.kernel_name == b"empty"


def test_no_cpu_fallback():
    """On a box without an sm_100 device a well-formed problem must FAIL loudly, never compute."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    rc, plan, msg = _query(_problem())
    assert rc == capi.ERR_NO_DEVICE and plan.supported == 0 and "sm_100" in msg
    assert capi.lib().ckb_device_sm_count() < 0
    assert capi.lib().ckb_prepack_b_u8(0x1000, 0x2000, 16, None) == capi.ERR_NO_DEVICE


def test_algorithmic_byte_and_flop_accounting():
    # cfg2 (N,K) = (4096,4096), M = 1/4/16 -> 8,667,136 / 8,716,288 / 8,912,896 bytes
    for M, want in [(1, 8667136), (4, 8716288), (16, 8912896)]:
        p = _problem(M=M, N=4096, K=4096, lda=4096, lde=4096, b_dtype=capi.I4, b_layout=capi.PREPACKED,
                     b_op=capi.BOP_GROUP, group_size=128)
        rc, plan, _ = _query(p)
        assert plan.algo_bytes == want and plan.flops == 2 * M * 4096 * 4096
    assert capi.lib().ckb_prepack_b_int4_bytes(4096, 4096) == 4096 * 4096 // 2
    assert capi.lib().ckb_prepack_b_int4_bytes(1000, 130) == 63 * 3 * 512  # padded to whole 16-channel x 64-k fragments blocks


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under composable_kernel_b200/ or include/ may reference it."""
    bad = []
    for base in ("composable_kernel_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            if "build" in dp:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"ck_oracle|oracle/|import oracle|from oracle", txt):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_grouped_argument_layout_and_query_without_gpu():
    """GroupedGemmMultiABDKernelArgument<1,NumB,NumD> sizes (pointers, then index_t fields, padded to 8) and the host-side
    IsSupportedArgument checks of the grouped op need no device."""
    import ctypes as C
    from composable_kernel_b200 import capi
    lib = capi.lib()
    assert lib.ckb_grouped_gemm_arg_bytes(2, 1) == 72 == C.sizeof(C.c_void_p) * 5 + 4 * 8
    assert lib.ckb_grouped_gemm_arg_bytes(2, 0) == 72   # a zero-length std::array still occupies storage
    assert lib.ckb_grouped_gemm_arg_bytes(1, 0) == 64
    assert lib.ckb_grouped_gemm_arg_bytes(1, 1) == 64
    assert lib.ckb_grouped_gemm_arg_bytes(3, 0) == 0
    p = capi.GemmProblem()
    p.N, p.K, p.lda, p.ldb, p.lde = 1024, 512, 512, 1024, 1024
    p.a_dtype, p.a_layout, p.b_dtype, p.b_layout, p.e_dtype = capi.BF16, capi.ROW, capi.I8, capi.ROW, capi.BF16
    p.b_op, p.s_dtype, p.cde_op, p.num_d = capi.BOP_MULTIPLY, capi.BF16, capi.CDE_ADD_FASTGELU, 1
    p.d_dtype[0] = capi.BF16
    assert lib.ckb_grouped_gemm_query(C.byref(p), 16) == capi.OK
    assert lib.ckb_grouped_gemm_query(C.byref(p), 0) == capi.ERR_UNSUPPORTED
    p.K = 520
    assert lib.ckb_grouped_gemm_query(C.byref(p), 16) == capi.ERR_UNSUPPORTED and "multiples of 16" in capi.last_error()
    p.K, p.a_layout = 512, capi.COL
    p.lda = 64
    assert lib.ckb_grouped_gemm_query(C.byref(p), 16) == capi.OK                  # km_kn_mn: column-major A, pitch >= M
    p.a_layout = capi.PREPACKED
    assert lib.ckb_grouped_gemm_query(C.byref(p), 16) == capi.ERR_UNSUPPORTED
    p.a_layout, p.num_d = capi.ROW, 0
    assert lib.ckb_grouped_gemm_query(C.byref(p), 16) == capi.ERR_UNSUPPORTED   # AddFastGelu needs its D tensor


def test_reference_instance_declarations_compile_and_construct():
    """tests/cpp/verbatim_instances.cpp holds the reference's own instance declarations (full template argument lists of
    examples 64 / 14 / 60 and the V3 order of example 65); built by __graft_entry__.build(), it must construct the four
    operators without a GPU and report the reference class names."""
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "bin", "verbatim_instances")
    if not os.path.exists(exe):
        pytest.skip("tests/cpp not built")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr[-1000:]
    for cls in ("DeviceFpAintBGemm_Wmma_CShuffle", "DeviceGemmMultipleD_Xdl_CShuffle", "DeviceGemmMultipleABD_Xdl_CShuffle", "DeviceGemmMultiD_Xdl_CShuffle_V3"):
        assert cls + "<sm_100a" in r.stdout
