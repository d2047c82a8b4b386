"""GPU bring-up probe: runs each kernel family on small problems in a separate subprocess (with a
timeout, so a hang cannot take the box down) and prints a compact diagnosis.  Not a pytest file.
usage: python tests/gpu_probe.py [case ...]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = {}


def case(fn):
    CASES[fn.__name__] = fn
    return fn


def _report(name, got, ref, exact=False, rtol=1e-2, atol=1e-2):
    import numpy as np
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    diff = np.abs(got - ref)
    bad = diff > (0 if exact else atol + rtol * np.abs(ref))
    print(f"[{name}] shape={got.shape} max|diff|={diff.max():.6g} bad={int(bad.sum())}/{bad.size} "
          f"ref_absmax={np.abs(ref).max():.4g} got_absmax={np.abs(got).max():.4g}")
    if bad.any():
        idx = np.argwhere(bad)
        print("   first bad idx:", idx[:8].tolist())
        rows = np.unique(idx[:, 0])[:16].tolist()
        cols = np.unique(idx[:, 1])[:32].tolist()
        print("   bad rows:", rows, " bad cols:", cols)
        i, j = idx[0]
        print(f"   got[{i},{j}]={got[i, j]} ref={ref[i, j]}")
        return False
    return True


def _i8_case(M, N, K, name):
    import numpy as np
    import torch
    from composable_kernel_b200 import capi, ops
    rng = np.random.default_rng(1)
    a = rng.integers(-128, 128, (M, K)).astype(np.int8)
    b = rng.integers(-128, 128, (N, K)).astype(np.int8)
    ref = a.astype(np.int64) @ b.astype(np.int64).T
    ad, bd = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    e = torch.zeros(M, N, dtype=torch.int32, device="cuda")
    op = ops.GemmOp(ad, bd, e)
    print("   kernel:", op.kernel_name, "supported:", op.supported, op.why)
    op.run()
    torch.cuda.synchronize()
    return _report(name, e.cpu().numpy(), ref, exact=True)


@case
def ss_i8_one_tile():
    return _i8_case(128, 256, 128, "ss_i8 128x256x128")


@case
def ss_i8_multi_k():
    return _i8_case(128, 256, 1024, "ss_i8 128x256x1024")


@case
def ss_i8_multi_tile():
    return _i8_case(640, 1280, 512, "ss_i8 640x1280x512")


@case
def ss_i8_ragged():
    return _i8_case(200, 300, 144, "ss_i8 200x300x144")


@case
def ss_bf16():
    import numpy as np
    import torch
    from composable_kernel_b200 import ops
    torch.manual_seed(0)
    M, N, K = 256, 512, 256
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    e = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    op = ops.GemmOp(a, b, e)
    print("   kernel:", op.kernel_name)
    op.run()
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    return _report("ss_bf16", e.float().cpu().numpy(), ref.cpu().numpy())


def _w4_case(M, N, K, name, with_zero=False, f16=False, onehot=False):
    import numpy as np
    import torch
    from composable_kernel_b200 import capi, ops
    from oracle import ck_oracle as orc
    rng = np.random.default_rng(2)
    G = 128
    q = rng.integers(0, 16, (K, N)).astype(np.uint8)
    if f16:
        a = (rng.standard_normal((M, K))).astype(np.float16)
        s = (rng.random((K // G, N)) * 0.015 + 0.005).astype(np.float16)
        a_np, s_np, adt, tdt = a, s, orc.F16, torch.float16
    else:
        a_np = orc.bf16_trunc(rng.standard_normal((M, K)).astype(np.float32))
        s_np = orc.bf16_trunc((rng.random((K // G, N)) * 0.015 + 0.005).astype(np.float32))
        adt, tdt = orc.BF16, torch.bfloat16
    if onehot:
        a_np[:] = 0
        for m in range(M):
            a_np[m, (m * 37) % K] = np.float16(1.0) if f16 else 0x3F80
    z = rng.integers(0, 16, (K // G, N)).astype(np.uint8) if with_zero else None
    ref = orc.gemm(a_np, adt, q, orc.U8, orc.F32, b_op=orc.BOP_GROUP_F32, scale=s_np, s_dtype=adt, zero=z,
                   zero_const=8, group_size=G)
    to_t = lambda x: (torch.from_numpy(x).cuda() if f16 else torch.from_numpy(x.view(np.int16)).cuda().view(torch.bfloat16))
    ad, sd = to_t(a_np), to_t(s_np)
    qd = torch.from_numpy(q).cuda()
    packed = ops.prepack_int4(qd)
    zd = torch.from_numpy(z).cuda() if with_zero else None
    e = torch.zeros(M, N, dtype=torch.float32, device="cuda")
    op = ops.GemmOp(ad, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=N, K=K,
                    scale=sd, zero=zd, zero_const=8, group_size=G)
    print("   kernel:", op.kernel_name, "grid", op.plan.grid, "smem", op.plan.smem_bytes, "supported:", op.supported, op.why)
    op.run()
    torch.cuda.synchronize()
    ok = _report(name, e.cpu().numpy(), ref, rtol=2e-3, atol=2e-3)
    op.run()
    torch.cuda.synchronize()
    ok2 = _report(name + " (2nd run: workspace restored?)", e.cpu().numpy(), ref, rtol=2e-3, atol=2e-3)
    return ok and ok2


@case
def w4_onehot_one_unit():
    return _w4_case(16, 128, 128, "w4 onehot 16x128x128", onehot=True)


@case
def w4_one_unit():
    return _w4_case(16, 128, 128, "w4 16x128x128")


@case
def w4_multi_k():
    return _w4_case(4, 128, 1024, "w4 4x128x1024")


@case
def w4_streamk():
    return _w4_case(16, 4096, 4096, "w4 16x4096x4096", with_zero=True)


@case
def w4_m1_f16():
    return _w4_case(1, 1024, 512, "w4 f16 1x1024x512", f16=True)


@case
def w4_m64():
    return _w4_case(48, 512, 1024, "w4 48x512x1024")


def _w8_case(M, N, K, mode, name):
    import numpy as np
    import torch
    from composable_kernel_b200 import capi, ops
    from oracle import ck_oracle as orc
    rng = np.random.default_rng(3)
    qn = rng.integers(-128, 128, (N, K)).astype(np.int8)      # [N,K] storage, K contiguous (Col-major B)
    q_kn = qn.T
    if mode == "fpaintb":
        a = rng.standard_normal((M, K)).astype(np.float16)
        s = (rng.random(N) * 0.02 + 0.001).astype(np.float16)
        ref = orc.gemm(a, orc.F16, q_kn, orc.I8, orc.F16, b_op=orc.BOP_FPAINTB, scale=np.broadcast_to(s, (K, N)),
                       s_dtype=orc.F16).astype(np.float32)
        ad, sd = torch.from_numpy(a).cuda(), torch.from_numpy(s).cuda()
        bd = ops.prepack_u8(torch.from_numpy(qn).cuda())
        e = torch.zeros(M, N, dtype=torch.float16, device="cuda")
        op = ops.GemmOp(ad, bd, e, b_op=capi.BOP_FPAINTB, scale=sd)
    else:
        a = orc.bf16_trunc(rng.standard_normal((M, K)).astype(np.float32))
        s = orc.bf16_trunc((rng.random(N) * 0.02 + 0.001).astype(np.float32))
        bias = orc.bf16_trunc(rng.standard_normal(N).astype(np.float32))
        bop = orc.BOP_MULTIPLY if mode == "mul" else orc.BOP_PASSTHROUGH
        ref = orc.bf16_to_f32(orc.gemm(a, orc.BF16, q_kn, orc.I8, orc.BF16, b_op=bop,
                                       scale=np.broadcast_to(s, (K, N)), s_dtype=orc.BF16,
                                       ds=(np.broadcast_to(bias, (M, N)),), d_dtypes=(orc.BF16,), cde_op=orc.CDE_ADD))
        tb = lambda x: torch.from_numpy(x.view(np.int16)).cuda().view(torch.bfloat16)
        ad, sd, biasd = tb(a), tb(s), tb(bias)
        bd = torch.from_numpy(qn).cuda()
        e = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
        op = ops.GemmOp(ad, bd, e, b_op=capi.BOP_MULTIPLY if mode == "mul" else capi.BOP_PASSTHROUGH,
                        scale=sd if mode == "mul" else None, ds=(biasd,), d_strides=((0, 1),), cde_op=capi.CDE_ADD)
    print("   kernel:", op.kernel_name, "grid", op.plan.grid, "supported:", op.supported, op.why)
    op.run()
    torch.cuda.synchronize()
    return _report(name, e.float().cpu().numpy(), ref)


@case
def w8_fpaintb_small():
    return _w8_case(16, 256, 512, "fpaintb", "w8 fpaintb 16x256x512")


@case
def w8_mul_small():
    return _w8_case(8, 256, 512, "mul", "w8 mul+bias 8x256x512")


@case
def w8_pass_small():
    return _w8_case(8, 128, 256, "pass", "w8 pass+bias 8x128x256")


@case
def w8_mul_m200():
    return _w8_case(200, 384, 512, "mul", "w8 mul+bias 200x384x512 (slabs)")


def main():
    names = sys.argv[1:] or list(CASES)
    if len(names) == 1 and os.environ.get("CKB_PROBE_CHILD") == "1":
        ok = CASES[names[0]]()
        sys.exit(0 if ok else 1)
    summary = {}
    for n in names:
        env = dict(os.environ, CKB_PROBE_CHILD="1")
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), n], env=env, timeout=120,
                               capture_output=True, text=True)
            out = (r.stdout + r.stderr).strip().splitlines()
            print(f"=== {n}: rc={r.returncode}")
            for line in out[-25:]:
                print("   ", line)
            summary[n] = "PASS" if r.returncode == 0 else f"FAIL({r.returncode})"
        except subprocess.TimeoutExpired:
            print(f"=== {n}: TIMEOUT")
            summary[n] = "TIMEOUT"
    print("SUMMARY", summary)


if __name__ == "__main__":
    main()
