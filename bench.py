#!/usr/bin/env python
"""bench.py -- headline benchmark: Llama-2-7B decode, W4A16 group-128 GEMMs (BASELINE.json configs[1]).

One STEP = one decode pass over all 32 layers' weight GEMMs at each M in {1,4,16}:
  per layer 4 x (N=4096,K=4096) [q,k,v,o], 2 x (N=11008,K=4096) [gate,up], 1 x (N=4096,K=11008) [down]
  -> 224 GEMMs per M, 672 launches per step, every GEMM on its own layer's weights (3.4 GB of packed
  int4 + scales per rank: far larger than the 126 MB L2, so no flush is needed between iterations).
metric = algorithmic HBM GB/s (A + packed B + scales + E bytes per GEMM) over
all launches; TFLOP/s and per-shape numbers ride along in "detail" (every entry timed under CUDA-graph replay).
The other half of BASELINE.json's metric rides on the same JSON line:
  "prefill"  W4A16 g128 at M=4096 on (11008,4096), (4096,11008), (28672,8192): TFLOP/s and its own tensor-bound roofline;
  "configs"  BASELINE.json configs: cfg1 fpAintB 3840x4096x4096, cfg3 bf16 x int8 Multiply + bias 4096x14336x4096,
             cfg4 int8 8192^3 MultiplyMultiply (against a cuBLASLt int8 peak measured in the same run), cfg5 28672x8192 at
             M in {1,16,4096};
  "cfg5_sharded" (N > 1): ONE global N=28672 weight column-sharded over the ranks (strong scaling), with and without the
             all-gather of the E slices.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--no-graph]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LAYERS = 32
LAYER_SHAPES = [(4096, 4096)] * 4 + [(11008, 4096)] * 2 + [(4096, 11008)]  # (N, K)
MS = (1, 4, 16)
GROUP = 128
METRIC = "w4a16_g128_decode_gemm_hbm_gbps"


def algo_bytes(M, N, K):
    """A + packed int4 B + bf16 group scales + E (no zero points)."""
    return 2 * M * K + N * K // 2 + (K // GROUP) * N * 2 + 2 * M * N


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def tensor_peaks():
    """dense bf16 TFLOP/s: (burst, sustained, source)."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["bf16_tflops"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured (MEASURED_PEAKS.json)"
    return 1650.0, 1400.0, "fallback (B200_PROFILING.md)"


def time_graph(torch, ops_list, reps=3, stream=None):
    """us per launch of `ops_list` replayed `reps` times as ONE CUDA graph (programmatic dependent launch between the
    decode kernels, no host launch cost) -- the same way the headline step is timed."""
    st = stream or torch.cuda.Stream()
    with torch.cuda.stream(st):
        for op in ops_list:
            op.run(st.cuda_stream)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for op in ops_list:
                op.run(st.cuda_stream)
        g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(reps):
            g.replay()
        b.record(st)
        torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / (reps * len(ops_list))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def workload_config(nl, world):
    """`config` of the JSON line: the same dict on both arms (--impl ours / reference), so that the driver compares like with like."""
    return {"workload": f"llama2-7b decode W4A16 g128: M in {{1,4,16}} x {nl} layers x 7 GEMMs "
                        "((4096,4096)x4,(11008,4096)x2,(4096,11008)x1)",
            "l2": "inputs larger than L2 (3.4 GB of distinct weights per step)",
            "parallelism": f"x{world} independent replicas of the GEMM list, one per GPU, own weights, no collective (weak scaling); "
                           "the N-sharded strong-scaling measurement of one global weight is the cfg5_sharded block"}


# --------------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (the restated ReferenceGemm oracle, all host threads) on a bounded sample of
    the same workload and the same `config`: per step ONE layer's 7 GEMMs at every M of the mix (21 GEMMs = 1/32 of the GPU arm's
    step, same shape and M mix, so the GB/s figures are like for like)."""
    if rank != 0:
        return
    import numpy as np
    from oracle import ck_oracle as orc
    rng = np.random.default_rng(0)
    threads = orc.lib().orc_hardware_threads()
    weights = []
    for (N, K) in LAYER_SHAPES:
        q = rng.integers(0, 16, (K, N), dtype=np.uint8)
        s = orc.bf16_trunc((rng.random((K // GROUP, N)) * 0.015 + 0.005).astype(np.float32))
        weights.append((q, s, N, K))
    acts = {(M, K): orc.bf16_trunc(rng.standard_normal((M, K)).astype(np.float32)) for M in MS for K in (4096, 11008)}

    def step():
        for M in MS:
            for q, s, N, K in weights:
                orc.gemm(acts[(M, K)], orc.BF16, q, orc.U8, orc.BF16, b_op=orc.BOP_GROUP_F32, scale=s, s_dtype=orc.BF16,
                         zero_const=8, group_size=GROUP)
    for _ in range(min(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    nbytes = sum(algo_bytes(M, N, K) for M in MS for (N, K) in LAYER_SHAPES)
    val = nbytes / dt / 1e9
    sample = (f"one layer (7 GEMMs) at every M in {{1,4,16}} per step = 21 GEMMs = 1/{args.layers} of the GPU arm's step "
              "(same shape and M mix; restated ReferenceGemm, ab_scale host pattern, all host threads)")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(args.layers, world),
        "launch": "host threads",
        "cpu_baseline": {"value": val, "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="launch eagerly instead of replaying a CUDA graph")
    ap.add_argument("--layers", type=int, default=LAYERS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the prefill / BASELINE-config blocks (headline only)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from composable_kernel_b200 import capi, ops

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"    # the version banner goes to stdout, in front of the one JSON line
        dist.init_process_group("nccl", device_id=dev)

    torch.manual_seed(1234 + rank)
    nl = args.layers
    # ---- synthetic weights: every byte pattern is a valid pair of 4-bit codes ---------------------
    weights = []
    for l in range(nl):
        for (N, K) in LAYER_SHAPES:
            nbytes = capi.lib().ckb_prepack_b_int4_bytes(N, K)
            packed = torch.randint(0, 256, (int(nbytes),), dtype=torch.uint8, device=dev)
            scales = (torch.rand(K // GROUP, N, device=dev) * 0.015 + 0.005).bfloat16()
            weights.append((N, K, packed, scales))
    acts = {(M, K): torch.randn(M, K, device=dev).bfloat16() for M in MS for K in (4096, 11008)}
    outs = {(M, N): torch.zeros(M, N, dtype=torch.bfloat16, device=dev) for M in MS for N in (4096, 11008)}
    gemm_ops, step_bytes, step_flops = [], 0, 0
    for M in MS:
        for (N, K, packed, scales) in weights:
            op = ops.GemmOp(acts[(M, K)], packed, outs[(M, N)], b_dtype=capi.I4, b_layout=capi.PREPACKED,
                            b_op=capi.BOP_GROUP, N=N, K=K, scale=scales, zero_const=8, group_size=GROUP)
            assert op.supported, op.why
            gemm_ops.append(op)
            step_bytes += algo_bytes(M, N, K)
            step_flops += 2 * M * N * K
    launches_per_step = len(gemm_ops)
    stream = torch.cuda.Stream(device=dev)

    def step_eager():
        s = torch.cuda.current_stream().cuda_stream
        for op in gemm_ops:
            op.run(s)

    graph = None
    if not args.no_graph:
        try:
            with torch.cuda.stream(stream):
                step_eager()
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    step_eager()
                graph = g
        except Exception as e:  # capture unsupported -> eager launches
            print(f"# graph capture failed ({e}); using eager launches", file=sys.stderr)
            graph = None
            torch.cuda.synchronize()

    def step():
        if graph is not None:
            graph.replay()
        else:
            step_eager()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step()
        barrier()
        clocks = ClockSampler(local_rank)
        if rank == 0:
            clocks.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(args.steps):
            step()
        ev1.record()
        barrier()
        ms_total = ev0.elapsed_time(ev1)
        clk = clocks.stop() if rank == 0 else None

        # per-shape detail: every (M, shape) family replayed as its own CUDA graph over all layers' weights (> L2)
        detail = {"launch": "cuda-graph replay (same as the headline step)"}
        if rank == 0:
            for M in MS:
                for (N, K) in sorted(set(LAYER_SHAPES)):
                    sel = [op for op in gemm_ops if op.problem.M == M and op.problem.N == N and op.problem.K == K]
                    us = time_graph(torch, sel, 3, stream)
                    detail[f"M{M}_N{N}_K{K}"] = {"us": round(us, 3), "gbps": round(algo_bytes(M, N, K) / us / 1e3, 1),
                                                 "frac_hbm": round(algo_bytes(M, N, K) / us / 1e3 / peaks()[0], 4),
                                                 "tflops": round(2 * M * N * K / us / 1e6, 3), "kernel": sel[0].kernel_name}

    # ---- the same 672 GEMMs with Llama's real dependencies declared (rank 0, 1 GPU): the k / v projections read the same layer
    # input as q, and `up` the same input as `gate`, so those three GEMMs of a layer carry CKB_FLAG_INPUTS_READY -- they do not
    # wait for their predecessor before reading A (they still complete in stream order).  The headline above keeps EVERY GEMM
    # dependent on its predecessor (strict chain, comparable with round 1); this block shows what the hint buys.
    dataflow = None
    if rank == 0 and world == 1 and graph is not None:
        df_ops = []
        i = 0
        for M in MS:
            for (N, K, packed, scales) in weights:
                pos = i % len(LAYER_SHAPES)          # q k v o gate up down
                fl = capi.FLAG_INPUTS_READY if pos in (1, 2, 5) else 0
                df_ops.append(ops.GemmOp(acts[(M, K)], packed, outs[(M, N)], b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP,
                                         N=N, K=K, scale=scales, zero_const=8, group_size=GROUP, flags=fl))
                i += 1
        us = time_graph(torch, df_ops, max(3, args.steps // 4), stream)
        dataflow = {"ms_per_step": round(us * len(df_ops) / 1e3, 4), "value": round(step_bytes / (us * len(df_ops)) / 1e3, 1), "unit": "GB/s",
                    "frac_hbm": round(step_bytes / (us * len(df_ops)) / 1e3 / peaks()[0], 4),
                    "declared_independent": "k, v, up (3 of the 7 GEMMs of a layer): same input as their predecessor",
                    "note": "CKB_FLAG_INPUTS_READY; every other GEMM waits for the one before it, as in the headline"}
        del df_ops

    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    value = world * step_bytes / (ms_per_step * 1e-3) / 1e9

    # ---- e2e: same step through the host-buffer C-ABI call (H2D of A, D2H of E inside the timing) ----
    e2e = None
    if rank == 0 or world > 1:
        import ctypes as C
        lib = capi.lib()
        NS = 4  # streams the independent GEMMs of a step are spread over (each with its own staging buffer)
        host_a = {k: v.cpu().pin_memory() for k, v in acts.items()}
        streams = [torch.cuda.Stream(device=dev) for _ in range(NS)]
        stagings = [torch.empty(1 << 20, dtype=torch.uint8, device=dev) for _ in range(NS)]
        probs, host_e = [], []
        h2d = d2h = 0
        for i, op in enumerate(gemm_ops):
            p = capi.GemmProblem()
            C.memmove(C.byref(p), C.byref(op.problem), C.sizeof(p))
            he = torch.zeros(p.M, p.N, dtype=torch.bfloat16).pin_memory()  # every GEMM of the step owns its result buffer
            host_e.append(he)
            p.a = host_a[(p.M, p.K)].data_ptr()
            p.e = he.data_ptr()
            probs.append((p, op.workspace.data_ptr() if op.workspace is not None else None, i % NS))
            h2d += 2 * p.M * p.K
            d2h += 2 * p.M * p.N
        e2e_steps = max(10, min(args.steps, 20))

        def enqueue(strs):
            for p, ws, k in probs:
                capi.check(lib.ckb_gemm_run_host_async(C.byref(p), stagings[k].data_ptr(), ws, strs[k].cuda_stream))

        def e2e_step():
            # every GEMM: A from pinned host memory -> device, kernel, E -> pinned host memory; the GEMMs of a step are
            # independent, so they are enqueued round-robin on NS streams (stream order protects each staging buffer) and
            # ONE device synchronisation ends the step, when the caller owns all 672 results.  The plan of each distinct
            # problem is cached inside the library.
            enqueue(streams)
            torch.cuda.synchronize()
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        # the same chain of C-ABI calls captured once and replayed as a CUDA graph (the copies are graph nodes: still inside
        # the timed region, every step moves the same bytes)
        dt_graph = None
        try:
            st2 = torch.cuda.Stream(device=dev)
            side = [torch.cuda.Stream(device=dev) for _ in range(NS)]
            with torch.cuda.stream(st2):
                g2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g2, stream=st2):
                    for sd in side:
                        sd.wait_stream(st2)
                    enqueue(side)
                    for sd in side:
                        st2.wait_stream(sd)
                g2.replay()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for _ in range(e2e_steps):
                    g2.replay()
                    torch.cuda.synchronize()
                dt_graph = (time.perf_counter() - t0) / e2e_steps
        except Exception as e:  # noqa: BLE001
            print(f"# e2e graph capture failed: {e}", file=sys.stderr)
            torch.cuda.synchronize()
        tt = torch.tensor([dt, dt_graph if dt_graph is not None else dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt, dtg = float(tt[0].item()), float(tt[1].item())
        best = min(dt, dtg)
        e2e = {"value": world * step_bytes / best / 1e9, "unit": "GB/s", "mode": "graph_replay" if dtg < dt else "eager_calls",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": best * 1e3, "steps": e2e_steps,
               "eager_calls": {"value": world * step_bytes / dt / 1e9, "ms_per_step": dt * 1e3},
               "graph_replay": None if dt_graph is None else {"value": world * step_bytes / dtg / 1e9, "ms_per_step": dtg * 1e3},
               "copies_declared": True,
               "note": "every step moves A of every GEMM from pinned host memory to the device and E of every GEMM back to pinned host "
                       "memory through ckb_gemm_run_host_async (GEMMs round-robin on 4 streams, one device sync per step, plans cached "
                       "in the library). eager_calls: the C-ABI is called per GEMM every step; graph_replay: the same call chain "
                       "captured once into a CUDA graph (copies are graph nodes) and replayed per step; value = the faster of the two. "
                       "Weights stay device-resident (the reference also uploads weights once, outside timing)"}

    # ---- the other half of the metric (rank 0, single-GPU runs): prefill M=4096 and the BASELINE configs ----------------
    prefill = configs = None
    if world == 1 and not args.no_extras:
        prefill, configs = extras(torch, capi, ops, dev)

    # ---- cfg5 column-sharded over the ranks (strong scaling), with and without the E all-gather ----------------------
    sharded = None
    if world > 1:
        sharded = cfg5_sharded(torch, dist, capi, ops, dev, rank, world)

    if rank == 0:
        peak, peak_src = peaks()
        us_per_launch = ms_per_step * 1e3 / launches_per_step
        achieved = step_bytes / launches_per_step / us_per_launch / 1e3  # GB/s per launch average
        cpu = None
        if not args.no_cpu_baseline:
            cpu = cpu_baseline()
        # DRAM bytes per launch cannot be measured inside this run (it needs a profiler): the committed ncu --set full capture of
        # the same kernels gives dram__bytes / algorithmic bytes, applied here to this run's algorithmic bytes -- labelled as such
        traffic, traffic_src = None, None
        for name in ("r2_traffic.json", "r1_traffic.json"):
            tp = os.path.join(ROOT, "profiles", name)
            if os.path.exists(tp):
                with open(tp) as f:
                    traffic = json.load(f)["traffic_over_algorithmic"] * step_bytes / launches_per_step
                traffic_src = f"profiles/{name} (ncu dram__bytes_read+write of these kernels / algorithmic bytes) x this run's algorithmic bytes; not measured live"
                break
        line = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(nl, world),
            "launch": "cuda-graph replay" if graph is not None else "eager",
            "tflops": world * step_flops / (ms_per_step * 1e-3) / 1e12,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes_per_launch": step_bytes / launches_per_step,
                         "peak_source": peak_src,
                         "kernel": gemm_ops[0].kernel_name, "kernels": sorted({op.kernel_name for op in gemm_ops}),
                         "us_per_launch": us_per_launch},
            "cpu_baseline": cpu,
            "e2e": e2e,
            "prefill": prefill,
            "configs": configs,
            "cfg5_sharded": sharded,
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clk,
            "dataflow": dataflow,
            "detail": detail,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _rot(nbytes_per_set, target=300e6, cap=6):
    """operand sets to rotate through so that consecutive launches never find their weights in the 126 MB L2"""
    return int(max(2, min(cap, target // max(1, nbytes_per_set) + 1)))


def extras(torch, capi, ops, dev):
    """Prefill M=4096 (W4A16 g128) and the BASELINE.json configs, each timed as a CUDA graph over rotating operand sets."""
    burst, sustained, src = tensor_peaks()
    hbm = peaks()[0]
    lib = capi.lib()

    def w4_ops(M, N, K, nrot=None):
        nrot = nrot or _rot(N * K // 2)
        a = torch.randn(M, K, device=dev).bfloat16()
        lst = []
        for _ in range(nrot):
            packed = torch.randint(0, 256, (int(lib.ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device=dev)
            sc = (torch.rand(K // GROUP, N, device=dev) * 0.015 + 0.005).bfloat16()
            e = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
            # numerics contract per regime: decode rows use the ab_scale form (fp32 group scale), prefill rows the one-rounding form;
            # each contract is served at every M, these are the fast sides of each
            op = ops.GemmOp(a, packed, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP_RTN if M > 64 else capi.BOP_GROUP,
                            N=N, K=K, scale=sc, zero_const=8, group_size=GROUP)
            assert op.supported, op.why
            lst.append(op)
        return lst

    def entry(us, M, N, K, nbytes, kernel, tensor=True):
        tf = 2 * M * N * K / us / 1e6
        d = {"us": round(us, 2), "tflops": round(tf, 2), "gbps": round(nbytes / us / 1e3, 1), "kernel": kernel}
        if tensor:
            d["frac_bf16_burst"] = round(tf / burst, 4)
            d["frac_bf16_sustained"] = round(tf / sustained, 4)
        else:
            d["frac_hbm"] = round(nbytes / us / 1e3 / hbm, 4)
        return d

    prefill = {"dtype": "bf16 A x int4 g128 B, fp32 accumulate", "M": 4096, "launch": "cuda-graph replay, rotating weight sets > L2",
               "peak_source": src, "shapes": {}}
    best = None
    for (N, K) in [(11008, 4096), (4096, 11008), (28672, 8192)]:
        lst = w4_ops(4096, N, K)
        us = time_graph(torch, lst, 3)
        prefill["shapes"][f"N{N}_K{K}"] = entry(us, 4096, N, K, algo_bytes(4096, N, K), lst[0].kernel_name)
        tf = 2 * 4096 * N * K / us / 1e6
        if best is None or N == 28672:
            best = (tf, lst[0].kernel_name, us)
        del lst
        torch.cuda.empty_cache()
    prefill["roofline"] = {"bound": "tensor", "achieved": round(best[0], 2), "peak": burst, "unit": "TFLOP/s",
                           "frac": round(best[0] / burst, 4), "frac_sustained": round(best[0] / sustained, 4),
                           "kernel": best[1], "shape": "M4096_N28672_K8192", "us_per_launch": round(best[2], 2), "traffic": None}

    configs = {}
    # cfg1: example/64_fpAintB_gemm default problem 3840 x 4096 x 4096, fp16 A x biased-u8 B, per-channel fp16 scale
    M, N, K = 3840, 4096, 4096
    a = torch.randn(M, K, device=dev).half()
    lst = []
    for _ in range(_rot(N * K)):
        b = ops.prepack_u8(torch.randint(-128, 128, (N, K), dtype=torch.int8, device=dev))
        sc = (torch.rand(N, device=dev) * 0.02 + 0.001).half()
        e = torch.empty(M, N, dtype=torch.float16, device=dev)
        lst.append(ops.GemmOp(a, b, e, b_op=capi.BOP_FPAINTB, scale=sc))
    us = time_graph(torch, lst, 3)
    configs["cfg1_fpAintB_fp16_i8_3840x4096x4096"] = entry(us, M, N, K, 2 * M * K + N * K + 2 * N + 2 * M * N, lst[0].kernel_name)
    del lst
    # cfg3: client_example/30 bf16 A x int8 B (Multiply by a bf16 scale) + bias, 4096 x 14336 x 4096
    M, N, K = 4096, 14336, 4096
    a = torch.randn(M, K, device=dev).bfloat16()
    lst = []
    for _ in range(_rot(N * K)):
        b = torch.randint(-128, 128, (N, K), dtype=torch.int8, device=dev)
        sc = (torch.rand(N, device=dev) * 0.01).bfloat16()
        bias = torch.randn(N, device=dev).bfloat16()
        e = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        lst.append(ops.GemmOp(a, b, e, b_op=capi.BOP_MULTIPLY, scale=sc, ds=(bias,), d_strides=((0, 1),), cde_op=capi.CDE_ADD))
    us = time_graph(torch, lst, 3)
    configs["cfg3_bf16_i8_multiply_bias_4096x14336x4096"] = entry(us, M, N, K, 2 * M * K + N * K + 4 * N + 2 * M * N, lst[0].kernel_name)
    del lst
    torch.cuda.empty_cache()
    # cfg4: example/65 int8 x int8 -> int32, MultiplyMultiply (column scale x row scale) -> bf16, 8192^3
    M = N = K = 8192
    lst = []
    for _ in range(3):
        a8 = torch.randint(-128, 128, (M, K), dtype=torch.int8, device=dev)
        b8 = torch.randint(-128, 128, (N, K), dtype=torch.int8, device=dev)
        d0 = torch.rand(N, device=dev) * 1e-3
        d1 = torch.rand(M, device=dev) * 1e-3
        e = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        lst.append(ops.GemmOp(a8, b8, e, ds=(d0, d1), d_strides=((0, 1), (1, 0)), cde_op=capi.CDE_MULTIPLY_MULTIPLY))
    us = time_graph(torch, lst, 3)
    c4 = {"us": round(us, 2), "tops": round(2 * M * N * K / us / 1e6, 1), "kernel": lst[0].kernel_name}
    # int8 dense peak of this GPU, measured here: cuBLASLt int8 8192^3 through torch._int_mm (best of 10)
    try:
        y = b8.t()  # [K, N] column-major view
        torch._int_mm(a8, y)
        torch.cuda.synchronize()
        bestt = 1e30
        for _ in range(10):
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            torch._int_mm(a8, y)
            t1.record()
            torch.cuda.synchronize()
            bestt = min(bestt, t0.elapsed_time(t1))
        c4["int8_peak_tops"] = round(2 * M * N * K / (bestt * 1e-3) / 1e12, 1)
        c4["int8_peak_source"] = "torch._int_mm (cuBLASLt int8) 8192^3, best of 10, this run"
        c4["frac_int8_peak"] = round(c4["tops"] / c4["int8_peak_tops"], 4)
    except Exception as ex:  # noqa: BLE001
        c4["int8_peak_tops"] = None
        c4["int8_peak_source"] = f"unavailable: {type(ex).__name__}: {str(ex)[:120]}"
    configs["cfg4_i8_i8_multiply_multiply_8192x8192x8192"] = c4
    del lst
    torch.cuda.empty_cache()
    # cfg5: Llama-2-70B FFN 28672 x 8192, W4A16 g128, decode M in {1,16} (HBM-bound) and prefill M = 4096 (tensor-bound)
    for M in (1, 16):
        lst = w4_ops(M, 28672, 8192, 4)
        us = time_graph(torch, lst, 3)
        configs[f"cfg5_w4a16_M{M}_28672x8192"] = entry(us, M, 28672, 8192, algo_bytes(M, 28672, 8192), lst[0].kernel_name, tensor=False)
        del lst
        torch.cuda.empty_cache()
    configs["cfg5_w4a16_M4096_28672x8192"] = prefill["shapes"]["N28672_K8192"]
    return prefill, configs


def cfg5_sharded(torch, dist, capi, ops, dev, rank, world):
    """ONE global W4A16 weight N = 28672, K = 8192 (same seed on every rank), column-sharded: rank g owns channels
    [g N/G, (g+1) N/G) of the packed weight and scales.  Per M: the shard GEMM alone and the shard GEMM followed by the
    all-gather of the E slices into the pre-allocated E_all[G][M][N/G], both captured in ONE CUDA graph per rank
    (NCCL all_gather_into_tensor is graph-capturable); max over ranks."""
    N, K = 28672, 8192
    n_loc = N // world
    assert n_loc % 16 == 0
    lib = capi.lib()
    gen = torch.Generator(device=dev)
    gen.manual_seed(777)  # identical global weight on every rank
    nrot = int(max(2, min(16, 300e6 // (n_loc * K // 2) + 1)))
    shards = []
    for _ in range(nrot):
        packed = torch.randint(0, 256, (int(lib.ckb_prepack_b_int4_bytes(N, K)),), dtype=torch.uint8, device=dev, generator=gen)
        sc = (torch.rand(K // GROUP, N, device=dev, generator=gen) * 0.015 + 0.005).bfloat16()
        per_group = (K // 64) * 512  # bytes of one 16-channel group in the fragment-major layout: shards are contiguous
        g0 = rank * n_loc // 16
        shards.append((packed[g0 * per_group:(g0 + n_loc // 16) * per_group].clone(), sc[:, rank * n_loc:(rank + 1) * n_loc].contiguous()))
        del packed, sc
    torch.cuda.empty_cache()
    out = {"N": N, "K": K, "world": world, "scaling": "strong", "launch": "cuda-graph replay incl. the gather (gemm_gather_us: NCCL all_gather_into_tensor in the graph; gemm_fused_gather_us: epilogue peer stores + ckb_peer_wait, no collective call)", "M": {}}
    st = torch.cuda.Stream(device=dev)
    for M in (1, 16, 4096):
        a = torch.randn(M, K, device=dev, generator=gen).bfloat16()
        e_loc = [torch.empty(M, n_loc, dtype=torch.bfloat16, device=dev) for _ in range(nrot)]
        e_all = torch.empty(world * M, n_loc, dtype=torch.bfloat16, device=dev)
        lst = [ops.GemmOp(a, p, e, b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP_RTN if M > 64 else capi.BOP_GROUP, N=n_loc, K=K, scale=s,
                          zero_const=8, group_size=GROUP) for (p, s), e in zip(shards, e_loc)]

        def run(gather):
            for op, e in zip(lst, e_loc):
                op.run(st.cuda_stream)
                if gather:
                    dist.all_gather_into_tensor(e_all, e)
        res = {}
        for gather in (False, True):
            with torch.cuda.stream(st):
                run(gather)
                torch.cuda.synchronize()
                dist.barrier()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):
                    run(gather)
                g.replay()
                torch.cuda.synchronize()
                dist.barrier()
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record(st)
                for _ in range(3):
                    g.replay()
                t1.record(st)
                torch.cuda.synchronize()
                dist.barrier()
            t = torch.tensor([t0.elapsed_time(t1) * 1e3 / (3 * nrot)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res["gemm_gather_us" if gather else "gemm_us"] = round(float(t.item()), 2)
            del g
        if M <= 32:
            # fused form: the epilogue stores the slice into every rank's gathered buffer over NVLink peer memory and the
            # step ends with ckb_peer_wait (no collective call); buffers rotate over two slots (shard.PeerGather)
            from composable_kernel_b200 import shard as _shard
            pg = _shard.PeerGather(M, n_loc, world, rank, torch.bfloat16, slots=2)
            fl = [ops.GemmOp(a, p, e_loc[0], b_dtype=capi.I4, b_layout=capi.PREPACKED, b_op=capi.BOP_GROUP, N=n_loc, K=K, scale=s,
                             zero_const=8, group_size=GROUP, e_peers=pg.e_peers(i % 2), flag_peers=pg.flag_peers(), peer_rank=rank)
                  for i, (p, s) in enumerate(shards)]
            n_f = len(fl) // 2 * 2
            with torch.cuda.stream(st):
                def run_fused():
                    for op in fl[:n_f]:
                        op.run(st.cuda_stream)
                        pg.wait(op.plan.grid, st.cuda_stream)
                run_fused()
                torch.cuda.synchronize()
                dist.barrier()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):
                    run_fused()
                g.replay()
                torch.cuda.synchronize()
                dist.barrier()
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record(st)
                for _ in range(3):
                    g.replay()
                t1.record(st)
                torch.cuda.synchronize()
                dist.barrier()
            t = torch.tensor([t0.elapsed_time(t1) * 1e3 / (3 * n_f)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res["gemm_fused_gather_us"] = round(float(t.item()), 2)
            res["fused_kernel"] = fl[0].kernel_name
            res["fused_gather_over_gemm"] = round(res["gemm_us"] / res["gemm_fused_gather_us"], 3)
            res["gbps_gemm_fused_gather"] = round(algo_bytes(M, N, K) / res["gemm_fused_gather_us"] / 1e3, 1)
            del g, fl, pg
        gbytes = 2 * M * N  # bytes every rank holds after the gather
        res["tflops_gemm"] = round(2 * M * N * K / res["gemm_us"] / 1e6, 2)
        res["tflops_gemm_gather"] = round(2 * M * N * K / res["gemm_gather_us"] / 1e6, 2)
        res["gbps_gemm"] = round(algo_bytes(M, N, K) / res["gemm_us"] / 1e3, 1)
        res["gbps_gemm_gather"] = round(algo_bytes(M, N, K) / res["gemm_gather_us"] / 1e3, 1)
        res["gather_over_gemm"] = round(res["gemm_us"] / res["gemm_gather_us"], 3)
        dt_g = max(1e-3, res["gemm_gather_us"] - res["gemm_us"])
        res["gather_payload_bytes_per_rank"] = gbytes
        res["gather_gbps_vs_900"] = round(gbytes * (world - 1) / world / dt_g / 1e3, 1)
        out["M"][str(M)] = res
        del lst, e_loc, e_all
    return out



def cpu_baseline(min_seconds=10.0):
    """The oracle (kind 'port': restated ReferenceGemm, all host threads) on a bounded sample of the same workload:
    the three layer shapes at M in {1,4,16}, repeated until >= min_seconds of CPU work."""
    import numpy as np
    from oracle import ck_oracle as orc
    rng = np.random.default_rng(0)
    t_total, nbytes, n_gemm = 0.0, 0, 0
    probs = []
    for (N, K) in sorted(set(LAYER_SHAPES)):
        q = rng.integers(0, 16, (K, N), dtype=np.uint8)
        s = orc.bf16_trunc((rng.random((K // GROUP, N)) * 0.015 + 0.005).astype(np.float32))
        probs.append((N, K, q, s))
    while t_total < min_seconds:
        for M in MS:
            for (N, K, q, s) in probs:
                a = orc.bf16_trunc(rng.standard_normal((M, K)).astype(np.float32))
                t0 = time.perf_counter()
                orc.gemm(a, orc.BF16, q, orc.U8, orc.BF16, b_op=orc.BOP_GROUP_F32, scale=s, s_dtype=orc.BF16,
                         zero_const=8, group_size=GROUP)
                t_total += time.perf_counter() - t0
                nbytes += algo_bytes(M, N, K)
                n_gemm += 1
    return {"value": nbytes / t_total / 1e9, "unit": "GB/s", "cores": orc.lib().orc_hardware_threads(),
            "kind": "port", "sample": f"{n_gemm} GEMMs: the 3 layer shapes x M in {{1,4,16}}, repeated for >= {min_seconds:.0f} s "
                                      "(restated ReferenceGemm, ab_scale host pattern, all host threads)",
            "seconds": t_total}


if __name__ == "__main__":
    main()
