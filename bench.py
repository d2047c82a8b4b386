#!/usr/bin/env python
"""bench.py -- headline benchmark: Llama-2-7B decode, W4A16 group-128 GEMMs (BASELINE.json configs[1]).

One STEP = one decode pass over all 32 layers' weight GEMMs at each M in {1,4,16}:
  per layer 4 x (N=4096,K=4096) [q,k,v,o], 2 x (N=11008,K=4096) [gate,up], 1 x (N=4096,K=11008) [down]
  -> 224 GEMMs per M, 672 launches per step, every GEMM on its own layer's weights (3.4 GB of packed
  int4 + scales per rank: far larger than the 126 MB L2, so no flush is needed between iterations).
metric = algorithmic HBM GB/s (A + packed B + scales + E bytes per GEMM) over
all launches; TFLOP/s and per-shape numbers ride along in "detail".

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


# --------------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (the restated ReferenceGemm oracle, all host
    threads) on a bounded sample of the same workload: one layer's 7 GEMMs at M=4 per step."""
    if rank != 0:
        return
    import numpy as np
    from oracle import ck_oracle as orc
    rng = np.random.default_rng(0)
    threads = orc.lib().orc_hardware_threads()
    M = 4
    probs = []
    for (N, K) in LAYER_SHAPES:
        a = orc.bf16_trunc(rng.standard_normal((M, K)).astype(np.float32))
        q = rng.integers(0, 16, (K, N), dtype=np.uint8)
        s = orc.bf16_trunc((rng.random((K // GROUP, N)) * 0.015 + 0.005).astype(np.float32))
        probs.append((a, q, s, N, K))

    def step():
        for a, q, s, N, K in probs:
            orc.gemm(a, orc.BF16, q, orc.U8, orc.BF16, b_op=orc.BOP_GROUP_F32, scale=s, s_dtype=orc.BF16,
                     zero_const=8, group_size=GROUP)
    for _ in range(min(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    nbytes = sum(algo_bytes(M, N, K) for (N, K) in LAYER_SHAPES)
    val = nbytes / dt / 1e9
    sample = "one layer (7 GEMMs) at M=4 per step = 1/96 of the GPU step's GEMM list"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "llama2-7b decode W4A16 g128, M in {1,4,16}, 32 layers x 7 GEMMs", "sample": sample},
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

        # per-shape detail (eager back-to-back launches of one op family, its own events)
        detail = {}
        if rank == 0:
            for M in MS:
                for (N, K) in sorted(set(LAYER_SHAPES)):
                    sel = [op for op in gemm_ops if op.problem.M == M and op.problem.N == N and op.problem.K == K]
                    s = torch.cuda.current_stream().cuda_stream
                    for op in sel:
                        op.run(s)
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    for _ in range(3):
                        for op in sel:
                            op.run(s)
                    b.record()
                    torch.cuda.synchronize()
                    us = a.elapsed_time(b) * 1e3 / (3 * len(sel))
                    detail[f"M{M}_N{N}_K{K}"] = {"us": round(us, 3), "gbps": round(algo_bytes(M, N, K) / us / 1e3, 1),
                                                 "tflops": round(2 * M * N * K / us / 1e6, 3)}

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
        host_a = {k: v.cpu().pin_memory() for k, v in acts.items()}
        host_e = {k: torch.zeros_like(v, device="cpu").pin_memory() for k, v in outs.items()}
        staging = torch.empty(1 << 20, dtype=torch.uint8, device=dev)
        probs = []
        h2d = d2h = 0
        for op in gemm_ops:
            p = capi.GemmProblem()
            C.memmove(C.byref(p), C.byref(op.problem), C.sizeof(p))
            p.a = host_a[(p.M, p.K)].data_ptr()
            p.e = host_e[(p.M, p.N)].data_ptr()
            probs.append((p, op.workspace.data_ptr() if op.workspace is not None else None))
            h2d += 2 * p.M * p.K
            d2h += 2 * p.M * p.N
        s = torch.cuda.current_stream().cuda_stream
        e2e_steps = max(1, min(args.steps, 3))

        def e2e_step():
            # every GEMM: A from pinned host memory -> device, kernel, E -> pinned host memory, all enqueued on one
            # stream (stream order protects the shared staging buffer); ONE synchronisation ends the step, when the
            # caller owns all 672 results
            for p, ws in probs:
                capi.check(lib.ckb_gemm_run_host_async(C.byref(p), staging.data_ptr(), ws, s))
            torch.cuda.synchronize()
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e = {"value": world * step_bytes / float(tt.item()) / 1e9, "unit": "GB/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": float(tt.item()) * 1e3,
               "note": "ckb_gemm_run_host_async per GEMM: pinned-host A -> device, kernel, E -> pinned host, one stream sync per step; "
                       "weights stay device-resident (the reference also uploads weights once, outside timing)"}

    # ---- N-sharded path with the output all-gather (NCCL over NVLink), timed separately ----------------
    with_gather = None
    if world > 1:
        from composable_kernel_b200 import shard
        sel = gemm_ops[::8]  # bounded sample of the step's GEMM list (every 8th launch)
        def gather_step():
            s_ = torch.cuda.current_stream().cuda_stream
            for op in sel:
                op.run(s_)
                shard.all_gather_columns(outs[(op.problem.M, op.problem.N)], world, row_major=False)
        gather_step()
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(3):
            gather_step()
        g1.record()
        barrier()
        tg = torch.tensor([g0.elapsed_time(g1) / 3], dtype=torch.float64, device=dev)
        dist.all_reduce(tg, op=dist.ReduceOp.MAX)
        sel_bytes = sum(algo_bytes(op.problem.M, op.problem.N, op.problem.K) for op in sel)
        with_gather = {"value": world * sel_bytes / (float(tg.item()) * 1e-3) / 1e9, "unit": "GB/s",
                       "sample": f"every 8th GEMM of the step ({len(sel)} launches), eager, all_gather_into_tensor of "
                                 "each E slice into E_all[G][M][N/G] after every GEMM"}

    if rank == 0:
        peak, peak_src = peaks()
        us_per_launch = ms_per_step * 1e3 / launches_per_step
        achieved = step_bytes / launches_per_step / us_per_launch / 1e3  # GB/s per launch average
        cpu = None
        if not args.no_cpu_baseline:
            cpu = cpu_baseline()
        traffic = None  # dram bytes per launch from the committed ncu --set full capture (ratio x algorithmic bytes)
        tp = os.path.join(ROOT, "profiles", "r1_traffic.json")
        if os.path.exists(tp):
            with open(tp) as f:
                traffic = json.load(f)["traffic_over_algorithmic"] * step_bytes / launches_per_step
        line = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"llama2-7b decode W4A16 g128: M in {{1,4,16}} x {nl} layers x 7 GEMMs "
                                   "((4096,4096)x4,(11008,4096)x2,(4096,11008)x1)",
                       "launch": "cuda-graph replay" if graph is not None else "eager",
                       "l2": "inputs larger than L2 (3.4 GB of distinct weights per step)",
                       "parallelism": f"column-shard x{world}: every rank owns an N-shard of the shapes above "
                                      "(independent shards, no collective in the GEMM; all-gather reported in with_allgather)"},
            "tflops": world * step_flops / (ms_per_step * 1e-3) / 1e12,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "algorithmic_bytes_per_launch": step_bytes / launches_per_step,
                         "peak_source": peak_src,
                         "kernel": gemm_ops[0].kernel_name, "us_per_launch": us_per_launch},
            "cpu_baseline": cpu,
            "e2e": e2e,
            "with_allgather": with_gather,
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clk,
            "detail": detail,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


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
