#!/usr/bin/env python
"""bench.py — headline benchmark: GStencil/s of the 3-D 7-point fp64 heat stencil.

Workload (BASELINE.json configs[1]): 512^3 fp64, radius-1 star, T = 100 double-buffered
steps, single B200.  One bench "step" = one full T-step sweep of the grid.  With
--gpus N (torchrun, one rank per GPU) the grid is weak-scaled: N slabs of 512^3 stacked
along z (512 x 512 x 512N) with ghost-plane halo exchange every fused time block.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = device-resident throughput (inputs already in
HBM); `e2e` = the same metric through the host-buffer C-ABI call sb200_run_host with the
H2D/D2H copies inside the timed region; `roofline` = algorithmic bytes of one launch of
the sweep kernel / its average duration vs the measured HBM peak; `cpu_baseline` = the
CPU oracle port (OpenMP, all host cores) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_SIDE = 512
T_STEPS = 100
METRIC = "GStencil/s (grid-point updates/s), 3D 7-point fp64 heat stencil 512^3, 100 steps"
UNIT = "GStencil/s"


def ncu_traffic(key):
    """DRAM bytes per launch of the dominant kernel, from the committed ncu capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            e = json.load(f)[key]
        return float(e["bytes"]), e["source"]
    except Exception:  # noqa: BLE001
        return None, None


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


class ClockSampler:
    """Samples nvidia-smi SM clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            p = [x.strip() for x in s.split(",")]
            if len(p) < 6:
                continue
            try:
                sm.append(float(p[0]))
                mx = float(p[1])
            except ValueError:
                continue
            for nm, v in zip(names, p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def host_threads() -> int:
    """Host cores this process may use.  Not omp_get_max_threads(): torchrun exports
    OMP_NUM_THREADS=1 to its workers, which would silently time a 1-thread baseline."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_sample(threads: int, budget_s: float = 12.0):
    """Time the oracle port on the workload grid for a bounded number of steps."""
    import oracle
    n = (N_SIDE, N_SIDE, N_SIDE)
    a = oracle.fill(n, 1, oracle.F64, oracle.INIT_RANDOM, seed=1)
    b = oracle.fill(n, 1, oracle.F64, oracle.INIT_REF)
    t0 = time.perf_counter()
    oracle.sweep(n, 1, oracle.STAR, 1, a, b, threads)
    one = time.perf_counter() - t0
    steps = max(1, min(2000, int(budget_s / max(one, 1e-3))))
    t0 = time.perf_counter()
    oracle.sweep(n, 1, oracle.STAR, steps, a, b, threads)
    dt = time.perf_counter() - t0
    return N_SIDE ** 3 * steps / dt / 1e9, steps, dt


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (oracle port of
    src/stencil/stencil.cpp:94-131 generalised to 3-D fp64; the reference itself is 2-D
    fp32 only) on all host cores; each bench step is a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    oracle.lib()
    threads = host_threads()
    n = (N_SIDE, N_SIDE, N_SIDE)
    a = oracle.fill(n, 1, oracle.F64, oracle.INIT_RANDOM, seed=1)
    b = oracle.fill(n, 1, oracle.F64, oracle.INIT_REF)
    sample_steps = 10
    for _ in range(args.warmup):
        oracle.sweep(n, 1, oracle.STAR, 1, a, b, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.sweep(n, 1, oracle.STAR, sample_steps, a, b, threads)
    dt = time.perf_counter() - t0
    val = N_SIDE ** 3 * sample_steps * args.steps / dt / 1e9
    sample = f"{sample_steps} of {T_STEPS} time steps of the 512^3 fp64 grid per bench step"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": "3D 7-point star r=1, 512^3 fp64, T=100 (BASELINE configs[1])",
                   "sample": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--time-block", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--exchange", default="auto", choices=["auto", "p2p", "nccl"],
                    help="multi-GPU halo exchange: fused peer stores (p2p) or NCCL send/recv")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3

    import numpy as np
    import torch
    import torch.distributed as dist

    from stencil_b200 import capi, slab, sweep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    capi.check(capi.lib().sb200_init(local))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    n = (N_SIDE, N_SIDE, N_SIDE)
    pts_rank = N_SIDE ** 3
    desc = capi.make_desc(n, 1, capi.STAR, capi.F64, T_STEPS, time_block=args.time_block)
    runner = slab.SlabSweep(desc, rank, world, exchange=args.exchange)
    runner.fill(capi.INIT_RANDOM, seed=1)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        runner.sweep(T_STEPS)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    barrier()
    e0.record()
    for _ in range(args.steps):
        launches += runner.sweep(T_STEPS)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    clocks = sampler.stop() if rank == 0 else None
    value = pts_rank * world * T_STEPS * args.steps / (ms_total * 1e-3) / 1e9

    # roofline of the dominant kernel: one launch fuses k steps over the whole slab
    k = runner.time_block
    sweep_launches = launches // runner.kernels_per_launch_unit
    avg_launch_ms = ms_total / max(sweep_launches, 1)
    alg_bytes = 16.0 * pts_rank * k
    peak, peak_kind = measured_peak_gbs()
    achieved = alg_bytes / (avg_launch_ms * 1e-3) / 1e9
    traffic, traffic_src = (ncu_traffic(f"tma3d_f64_star_k{k}_512") if world == 1 else (None, None))

    # e2e through the host-buffer C-ABI call (pinned host memory), rank-local grid
    e2e = None
    if not args.no_e2e:
        lay_host = (N_SIDE + 2) ** 3 * 8
        hp = [ctypes.c_void_p(), ctypes.c_void_p()]
        for p in hp:
            capi.check(capi.lib().sb200_host_alloc(ctypes.byref(p), lay_host))
        ha = np.ctypeslib.as_array(ctypes.cast(hp[0], ctypes.POINTER(ctypes.c_double)), shape=(N_SIDE + 2,) * 3)
        hb = np.ctypeslib.as_array(ctypes.cast(hp[1], ctypes.POINTER(ctypes.c_double)), shape=(N_SIDE + 2,) * 3)
        runner.download_initial_into(ha, hb, seed=1)
        tm = capi.Timing()
        e2e_steps = max(2, min(args.steps, 3))
        capi.check(capi.lib().sb200_run_host(ctypes.byref(desc), hp[0], hp[1], ctypes.byref(tm)))
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            capi.check(capi.lib().sb200_run_host(ctypes.byref(desc), hp[0], hp[1], ctypes.byref(tm)))
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": pts_rank * world * T_STEPS * e2e_steps / float(dt.item()) / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": lay_host, "d2h_bytes_per_step": N_SIDE ** 3 * 8,
               "steps": e2e_steps, "h2d_ms": tm.h2d_ms, "sweep_ms": tm.sweep_ms, "d2h_ms": tm.d2h_ms,
               "note": "sb200_run_host on pinned host buffers, each rank its own 512^3 grid; upload, sweep and "
                       "download overlap (z-t parallelograms on separate streams): h2d_ms = all upload chunks, "
                       "sweep_ms = the whole device pipeline, d2h_ms = first to last download chunk"}
        for p in hp:
            capi.lib().sb200_host_free(p)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = host_threads()
        v, s, dt = cpu_sample(threads)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{s} time steps of the full 512^3 fp64 grid ({dt:.1f} s; the workload runs {T_STEPS}), "
                         f"oracle/stencil_oracle.c with OpenMP"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "3D 7-point star r=1, 512^3 fp64 per GPU, T=100 (BASELINE configs[1])",
                       "grid": [N_SIDE, N_SIDE, N_SIDE * world], "time_block": k,
                       "kernel": runner.kernel_name, "parallelism": f"z-slab x{world}", "halo_exchange": runner.exchange,
                       "l2": "inputs (2 x 1.15 GB per GPU) larger than the 126 MB L2",
                       "init": "counter-based uniform [0,1) seed 1"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_kind": peak_kind,
                         "alg_bytes_per_launch": alg_bytes, "avg_launch_ms": avg_launch_ms},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        }))
    runner.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
