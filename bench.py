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


def source_sha16(rel):
    import hashlib
    with open(os.path.join(ROOT, rel), "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()[:16]


def ncu_traffic(key):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture — only if
    the kernel source is still the one that was profiled (hash stored at capture time by
    tools/update_traffic.py); otherwise (None, reason)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            e = json.load(f)[key]
        now = source_sha16(e["kernel_source"])
        if now != e["kernel_source_sha16"]:
            return None, f"stale: {e['kernel_source']} changed since {e['source']} was captured"
        return float(e["bytes"]), e["source"]
    except Exception as ex:  # noqa: BLE001
        return None, f"unavailable: {type(ex).__name__}"


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


def bench_config(world: int) -> dict:
    """The `config` object — identical in both arms (b200 / reference) for the same N."""
    return {"workload": "3D 7-point star r=1, 512^3 fp64 per GPU, T=100 (BASELINE configs[1])",
            "grid": [N_SIDE, N_SIDE, N_SIDE * world], "time_steps": T_STEPS,
            "parallelism": f"z-slab x{world}",
            "l2": "inputs (2 x 1.15 GB per GPU) larger than the 126 MB L2",
            "init": "counter-based uniform [0,1) seed 1"}


PARITY = {"pinned_by_reference": False,
          "oracle": "definition (SURVEY.md 8(c)): the reference has no 3-D / fp64 code; the oracle generalises "
                    "stencil.cpp:94-131 and is pinned bit for bit on the 2-D star it does have",
          "checked_by": "tests/test_gpu_fullsize.py (full 512^3 checksum == oracle), tests/test_gpu_parity.py"}


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
    sample = (f"{sample_steps} of {T_STEPS} time steps of one 512^3 fp64 block of the grid per bench step "
              f"(throughput does not depend on T or on the number of blocks)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": bench_config(args.gpus),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "parity": PARITY,
    }), flush=True)


# ---- BASELINE configs[4]: 2048 x 1024 x 1024 fp64, strong scaling -----------------------------
C5_NZ, C5_NXY, C5_T = 2048, 1024, 40
# The 1-GPU run of this very leg (same seed, same T): the checksum every N must reproduce bit
# for bit, and the base of the strong-scaling efficiency together with the SM clock it ran at.
C5_BASE = {"checksum": "0xcc6b79639046d6e0", "value": 661.2, "sm_mhz": 1770.0,
           "source": "profiles/r02g_bench_n1.json (bench.py --steps 10, 1 x B200, sw_power_cap)"}


def _allreduce_u64_sum(torch, dist, world, c: int) -> int:
    cs = torch.tensor([c & 0x7FFFFFFF, (c >> 31) & 0x7FFFFFFF, c >> 62], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(cs)
    return (int(cs[0]) + (int(cs[1]) << 31) + (int(cs[2]) << 62)) % (1 << 64)


def config5_leg(args, torch, dist, capi, slab, rank, world, local, barrier):
    """Strong scaling of 2048x1024x1024 fp64 (z-slabs of 2048/N planes, fused peer-store halo
    exchange), the whole-grid checksum against the 1-GPU run, and a reduced grid through the
    same slab machinery against the CPU oracle."""
    import oracle
    out = {"workload": "3D 7-point star r=1, 2048x1024x1024 fp64 (z x y x x), T=40 (BASELINE configs[4])",
           "scaling": "strong", "n_gpus": world, "steps": C5_T}
    # (1) oracle check on a reduced grid, same decomposition and exchange
    small_nz = 24 * world
    ds = capi.make_desc((96, 80, small_nz // world), 1, capi.STAR, capi.F64, 9, time_block=args.time_block)
    sw = slab.SlabSweep(ds, rank, world, exchange=args.exchange)
    sw.fill(capi.INIT_RANDOM, seed=7)
    sw.sweep(9)
    got = _allreduce_u64_sum(torch, dist, world, sw.checksum())
    sw.close()
    if rank == 0:
        n = (96, 80, small_nz)
        a = oracle.fill(n, 1, oracle.F64, oracle.INIT_RANDOM, seed=7)
        b = oracle.fill(n, 1, oracle.F64, oracle.INIT_REF)
        want = oracle.checksum(n, 1, b if oracle.sweep(n, 1, oracle.STAR, 9, a, b, host_threads()) else a)
        out["oracle_check"] = {"grid": [96, 80, small_nz], "steps": 9, "checksum": f"{got:#018x}",
                               "oracle_checksum": f"{want:#018x}", "match": got == want}
    # (2) the full grid
    nz = C5_NZ // world
    d = capi.make_desc((C5_NXY, C5_NXY, nz), 1, capi.STAR, capi.F64, C5_T, time_block=args.time_block)
    sw = slab.SlabSweep(d, rank, world, exchange=args.exchange)
    sw.fill(capi.INIT_RANDOM, seed=1)
    sw.sweep(C5_T)
    total = _allreduce_u64_sum(torch, dist, world, sw.checksum())   # after exactly T steps from the seed
    sw.sweep(C5_T)                                                  # warm
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    reps = 3
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    launches = 0
    for _ in range(reps):
        launches += sw.sweep(C5_T)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    sw.check_link()
    ms_total = float(ms.item())
    clocks = sampler.stop() if rank == 0 else None
    k, ghost, exchange = sw.time_block, sw.ghost, sw.exchange
    sw.close()
    if rank != 0:
        return None
    pts = C5_NXY * C5_NXY * C5_NZ
    value = pts * C5_T * reps / (ms_total * 1e-3) / 1e9
    peak, _ = measured_peak_gbs()
    blocks = -(-C5_T // k) * reps
    launch_ms = ms_total / blocks
    face_bytes = 2 * ghost * C5_NXY * C5_NXY * 8 if world > 1 else 0   # out per interior GPU per block
    nvlink = face_bytes / (launch_ms * 1e-3) / 1e9
    compulsory = (nz + 2 * ghost + nz) * C5_NXY * C5_NXY * 8           # read slab + ghosts, write slab
    out.update({
        "value": value, "unit": UNIT, "ms_per_sweep": ms_total / reps, "time_block": k, "halo_exchange": exchange,
        "gpu_launches": launches,
        "roofline_per_gpu": {"achieved": value * 16 / world, "peak": peak, "unit": "GB/s",
                             "frac": value * 16 / world / peak,
                             "dram_frac_compulsory": compulsory / (launch_ms * 1e-3) / 1e9 / peak,
                             "note": "achieved = algorithmic bytes (16 B per update); dram_frac_compulsory = the "
                                     "bytes one launch must move (slab + ghost planes in, slab out) / its time"},
        "nvlink": {"halo_bytes_out_per_block_per_gpu": face_bytes, "out_GBps_per_gpu": nvlink,
                   "frac_of_900GBps": nvlink / 900.0},
        "checksum": f"{total:#018x}", "checksum_1gpu": C5_BASE["checksum"],
        "checksum_matches_1gpu": (f"{total:#018x}" == C5_BASE["checksum"]) if C5_BASE["checksum"] else None,
        "clocks": clocks, "link_timeout": False,
    })
    if C5_BASE["value"]:
        eff = value / world / C5_BASE["value"]
        out["strong_efficiency_vs_1gpu"] = {
            "base_value": C5_BASE["value"], "base_sm_mhz": C5_BASE["sm_mhz"], "base_source": C5_BASE["source"],
            "efficiency": eff,
            "efficiency_at_matched_clock": (eff * C5_BASE["sm_mhz"] / clocks["sm_mhz"]
                                            if clocks and clocks.get("sm_mhz") and C5_BASE["sm_mhz"] else None),
            "note": "efficiency = value / N / base_value; the matched-clock figure rescales by the median SM "
                    "clock of the two runs (under the 1000 W cap the long 1-GPU run clocks lower than the short "
                    "N-GPU runs, and throughput follows the clock: profiles/README.md r02)"}
    return out


def e2e_leg(args, torch, dist, capi, rank, world, local, cpu_barrier):
    """The same metric end to end through ONE sb200_run_host call on pinned HOST buffers: at
    N > 1 rank 0 binds all N GPUs (sb200_init_multi) and the call spreads the 512 x 512 x 512N
    grid over them as z-slabs; the other ranks idle."""
    import numpy as np
    L = capi.lib()
    if rank != 0:
        cpu_barrier()   # rank 0 runs the leg
        return None
    nz = N_SIDE * world
    shape = (nz + 2, N_SIDE + 2, N_SIDE + 2)
    nbytes = shape[0] * shape[1] * shape[2] * 8
    # one 512^3 block generated on the device, replicated along z on the host
    from stencil_b200 import sweep
    d1 = capi.make_desc((N_SIDE,) * 3, 1, capi.STAR, capi.F64, T_STEPS)
    g = sweep.DeviceGrids(d1, 0)
    g.fill(capi.INIT_RANDOM, 1)
    blocks = [g.download(0), g.download(1)]
    del g
    torch.cuda.empty_cache()
    L.sb200_finalize()
    capi.check(L.sb200_init_multi(world, None))
    hp = [ctypes.c_void_p(), ctypes.c_void_p()]
    host = []
    desc = capi.make_desc((N_SIDE, N_SIDE, nz), 1, capi.STAR, capi.F64, T_STEPS, time_block=args.time_block, ngpus=world)
    nodes = ctypes.c_int(0)
    for p, blk in zip(hp, blocks):
        # page-locked, every slab's planes on the memory node of its GPU
        capi.check(L.sb200_host_alloc_for(ctypes.byref(desc), ctypes.byref(p), ctypes.byref(nodes)))
        h = np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_double)), shape=shape)
        h[0] = blk[0]
        for b in range(world):
            h[1 + b * N_SIDE:1 + (b + 1) * N_SIDE] = blk[1:-1]
        h[-1] = blk[-1]
        host.append(h)
    capi.check(L.sb200_prepare(ctypes.byref(desc)))
    tm = capi.Timing()
    capi.check(L.sb200_run_host(ctypes.byref(desc), hp[0], hp[1], ctypes.byref(tm)))   # warm
    e2e_steps = max(2, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        capi.check(L.sb200_run_host(ctypes.byref(desc), hp[0], hp[1], ctypes.byref(tm)))
    dt = time.perf_counter() - t0
    plan = capi.slab_plan(desc, world)
    up_planes = sum(plan.slab[i].sub_hi - plan.slab[i].sub_lo + 2 for i in range(plan.ngpus_used))
    e2e = {"value": N_SIDE ** 3 * world * T_STEPS * e2e_steps / dt / 1e9, "unit": UNIT,
           "h2d_bytes_per_step": int(up_planes) * shape[1] * shape[2] * 8, "d2h_bytes_per_step": N_SIDE ** 3 * world * 8,
           "steps": e2e_steps, "ms_per_call": dt / e2e_steps * 1e3, "host_memory_nodes": nodes.value,
           "h2d_ms": tm.h2d_ms, "sweep_ms": tm.sweep_ms, "d2h_ms": tm.d2h_ms, "gpus_used": tm.ngpus,
           "grid": [N_SIDE, N_SIDE, nz],
           "note": "ONE sb200_run_host call from one process on pinned host buffers holding the whole "
                   f"512 x 512 x {nz} grid; the call cuts it into {tm.ngpus} z-slab(s) with T*r = {plan.deep_halo} planes "
                   "of deep halo per internal side (read from the host array, no GPU-to-GPU traffic) and every slab "
                   "runs the overlapped upload / sweep / download schedule on its own PCIe link; h2d_ms / sweep_ms / "
                   "d2h_ms = max over GPUs of the upload span / whole device pipeline / download span"}
    for p in hp:
        L.sb200_host_free(p)
    L.sb200_finalize()
    capi.check(L.sb200_init(local))
    cpu_barrier()
    return e2e


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--time-block", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-config5", action="store_true", help="skip the 2048x1024x1024 strong-scaling leg")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="strong: the headline value is the config-5 leg (fixed 2048x1024x1024 grid)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "p2p", "nccl"],
                    help="multi-GPU halo exchange: fused peer stores (p2p) or NCCL send/recv")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3

    import torch
    import torch.distributed as dist

    from stencil_b200 import capi, slab

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    capi.check(capi.lib().sb200_init(local))
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        cpu_group = dist.new_group(backend="gloo")   # host-side rendezvous that launches nothing on the GPUs

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def cpu_barrier():
        if world > 1:
            dist.barrier(group=cpu_group)

    n = (N_SIDE, N_SIDE, N_SIDE)
    pts_rank = N_SIDE ** 3
    desc = capi.make_desc(n, 1, capi.STAR, capi.F64, T_STEPS, time_block=args.time_block)
    runner = slab.SlabSweep(desc, rank, world, exchange=args.exchange)
    runner.fill(capi.INIT_RANDOM, seed=1)
    for _ in range(args.warmup):
        runner.sweep(T_STEPS)
    barrier()
    runner.check_link()
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
    runner.check_link()   # a timed-out halo exchange would have produced garbage at full speed
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
    traffic, traffic_src = ncu_traffic(f"tma3d_f64_star_k{k}_512") if world == 1 else (None, "single-GPU capture only")
    impl_detail = {"time_block": k, "kernel": runner.kernel_name, "halo_exchange": runner.exchange}
    runner.close()
    del runner
    torch.cuda.empty_cache()

    c5 = None
    if not args.no_config5:
        c5 = config5_leg(args, torch, dist, capi, slab, rank, world, local, barrier)
    e2e = None
    if not args.no_e2e:
        e2e = e2e_leg(args, torch, dist, capi, rank, world, local, cpu_barrier)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = host_threads()
        v, s, dt = cpu_sample(threads)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{s} time steps of the full 512^3 fp64 grid ({dt:.1f} s; the workload runs {T_STEPS}), "
                         f"oracle/stencil_oracle.c with OpenMP"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": bench_config(world), "impl_detail": impl_detail,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "k": k, "traffic": traffic, "traffic_source": traffic_src,
                         "dram_frac": (traffic / (avg_launch_ms * 1e-3) / 1e9 / peak) if traffic else None,
                         "peak_kind": peak_kind,
                         "alg_bytes_per_launch": alg_bytes, "avg_launch_ms": avg_launch_ms,
                         "note": "frac counts ALGORITHMIC bytes (16 B per update x k fused steps per launch) and "
                                 "exceeds 1 when k > 1; dram_frac = physical DRAM bytes (ncu) / launch time / peak"},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "parity": PARITY, "config5": c5,
        }
        if args.scaling == "strong" and c5:
            line.update({"metric": "GStencil/s (grid-point updates/s), 3D 7-point fp64 heat stencil 2048x1024x1024, 40 steps",
                         "value": c5["value"], "scaling": "strong", "ms_per_step": c5["ms_per_sweep"],
                         "config": {"workload": c5["workload"], "grid": [C5_NXY, C5_NXY, C5_NZ], "time_steps": C5_T,
                                    "parallelism": f"z-slab x{world}"}, "headline_config2": value})
        print(json.dumps(line), flush=True)   # flushed here: a teardown hiccup must not eat the line
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
