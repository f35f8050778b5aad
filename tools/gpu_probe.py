"""Developer timing probe (run under gpurun): device-resident sweeps of each BASELINE
config with several kernels / time blocks, CUDA-event timed.  Prints one line per variant.
Not the bench (bench.py is); used to choose defaults."""
import ctypes
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from stencil_b200 import capi, sweep

capi.check(capi.lib().sb200_init(0))
PEAK = 6529.1


def run(name, n, shape, dtype, T, kernel, k, reps=3, r=1, flags=0, hint=0):
    d = capi.make_desc(n, r, shape, dtype, T, time_block=k, kernel=kernel, flags=flags | (capi.FLAG_BLOCK_TILE if hint else 0), block_hint=hint)
    try:
        g = sweep.DeviceGrids(d)
        g.fill(capi.INIT_RANDOM, seed=1)
        g.sweep(min(T, 2 * max(k, 1)))
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            nl = g.sweep(T)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        pts = 1
        for v in n:
            pts *= v
        gs = pts * T / best / 1e6
        es = 8 if dtype == capi.F64 else 4
        print(f"{name:10s} n={n} kern={kernel} k={k} T={T}: {best:9.3f} ms  {gs:9.1f} GStencil/s  "
              f"alg {gs * 2 * es:8.1f} GB/s  frac {gs * 2 * es / PEAK:5.3f}  launches {nl}", flush=True)
        del g
    except Exception as e:  # noqa: BLE001
        print(f"{name} n={n} kern={kernel} k={k}: FAILED {e}", flush=True)
    torch.cuda.empty_cache()


which = sys.argv[1:] or ["c1", "c2", "c3", "c4"]
if "c2" in which:
    for k in (1, 2, 3):
        run("c2-7pt", (512, 512, 512), capi.STAR, capi.F64, 24, capi.KERNEL_TMA3D, k)
    run("c2-7pt", (512, 512, 512), capi.STAR, capi.F64, 10, capi.KERNEL_GENERIC, 1)
    for k in (1, 2, 3):
        run("c2-7pt-f32", (512, 512, 512), capi.STAR, capi.F32, 24, capi.KERNEL_TMA3D, k)
if "c1" in which:
    for k in (1, 2, 4, 6):
        run("c1-5pt-f64", (4096, 4096), capi.STAR, capi.F64, 96, capi.KERNEL_WARP2D, k)
    run("c1-5pt-f64", (4096, 4096), capi.STAR, capi.F64, 20, capi.KERNEL_GENERIC, 1)
    for k in (1, 4, 6):
        run("c1-5pt-f32", (4096, 4096), capi.STAR, capi.F32, 96, capi.KERNEL_WARP2D, k)
if "c4" in which:
    for k in (1, 2, 4, 6):
        run("c4-9pt-f32", (16384, 16384), capi.BOX, capi.F32, 48, capi.KERNEL_WARP2D, k)
    run("c4-9pt-f32", (16384, 16384), capi.BOX, capi.F32, 6, capi.KERNEL_GENERIC, 1)
if "c1k" in which:  # config 1: every fused depth
    for dt, nm in ((capi.F64, "f64"), (capi.F32, "f32")):
        for k in (1, 2, 3, 4, 5, 6):
            run(f"c1-{nm}", (4096, 4096), capi.STAR, dt, 120, capi.KERNEL_WARP2D, k)
if "star64" in which:  # fp64 5-point star at other sizes: default depth 4 against 6
    for n in ((2048, 2048), (8192, 8192), (16384, 8192), (3000, 5000)):
        for k in (4, 6):
            run("5pt-f64", n, capi.STAR, capi.F64, 120, capi.KERNEL_WARP2D, k)
if "c1tile" in which:  # config 1 (and its fp32 twin): rows per warp (SB200_FLAG_BLOCK_TILE) against the planner's choice
    for dt, nm in ((capi.F64, "f64"), (capi.F32, "f32")):
        for k in (4, 6):
            for rows in (0, 32, 48, 64, 82, 96, 128, 164, 256):
                run(f"c1-{nm}-rows{rows}", (4096, 4096), capi.STAR, dt, 96, capi.KERNEL_WARP2D, k, hint=rows)
if "c2tile" in which:  # configs 2 / 3 / 5-slab: planes per CTA against the planner's makespan search
    for nm, n, shp, dt in (("c2", (512, 512, 512), capi.STAR, capi.F64), ("c3", (512, 512, 512), capi.BOX, capi.F64),
                           ("c5slab", (1024, 1024, 256), capi.STAR, capi.F64), ("c2f32", (512, 512, 512), capi.STAR, capi.F32)):
        for planes in (0, 16, 24, 32, 40, 48, 56, 64, 86, 103, 128, 171, 256):
            run(f"{nm}-planes{planes}", n, shp, dt, 24, capi.KERNEL_TMA3D, 2, hint=planes)
if "c4tile" in which:  # config 4: rows per warp against the planner's choice
    for k in (4, 6):
        for rows in (0, 64, 100, 128, 150, 200, 250, 300, 350, 400, 469, 600, 781, 1024, 1600):
            run(f"c4-rows{rows}", (16384, 16384), capi.BOX, capi.F32, 48, capi.KERNEL_WARP2D, k, hint=rows)
        for rows in (0, 150, 200, 250, 300, 400, 781):
            run(f"c4-sep-rows{rows}", (16384, 16384), capi.BOX, capi.F32, 48, capi.KERNEL_WARP2D, k, hint=rows, flags=capi.FLAG_REASSOCIATE)
if "sep9" in which:  # separable 9-point box (SB200_FLAG_REASSOCIATE) against the reference-order level
    for dt, nm, n in ((capi.F32, "f32", 16384), (capi.F64, "f64", 8192)):
        for k in (2, 4, 6):
            run(f"9pt-{nm}", (n, n), capi.BOX, dt, 48, capi.KERNEL_WARP2D, k)
            run(f"9pt-{nm}-sep", (n, n), capi.BOX, dt, 48, capi.KERNEL_WARP2D, k, flags=capi.FLAG_REASSOCIATE)
if "c3" in which:
    run("c3-27pt", (512, 512, 512), capi.BOX, capi.F64, 6, capi.KERNEL_GENERIC, 1)
    for k in (1, 2):
        run("c3-27pt", (512, 512, 512), capi.BOX, capi.F64, 12, capi.KERNEL_TMA3D, k)
        run("c3-27pt-f32", (512, 512, 512), capi.BOX, capi.F32, 12, capi.KERNEL_TMA3D, k)
if "rad" in which:  # general-radius 2-D star: fused warp kernel vs the one-thread-per-point kernel
    for dt, nm in ((capi.F32, "f32"), (capi.F64, "f64")):
        for r in (2, 3, 4):
            for k in (1, 2, 3, 4):
                if k <= capi.lib().sb200_max_time_block(ctypes.byref(
                        capi.make_desc((8192, 8192), r, capi.STAR, dt, 24, kernel=capi.KERNEL_WARP2D))):
                    run(f"r{r}-{nm}", (8192, 8192), capi.STAR, dt, 24, capi.KERNEL_WARP2D, k, r=r)
            run(f"r{r}-{nm}", (8192, 8192), capi.STAR, dt, 8, capi.KERNEL_GENERIC, 1, r=r)
if "star13" in which:  # 3-D radius 2..4 star: TMA kernel vs one-thread-per-vector kernel
    for dt, nm in ((capi.F32, "f32"), (capi.F64, "f64")):
        for r in (2, 3, 4):
            run(f"star3d-r{r}-{nm}", (512, 512, 512), capi.STAR, dt, 12, capi.KERNEL_TMA3D, 1, r=r)
            run(f"star3d-r{r}-{nm}", (512, 512, 512), capi.STAR, dt, 6, capi.KERNEL_GENERIC, 1, r=r)
if "box2" in which:  # 25-point box: fused warp kernel vs one launch per step
    for dt, nm in ((capi.F32, "f32"), (capi.F64, "f64")):
        for k in (1, 2):
            if k <= capi.lib().sb200_max_time_block(ctypes.byref(
                    capi.make_desc((8192, 8192), 2, capi.BOX, dt, 24, kernel=capi.KERNEL_WARP2D))):
                run(f"box2-{nm}", (8192, 8192), capi.BOX, dt, 24, capi.KERNEL_WARP2D, k, r=2)
        run(f"box2-{nm}", (8192, 8192), capi.BOX, dt, 8, capi.KERNEL_GENERIC, 1, r=2)
        for r in (3, 4):  # 49- and 81-point
            run(f"box{r}-{nm}", (8192, 8192), capi.BOX, dt, 8, capi.KERNEL_WARP2D, 1, r=r)
            run(f"box{r}-{nm}", (8192, 8192), capi.BOX, dt, 4, capi.KERNEL_GENERIC, 1, r=r)
if "box3" in which:  # 3-D box radius 2 (125-point): which kernel, and how far from the FP-add bound
    for dt, nm in ((capi.F32, "f32"), (capi.F64, "f64")):
        run(f"125pt-{nm}", (512, 512, 256), capi.BOX, dt, 4, capi.KERNEL_GENERIC, 1, r=2)
        run(f"125pt-{nm}", (512, 512, 256), capi.BOX, dt, 4, capi.KERNEL_TMA3D, 1, r=2)
        run(f"125pt-{nm}", (512, 512, 512), capi.BOX, dt, 4, capi.KERNEL_AUTO, 0, r=2)
        run(f"343pt-{nm}", (256, 256, 256), capi.BOX, dt, 4, capi.KERNEL_GENERIC, 1, r=3)
if "line" in which:  # 1-D: fused register kernel vs one launch per step
    for dt, nm in ((capi.F32, "f32"), (capi.F64, "f64")):
        for r in (1, 4):
            for k in (1, 4, 8, 16):
                run(f"1d-r{r}-{nm}", (1 << 27,), capi.STAR, dt, 48, capi.KERNEL_LINE1D, k, r=r)
            run(f"1d-r{r}-{nm}", (1 << 27,), capi.STAR, dt, 8, capi.KERNEL_GENERIC, 1, r=r)
