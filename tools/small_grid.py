"""Small-grid regime of the reference's run_expr.py sweep (N = 8*block, fp32, r=1, star):
device-resident time per step for each N at large T.  python tools/small_grid.py [T]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from stencil_b200 import capi, sweep

T = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
capi.check(capi.lib().sb200_init(0))
for N in (64, 128, 256, 384, 400, 512, 640, 800, 960):
    for k in (0, 4, 8):  # 0 = AUTO (register-resident cluster kernel when the grid fits)
        d = capi.make_desc((N, N), 1, capi.STAR, capi.F32, T, time_block=k)
        g = sweep.DeviceGrids(d)
        g.fill(capi.INIT_RANDOM, seed=1)
        g.sweep(64)
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            nl = g.sweep(T)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        print(f"N={N:4d} k={k} T={T}: {best:9.3f} ms  {best / T * 1e6:8.1f} ns/step  {best / nl * 1e3:7.2f} us/launch  "
              f"{N * N * T / best / 1e6:8.1f} GStencil/s", flush=True)
