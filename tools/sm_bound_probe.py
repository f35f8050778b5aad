"""SM-side rate of the tma3d kernels without the HBM limit (developer probe, run under gpurun).

Builds nothing itself: tools/sm_bound_probe.sh compiles a SECOND library with
-DSB200_PROBE_ZWRAP=8 (every CTA streams the same 8 planes, so loads and stores stay in L2;
results are garbage) and this script times it next to the product library's numbers."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from stencil_b200 import capi

if len(sys.argv) > 1 and sys.argv[1] == "probe":
    capi.LIB_PATH = os.path.join(ROOT, "stencil_b200", "libstencil_b200_probe.so")
from stencil_b200 import sweep  # noqa: E402

capi.check(capi.lib().sb200_init(0))


def run(name, n, shape, dtype, T, k):
    d = capi.make_desc(n, 1, shape, dtype, T, time_block=k, kernel=capi.KERNEL_TMA3D)
    g = sweep.DeviceGrids(d)
    g.fill(capi.INIT_RANDOM, seed=1)
    g.sweep(2 * k)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.sweep(T)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    pts = n[0] * n[1] * n[2]
    print(f"{sys.argv[1] if len(sys.argv) > 1 else 'product':8s} {name:12s} k={k}: {best:8.3f} ms "
          f"{pts * T / best / 1e6:8.1f} GStencil/s", flush=True)


for k in (1, 2):
    run("7pt-f64", (512, 512, 512), capi.STAR, capi.F64, 24, k)
    run("7pt-f32", (512, 512, 512), capi.STAR, capi.F32, 24, k)
    run("27pt-f64", (512, 512, 512), capi.BOX, capi.F64, 12, k)
    run("27pt-f32", (512, 512, 512), capi.BOX, capi.F32, 12, k)
