"""Time config-2-like sweeps with a variant library: python tools/variant_probe.py <tag|product> [cfg...]
(cfg: c2 c2k1 c3 c2f32)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from stencil_b200 import capi

tag = sys.argv[1] if len(sys.argv) > 1 else "product"
if tag != "product":
    capi.LIB_PATH = os.path.join(ROOT, "stencil_b200", f"libstencil_b200_{tag}.so")
from stencil_b200 import sweep  # noqa: E402

capi.check(capi.lib().sb200_init(0))
CFG = {"c2": ((512, 512, 512), capi.STAR, capi.F64, 2, 48), "c2k1": ((512, 512, 512), capi.STAR, capi.F64, 1, 24),
       "c2k3": ((512, 512, 512), capi.STAR, capi.F64, 3, 48),
       "c3": ((512, 512, 512), capi.BOX, capi.F64, 2, 24), "c3k1": ((512, 512, 512), capi.BOX, capi.F64, 1, 12),
       "c2f32": ((512, 512, 512), capi.STAR, capi.F32, 2, 48), "c3f32": ((512, 512, 512), capi.BOX, capi.F32, 2, 24),
       "c5": ((1024, 1024, 256), capi.STAR, capi.F64, 2, 24),
       # SB200_FLAG_REASSOCIATE: separable 27-point sums (tolerance-level parity)
       "c3sep": ((512, 512, 512), capi.BOX, capi.F64, 2, 24), "c3sepk1": ((512, 512, 512), capi.BOX, capi.F64, 1, 12),
       "c3sepf32": ((512, 512, 512), capi.BOX, capi.F32, 2, 24)}
for name in sys.argv[2:] or ["c2"]:
    n, shape, dtype, k, T = CFG[name]
    d = capi.make_desc(n, 1, shape, dtype, T, time_block=k, kernel=capi.KERNEL_TMA3D,
                       flags=capi.FLAG_REASSOCIATE if "sep" in name else 0)
    g = sweep.DeviceGrids(d)
    g.fill(capi.INIT_RANDOM, seed=1)
    g.sweep(2 * k)
    torch.cuda.synchronize()
    cs0 = None
    times = []
    for rep in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.sweep(T)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    pts = n[0] * n[1] * n[2]
    g.fill(capi.INIT_RANDOM, seed=1)
    g.sweep(T)
    print(f"{tag:10s} {name:6s} k={k}: best {min(times):8.3f} ms {pts * T / min(times) / 1e6:8.1f} GStencil/s  "
          f"median {sorted(times)[2]:8.3f} ms  checksum {g.checksum():#018x}", flush=True)
    del g
    torch.cuda.empty_cache()
