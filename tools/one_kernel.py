"""Run one config a few passes (for ncu captures):
tools/one_kernel.py c1|c1f32|c2|c3|c4|c2gen|res512|rad2|rad3f64|box3f32|box3f64|line|linef64 [k]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from stencil_b200 import capi, sweep

CFG = {"c1": ((4096, 4096), capi.STAR, capi.F64, 4), "c1f32": ((4096, 4096), capi.STAR, capi.F32, 8),
       "c2": ((512, 512, 512), capi.STAR, capi.F64, 2), "c3": ((512, 512, 512), capi.BOX, capi.F64, 2),
       "c4": ((16384, 16384), capi.BOX, capi.F32, 4),
       "c2gen": ((512, 512, 512), capi.STAR, capi.F64, 1), "res512": ((512, 512), capi.STAR, capi.F32, 0),
       "rad2": ((8192, 8192), capi.STAR, capi.F32, 4), "rad3f64": ((8192, 8192), capi.STAR, capi.F64, 2),
       "box3f32": ((512, 512, 512), capi.BOX, capi.F32, 1), "box3f64": ((512, 512, 512), capi.BOX, capi.F64, 1),
       "line": ((1 << 27,), capi.STAR, capi.F32, 16), "linef64": ((1 << 27,), capi.STAR, capi.F64, 16)}
RADIUS = {"rad2": 2, "rad3f64": 3, "box3f32": 2, "box3f64": 2}
name = sys.argv[1]
n, shape, dtype, k = CFG[name]
if len(sys.argv) > 2:
    k = int(sys.argv[2])
capi.check(capi.lib().sb200_init(0))
kern = capi.KERNEL_GENERIC if name == "c2gen" else capi.KERNEL_RESIDENT2D if name == "res512" else capi.KERNEL_AUTO
steps = 2000 if name == "res512" else 4 * k
d = capi.make_desc(n, RADIUS.get(name, 1), shape, dtype, steps, time_block=k, kernel=kern)
g = sweep.DeviceGrids(d)
g.fill(capi.INIT_RANDOM, seed=1)
for _ in range(3):
    g.sweep(steps)
torch.cuda.synchronize()
print("ok", name, k)
