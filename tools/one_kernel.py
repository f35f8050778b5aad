"""Run one BASELINE config a few passes (for ncu captures): tools/one_kernel.py c1|c2|c3|c4 [k]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from stencil_b200 import capi, sweep

CFG = {"c1": ((4096, 4096), capi.STAR, capi.F64, 4), "c1f32": ((4096, 4096), capi.STAR, capi.F32, 8),
       "c2": ((512, 512, 512), capi.STAR, capi.F64, 2), "c3": ((512, 512, 512), capi.BOX, capi.F64, 2),
       "c4": ((16384, 16384), capi.BOX, capi.F32, 6)}
name = sys.argv[1]
n, shape, dtype, k = CFG[name]
if len(sys.argv) > 2:
    k = int(sys.argv[2])
capi.check(capi.lib().sb200_init(0))
d = capi.make_desc(n, 1, shape, dtype, 4 * k, time_block=k)
g = sweep.DeviceGrids(d)
g.fill(capi.INIT_RANDOM, seed=1)
for _ in range(3):
    g.sweep(4 * k)
torch.cuda.synchronize()
print("ok", name, k)
