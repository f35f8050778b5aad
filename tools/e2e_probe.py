"""e2e probe of sb200_run_host on config 2 (pinned host buffers): python tools/e2e_probe.py [planes-per-CTA]
(env SB200_PIPELINE_CHUNKS / SB200_PIPELINE_STREAMS: parallelograms / sweep streams; the optional
argument sets the z chunk one CTA of a range launch streams, via SB200_FLAG_BLOCK_TILE)"""
import ctypes
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from stencil_b200 import capi, slab

N, T = 512, 100
capi.check(capi.lib().sb200_init(0))
hint = int(sys.argv[1]) if len(sys.argv) > 1 else 0
desc = capi.make_desc((N, N, N), 1, capi.STAR, capi.F64, T, block_hint=hint, flags=capi.FLAG_BLOCK_TILE if hint else 0)
runner = slab.SlabSweep(desc, 0, 1)
nbytes = (N + 2) ** 3 * 8
hp = [ctypes.c_void_p(), ctypes.c_void_p()]
for p in hp:
    capi.check(capi.lib().sb200_host_alloc(ctypes.byref(p), nbytes))
ha = np.ctypeslib.as_array(ctypes.cast(hp[0], ctypes.POINTER(ctypes.c_double)), shape=(N + 2,) * 3)
hb = np.ctypeslib.as_array(ctypes.cast(hp[1], ctypes.POINTER(ctypes.c_double)), shape=(N + 2,) * 3)
runner.download_initial_into(ha, hb, seed=1)
tm = capi.Timing()
for rep in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    capi.check(capi.lib().sb200_run_host(ctypes.byref(desc), hp[0], hp[1], ctypes.byref(tm)))
    dt = time.perf_counter() - t0
    if rep:
        print(f"chunks={os.environ.get('SB200_PIPELINE_CHUNKS', 'auto')} streams={os.environ.get('SB200_PIPELINE_STREAMS', '2')} "
              f"planes/CTA={hint or 'planner'}: "
              f"wall {dt * 1e3:7.2f} ms  {N ** 3 * T / dt / 1e9:7.1f} GStencil/s  h2d {tm.h2d_ms:.1f} pipeline {tm.sweep_ms:.1f} "
              f"d2h-span {tm.d2h_ms:.1f} launches {tm.launches}", flush=True)
