"""Would merging the pipeline's concurrently runnable range sweeps into one launch pay?  Device-resident,
512^3 fp64, k=2.  (a) today's shape: 51-plane range launches alternating on two streams, planner's chunking;
(b) merged-like: ONE launch per 51*m planes with 51-plane z chunks (SB200_FLAG_BLOCK_TILE), one stream.
Prints microseconds per output plane (a whole-grid launch: ~0.73)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from stencil_b200 import capi, sweep

capi.check(capi.lib().sb200_init(0))
L = capi.lib()
N, W = 512, 51
d0 = capi.make_desc((N, N, N), 1, capi.STAR, capi.F64, 2)
g = sweep.DeviceGrids(d0)
g.fill(capi.INIT_RANDOM, seed=1)
s = [torch.cuda.Stream(), torch.cuda.Stream()]


def rng(desc, lo, hi, stream):
    nl = ctypes.c_uint32(0)
    capi.check(L.sb200_sweep_device_range(ctypes.byref(desc), 0, 1, 1, g.a.data_ptr(), g.b.data_ptr(), 2, lo, hi,
                                          ctypes.c_void_p(stream.cuda_stream), ctypes.byref(nl)))


def timed(fn, planes, reps=20):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    for st in s:
        torch.cuda.current_stream().wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps / planes


def today(m):
    def f():
        for st in s:
            st.wait_stream(torch.cuda.current_stream())
        for c in range(m):
            rng(d0, c * W, (c + 1) * W, s[c & 1])
    return f


def merged(m, chunk):
    dm = capi.make_desc((N, N, N), 1, capi.STAR, capi.F64, 2, block_hint=chunk, flags=capi.FLAG_BLOCK_TILE)

    def f():
        s[0].wait_stream(torch.cuda.current_stream())
        rng(dm, 0, m * W, s[0])
    return f


print(f"whole grid, planner: {timed(lambda: rng(d0, 0, N, s[0]), N):.3f} us/plane")
for m in (2, 3, 4, 5, 6, 8, 10):
    t = timed(today(m), m * W)
    mm = timed(merged(m, W), m * W)
    m2 = timed(merged(m, 2 * W), m * W)
    mp = timed(lambda m=m: rng(d0, 0, m * W, s[0]), m * W)
    print(f"m={m:2d} ({m * W:3d} planes): today {t:.3f}  one launch, 51-plane chunks {mm:.3f}  102-plane chunks {m2:.3f}  "
          f"one launch, planner's chunks {mp:.3f} us/plane", flush=True)
