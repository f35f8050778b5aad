"""Single-GPU probe of the linked slab launch: the slab's neighbours are the slab itself
(periodic in z), so the whole face path — peer stores, arrival counters, waits — runs on one
GPU and can be timed / profiled without NVLink in the picture.
    python tools/self_link.py [nx ny nz] [--faces 0|1|2]"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from stencil_b200 import capi, slab

args = [a for a in sys.argv[1:] if not a.startswith("--")]
n = tuple(int(a) for a in args[:3]) if len(args) >= 3 else (512, 512, 512)
faces = 2
for a in sys.argv[1:]:
    if a.startswith("--faces="):
        faces = int(a.split("=")[1])
K, T = 2, 40
capi.check(capi.lib().sb200_init(0))
L = capi.lib()
d = capi.make_desc(n, 1, capi.STAR, capi.F64, T, time_block=K)
ghost = K
lay = capi.layout_for(d, ghost)
bufs = [slab._LibBuffer(lay.bytes) for _ in range(2)]
flags = slab._LibBuffer(16)
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for i, b in enumerate(bufs):
    capi.check(L.sb200_fill_device(ctypes.byref(d), ghost, b.data_ptr(), capi.INIT_RANDOM if i == 0 else capi.INIT_REF, 1, 0, 0, st))
assert faces == 2, "a slab that is its own neighbour has two internal faces"
lo_b = hi_b = 0


def run(linked):
    arrivals, cur = run.arr, 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(T // K):
        if linked:
            lk = capi.SlabLink()
            lk.peer_dst_lo = lk.peer_dst_hi = bufs[cur ^ 1].data_ptr()
            lk.flags_local = lk.flags_lo = lk.flags_hi = flags.data_ptr()
            lk.wait_arrivals = arrivals
            arr = ctypes.c_uint32(0)
            capi.check(L.sb200_sweep_slab_linked(ctypes.byref(d), ghost, lo_b, hi_b, bufs[cur].data_ptr(),
                                                 bufs[cur ^ 1].data_ptr(), K, ctypes.byref(lk), st, ctypes.byref(arr)))
            arrivals += arr.value
        else:
            nl = ctypes.c_uint32(0)
            capi.check(L.sb200_sweep_device_range(ctypes.byref(d), ghost, lo_b, hi_b, bufs[cur].data_ptr(),
                                                  bufs[cur ^ 1].data_ptr(), K, 0, n[2], st, ctypes.byref(nl)))
        cur ^= 1
    e1.record()
    torch.cuda.synchronize()
    run.arr = arrivals
    return e0.elapsed_time(e1)


run.arr = 0
for linked in (False, True, False, True):
    run(linked)
    ms = min(run(linked) for _ in range(3))
    pts = n[0] * n[1] * n[2]
    print(f"n={n} faces={faces} linked={linked}: {ms / (T // K) * 1e3:8.1f} us/launch  {pts * T / ms / 1e6:8.1f} GStencil/s", flush=True)
w = (ctypes.c_uint32 * 4)()
capi.check(L.sb200_link_status(flags.data_ptr(), w))
print("flags", list(w))
