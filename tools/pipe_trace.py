"""Timeline of one pipelined sb200_run_host call (config 2 by default) from SB200_PIPELINE_TRACE:
per-stream busy time and gaps, when each parallelogram's sweeps start and end against its upload,
the slowest / fastest sweep launches.   python tools/pipe_trace.py [nx ny nz T]"""
import csv
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

from stencil_b200 import capi

n = tuple(int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (512, 512, 512)
T = int(sys.argv[4]) if len(sys.argv) > 4 else 100
L = capi.lib()
capi.check(L.sb200_init(0))
shape = tuple(v + 2 for v in reversed(n))
nbytes = int(np.prod(shape)) * 8
hp = [ctypes.c_void_p(), ctypes.c_void_p()]
host = []
for p in hp:
    capi.check(L.sb200_host_alloc(ctypes.byref(p), nbytes))
    h = np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_double)), shape=shape)
    h[:] = 0
    h[..., 0] = h[..., -1] = 1
    host.append(h)
rng = np.random.default_rng(1)
host[0][1:-1, 1:-1, 1:-1] = rng.random((n[2], n[1], n[0]))
d = capi.make_desc(n, 1, capi.STAR, capi.F64, T)
tm = capi.Timing()
capi.check(L.sb200_prepare(ctypes.byref(d)))
for _ in range(2):
    capi.check(L.sb200_run_host(ctypes.byref(d), hp[0], hp[1], ctypes.byref(tm)))
trace = os.path.join(ROOT, "gpurun_out", "pipe_trace.csv")
os.makedirs(os.path.dirname(trace), exist_ok=True)
os.environ["SB200_PIPELINE_TRACE"] = trace
capi.check(L.sb200_run_host(ctypes.byref(d), hp[0], hp[1], ctypes.byref(tm)))
print(f"call: pipeline {tm.sweep_ms:.2f} ms, upload span {tm.h2d_ms:.2f}, download span {tm.d2h_ms:.2f}, {tm.launches} sweep launches")
rows = list(csv.DictReader(open(trace)))
KIND = {0: "upload", 1: "unpack", 2: "sweep", 3: "download"}
for r in rows:
    for k in ("kind", "stream", "parallelogram", "block", "steps", "lo", "hi"):
        r[k] = int(r[k])
    r["b"], r["e"] = float(r["begin_ms"]), float(r["end_ms"])
for st in sorted({r["stream"] for r in rows}):
    ops = sorted((r for r in rows if r["stream"] == st), key=lambda r: r["b"])
    busy = sum(r["e"] - r["b"] for r in ops)
    print(f"stream {st}: {len(ops):4d} ops, first begin {ops[0]['b']:7.2f}, last end {ops[-1]['e']:7.2f}, busy {busy:7.2f} ms")
print("parallelogram: upload end | first sweep begin .. last sweep end (sum of launch times) | download begin .. end")
for p in sorted({r["parallelogram"] for r in rows}):
    up = [r for r in rows if r["parallelogram"] == p and r["kind"] == 0]
    sw = [r for r in rows if r["parallelogram"] == p and r["kind"] == 2]
    dn = [r for r in rows if r["parallelogram"] == p and r["kind"] == 3]
    print(f"  {p:2d}: {max((r['e'] for r in up), default=0):7.2f} | {min(r['b'] for r in sw):7.2f} .. {max(r['e'] for r in sw):7.2f} "
          f"({sum(r['e'] - r['b'] for r in sw):6.2f} in {len(sw)} launches, planes {sw[0]['hi'] - sw[0]['lo']}..{sw[-1]['hi'] - sw[-1]['lo']}) | "
          + (f"{dn[0]['b']:7.2f} .. {dn[-1]['e']:7.2f}" if dn else "-"))
sw = sorted((r for r in rows if r["kind"] == 2 and r["steps"] == 2), key=lambda r: r["e"] - r["b"])
per_plane = [(r["e"] - r["b"]) * 1e3 / (r["hi"] - r["lo"]) for r in sw if r["hi"] - r["lo"] > 20]
print(f"sweep launches (k=2): median {1e3 * (sw[len(sw) // 2]['e'] - sw[len(sw) // 2]['b']):.1f} us, "
      f"us per output plane: median {sorted(per_plane)[len(per_plane) // 2]:.2f}, p10 {sorted(per_plane)[len(per_plane) // 10]:.2f}, "
      f"p90 {sorted(per_plane)[9 * len(per_plane) // 10]:.2f}  (a whole-grid launch: 0.73 us per plane)")
for p in hp:
    L.sb200_host_free(p)
