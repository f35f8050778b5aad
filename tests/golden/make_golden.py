"""Generate tests/golden/ref_2d_f32.npz (and ref_2d_f64.npz) from the GENUINE reference naive loop,
and ref_emu_2d_f32.npz from the reference's accelerator kernels run on the athread emulator.

Runs only where /root/reference is mounted (the build container): oracle/_ref/libref.so
is the reference's own src/stencil/stencil.cpp compiled by oracle/Makefile, and
oracle.ref_naive_2d_f32 recovers the full grid Stencil::check_result computes
(stencil.cpp:75-131).  The fixture travels to the GPU box; the reference does not.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import oracle  # noqa: E402

# (N, radius, T): sizes from the reference's own sweep (run_expr.py:6,14: N = 8*block,
# block 8 -> N 64) down to tiny/ragged grids; radii the CLI exposes (-r).
CASES = [
    (1, 1, 1), (2, 1, 3), (3, 1, 2), (5, 2, 4), (8, 1, 1), (8, 1, 2), (13, 1, 9),
    (16, 1, 10), (16, 1, 33), (24, 2, 7), (32, 3, 5), (40, 4, 9), (48, 1, 50),
    (64, 1, 100), (64, 2, 16), (64, 1, 1000),
]

# fp64: radius 3 (avg = 1/12 is not a dyadic rational) and long runs are where double and
# float arithmetic part ways
CASES64 = [
    (1, 1, 1), (3, 1, 2), (5, 2, 4), (13, 1, 9), (16, 1, 33), (24, 2, 7), (32, 3, 5), (40, 4, 9),
    (48, 3, 50), (64, 1, 100), (64, 3, 16), (64, 1, 1000),
]

# The reference's ACCELERATOR kernels on the athread emulator (oracle/ref_shim/emu/):
# (method, N = 8*block, radius, T).  Block sizes from run_expr.py's sweep (8 .. 120); only
# the DMA kernel takes a radius (stencil_dma_slave_pack.cpp:380-382, stencil_rma.cpp:466-471).
CASES_EMU = [
    ("DMA", 64, 1, 1), ("DMA", 64, 1, 100), ("DMA", 64, 2, 9), ("DMA", 128, 3, 16), ("DMA", 128, 4, 7),
    ("DMA", 256, 1, 200), ("DMASlavePack", 64, 1, 33), ("DMASlavePack", 128, 1, 100),
    ("RMA", 64, 1, 10), ("RMA", 64, 1, 1000), ("RMA", 192, 1, 101),
]

if __name__ == "__main__":
    oracle.build()
    # the reference's own BMP writer on its own result (N=20, T=30: even, so the field is in
    # `matrix`, the buffer the reference's to_bmp reads, stencil.cpp:159): pins --bmp
    bmp = os.path.join(os.path.dirname(__file__), "ref_bmp_N20_T30.bmp")
    oracle.ref_bmp_2d_f32(oracle.ref_naive_2d_f32(20, 1, 30), bmp)
    print("bmp", os.path.getsize(bmp), "bytes")
    out = {}
    for N, r, T in CASES:
        out[f"N{N}_r{r}_T{T}"] = oracle.ref_naive_2d_f32(N, r, T)
        print(N, r, T, "%.15f" % out[f"N{N}_r{r}_T{T}"][0, 0])
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "ref_2d_f32.npz"), **out)
    # the same reference sources compiled with float -> double (oracle/Makefile: libref64.so)
    out64 = {}
    for N, r, T in CASES64:
        out64[f"N{N}_r{r}_T{T}"] = oracle.ref_naive_2d_f64(N, r, T)
        print("f64", N, r, T, "%.17g" % out64[f"N{N}_r{r}_T{T}"][0, 0])
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "ref_2d_f64.npz"), **out64)
    # the reference's own DMA / DMASlavePack / RMA kernels, emulated CPE array
    emu = {}
    for m, N, r, T in CASES_EMU:
        g, ok = oracle.ref_emu_run_2d_f32(m, N, N // 8, r, T)
        assert ok, "the reference's own check_result rejected its kernel"
        emu[f"{m}_N{N}_r{r}_T{T}"] = g
        print(m, N, r, T, "%.15f" % g[0, 0])
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "ref_emu_2d_f32.npz"), **emu)
