"""CPU oracle bindings (TEST INFRASTRUCTURE — never imported by stencil_b200/).

ctypes wrappers over ``oracle/liboracle.so`` (the C restatement, stencil_oracle.c, which
follows /root/reference/src/stencil/stencil.cpp:75-131,190-207) and, when it was built,
``oracle/_ref/libref.so`` (the reference's own stencil.cpp compiled against stub vendor
headers, see oracle/ref_shim/).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this package.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
STAR, BOX = 0, 1
F32, F64 = 0, 1
INIT_REF, INIT_RANDOM = 0, 1
_NP = {F32: np.float32, F64: np.float64}


def build(quiet: bool = True) -> None:
    """Compile liboracle.so (and _ref/libref.so when /root/reference is present)."""
    subprocess.run(["make", "-C", _HERE], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


_lib = None
_ref = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = ctypes.CDLL(path)
        u64p = ctypes.POINTER(ctypes.c_uint64)
        L.oracle_max_threads.restype = ctypes.c_int
        L.oracle_mix64.restype = ctypes.c_uint64
        L.oracle_mix64.argtypes = [ctypes.c_uint64]
        L.oracle_mix64_seeded.restype = ctypes.c_uint64
        L.oracle_mix64_seeded.argtypes = [ctypes.c_uint64, ctypes.c_uint64]
        L.oracle_padded_elems.restype = ctypes.c_size_t
        L.oracle_padded_elems.argtypes = [ctypes.c_int, u64p, ctypes.c_uint]
        for sfx in ("f32", "f64"):
            f = getattr(L, "oracle_fill_" + sfx)
            f.restype = None
            f.argtypes = [ctypes.c_int, u64p, ctypes.c_uint, ctypes.c_int, ctypes.c_uint64,
                          ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p]
            s = getattr(L, "oracle_sweep_" + sfx)
            s.restype = ctypes.c_int
            s.argtypes = [ctypes.c_int, u64p, ctypes.c_uint, ctypes.c_int, ctypes.c_uint,
                          ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
            c = getattr(L, "oracle_checksum_" + sfx)
            c.restype = ctypes.c_uint64
            c.argtypes = [ctypes.c_int, u64p, ctypes.c_uint, ctypes.c_void_p,
                          ctypes.c_uint64, ctypes.c_uint64]
        _lib = L
    return _lib


def ref_lib():
    """The genuine reference build, or None if oracle/_ref/libref.so is absent."""
    global _ref
    if _ref is None:
        path = os.path.join(_HERE, "_ref", "libref.so")
        if not os.path.exists(path):
            return None
        R = ctypes.CDLL(path)
        R.ref_naive_2d_f32.restype = ctypes.c_int
        R.ref_naive_2d_f32.argtypes = [ctypes.c_uint] * 3 + [ctypes.c_void_p]
        R.ref_check_2d_f32.restype = ctypes.c_int
        R.ref_check_2d_f32.argtypes = [ctypes.c_uint] * 3 + [ctypes.c_void_p]
        R.ref_time_check_2d_f32.restype = ctypes.c_double
        R.ref_time_check_2d_f32.argtypes = [ctypes.c_uint] * 3
        if hasattr(R, "ref_bmp_2d_f32"):
            R.ref_bmp_2d_f32.restype = ctypes.c_int
            R.ref_bmp_2d_f32.argtypes = [ctypes.c_uint, ctypes.c_void_p, ctypes.c_char_p]
        _ref = R
    return _ref


def _n3(n):
    n = tuple(int(v) for v in n)
    arr = (ctypes.c_uint64 * 3)(*(n + (1,) * (3 - len(n))))
    return len(n), arr


def padded_shape(n, r):
    """numpy shape (z, y, x order, outermost first) of the padded grid; n = (nx[,ny[,nz]])."""
    return tuple(int(v) + 2 * r for v in reversed(n))


def max_threads() -> int:
    return lib().oracle_max_threads()


def fill(n, r, dtype=F32, init=INIT_REF, seed=0, z_offset=0, z_total=0) -> np.ndarray:
    ndim, n3 = _n3(n)
    g = np.empty(padded_shape(n, r), dtype=_NP[dtype])
    fn = lib().oracle_fill_f32 if dtype == F32 else lib().oracle_fill_f64
    fn(ndim, n3, r, init, seed, z_offset, z_total, g.ctypes.data)
    return g


def sweep(n, r, shape, T, a: np.ndarray, b: np.ndarray, threads: int = 1) -> int:
    """T double-buffered sweeps a<->b in place; returns 1 if step T is in b else 0."""
    assert a.dtype == b.dtype and a.flags.c_contiguous and b.flags.c_contiguous
    ndim, n3 = _n3(n)
    fn = lib().oracle_sweep_f32 if a.dtype == np.float32 else lib().oracle_sweep_f64
    return fn(ndim, n3, r, shape, T, a.ctypes.data, b.ctypes.data, threads)


def run(n, r, shape, T, dtype=F32, init=INIT_REF, seed=0, threads: int = 1) -> np.ndarray:
    """Initialise both buffers like Stencil::generate_initialized_matrix and return the
    padded grid holding step T."""
    a = fill(n, r, dtype, init, seed)
    b = fill(n, r, dtype, INIT_REF, 0)
    return b if sweep(n, r, shape, T, a, b, threads) else a


def checksum(n, r, g: np.ndarray, z_offset=0, z_total=0) -> int:
    ndim, n3 = _n3(n)
    fn = lib().oracle_checksum_f32 if g.dtype == np.float32 else lib().oracle_checksum_f64
    return int(fn(ndim, n3, r, g.ctypes.data, z_offset, z_total))


def interior(g: np.ndarray, r: int) -> np.ndarray:
    return g[tuple(slice(r, s - r) for s in g.shape)]


def ref_naive_2d_f32(N: int, r: int, T: int) -> np.ndarray:
    """Interior N×N result of the GENUINE reference naive loop (needs _ref/libref.so)."""
    R = ref_lib()
    if R is None:
        raise RuntimeError("oracle/_ref/libref.so not built (needs /root/reference)")
    out = np.empty((N, N), dtype=np.float32)
    if R.ref_naive_2d_f32(N, r, T, out.ctypes.data) != 0:
        raise RuntimeError("reference probe protocol failed")
    return out


def ref_bmp_2d_f32(grid: np.ndarray, path: str) -> None:
    """The GENUINE reference heat-map writer (Stencil::to_bmp + BMPImage::save) on `grid`."""
    R = ref_lib()
    if R is None:
        raise RuntimeError("oracle/_ref/libref.so not built (needs /root/reference)")
    g = np.ascontiguousarray(grid, dtype=np.float32)
    assert g.ndim == 2 and g.shape[0] == g.shape[1]
    if R.ref_bmp_2d_f32(g.shape[0], g.ctypes.data, path.encode()) != 0:
        raise RuntimeError("reference BMP writer failed")


_ref64 = None


def ref64_lib():
    """The reference's own sources compiled with float -> double (oracle/Makefile:
    _ref/libref64.so), or None if absent."""
    global _ref64
    if _ref64 is None:
        path = os.path.join(_HERE, "_ref", "libref64.so")
        if not os.path.exists(path):
            return None
        R = ctypes.CDLL(path)
        R.ref_naive_2d_f64.restype = ctypes.c_int
        R.ref_naive_2d_f64.argtypes = [ctypes.c_uint] * 3 + [ctypes.c_void_p]
        R.ref_check_2d_f64.restype = ctypes.c_int
        R.ref_check_2d_f64.argtypes = [ctypes.c_uint] * 3 + [ctypes.c_void_p]
        _ref64 = R
    return _ref64


def ref_naive_2d_f64(N: int, r: int, T: int) -> np.ndarray:
    """Interior N×N result of the reference naive loop compiled at double precision."""
    R = ref64_lib()
    if R is None:
        raise RuntimeError("oracle/_ref/libref64.so not built (needs /root/reference)")
    out = np.empty((N, N), dtype=np.float64)
    if R.ref_naive_2d_f64(N, r, T, out.ctypes.data) != 0:
        raise RuntimeError("reference probe protocol failed")
    return out


_emu = None
EMU_METHODS = ("DMA", "DMASlavePack", "RMA")


def ref_emu_lib():
    """The reference's accelerator kernels on the athread emulator (oracle/Makefile:
    _ref/libref_emu.so, oracle/ref_shim/emu/), or None if absent."""
    global _emu
    if _emu is None:
        path = os.path.join(_HERE, "_ref", "libref_emu.so")
        if not os.path.exists(path):
            return None
        R = ctypes.CDLL(path)
        R.ref_emu_run_2d_f32.restype = ctypes.c_int
        R.ref_emu_run_2d_f32.argtypes = [ctypes.c_char_p] + [ctypes.c_uint] * 4 + \
            [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
        _emu = R
    return _emu


def ref_emu_run_2d_f32(method: str, N: int, block: int, r: int, T: int):
    """Interior N×N result of the reference's own DMA / DMASlavePack / RMA kernel
    (src/stencil/slave/*.cpp) run on the emulated 8x8 CPE array, and the reference's own
    check_result() verdict on it.  N must be 8 * block."""
    R = ref_emu_lib()
    if R is None:
        raise RuntimeError("oracle/_ref/libref_emu.so not built (needs /root/reference)")
    out = np.empty((N, N), dtype=np.float32)
    ok = ctypes.c_int(-1)
    rc = R.ref_emu_run_2d_f32(method.encode(), N, block, r, T, out.ctypes.data, ctypes.byref(ok))
    if rc == -2:
        raise RuntimeError("athread emulator: the emulated CPE stacks are not congruent on this host")
    if rc != 0:
        raise ValueError(f"emulated reference kernel cannot run {method} N={N} block={block} r={r} T={T}")
    return out, bool(ok.value)
