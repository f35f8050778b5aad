"""numpy / torch conveniences over the C ABI (no arithmetic happens here).

``run_host`` is the Python spelling of the drop-in call that replaces the reference's
``athread_spawn``/``athread_join`` pair (src/stencil/stencil.cpp:33-54): padded host
arrays in, T double-buffered sweeps on the GPU, step T lands in ``out`` if T is odd else
in ``in`` (stencil.cpp:88-92,129-134).
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import capi

_NP = {capi.F32: np.float32, capi.F64: np.float64}


def dtype_code(dt) -> int:
    return capi.F64 if np.dtype(dt) == np.float64 else capi.F32


def padded_shape(n, r):
    return tuple(int(v) + 2 * r for v in reversed(tuple(n)))


def run_host(n, a: np.ndarray, b: np.ndarray, iterations: int, radius=1, shape=capi.STAR,
             time_block=0, kernel=capi.KERNEL_AUTO, ngpus=0, flags=0):
    """In-place on the two padded host arrays; returns (result_array, Timing)."""
    assert a.dtype == b.dtype and a.flags.c_contiguous and b.flags.c_contiguous
    assert a.shape == padded_shape(n, radius) == b.shape, (a.shape, padded_shape(n, radius))
    d = capi.make_desc(n, radius, shape, dtype_code(a.dtype), iterations, time_block, 0, kernel, ngpus, flags)
    t = capi.Timing()
    capi.check(capi.lib().sb200_run_host(ctypes.byref(d), a.ctypes.data, b.ctypes.data,
                                         ctypes.byref(t)))
    return (b if iterations & 1 else a), t


class DeviceGrids:
    """Two device buffers (torch-owned memory) in the library's device layout."""

    def __init__(self, desc: capi.Desc, ghost: int = 0, device=None):
        import torch
        self.torch = torch
        self.desc, self.ghost = desc, ghost
        self.layout = capi.layout_for(desc, ghost)
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.a = torch.zeros(self.layout.bytes, dtype=torch.uint8, device=self.device)
        self.b = torch.zeros(self.layout.bytes, dtype=torch.uint8, device=self.device)
        self.cur = 0  # which buffer holds the current state

    def _stream(self):
        return ctypes.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def bufs(self):
        return (self.a, self.b) if self.cur == 0 else (self.b, self.a)

    def fill(self, init=capi.INIT_REF, seed=0, z_offset=0, z_total=0):
        L = capi.lib()
        for i, t in enumerate((self.a, self.b)):
            # only the current-state buffer carries the random interior; the other one
            # only needs the halo (reference: both matrices share the boundary)
            capi.check(L.sb200_fill_device(ctypes.byref(self.desc), self.ghost, t.data_ptr(),
                                           init if i == 0 else capi.INIT_REF, seed, z_offset,
                                           z_total, self._stream()))
        self.cur = 0

    def upload(self, host: np.ndarray, which=0):
        t = (self.a, self.b)[which]
        capi.check(capi.lib().sb200_upload(ctypes.byref(self.desc), self.ghost, host.ctypes.data,
                                           t.data_ptr(), self._stream()))
        self.torch.cuda.current_stream(self.device).synchronize()

    def download(self, which=None) -> np.ndarray:
        d = self.desc
        n = [int(d.n[i]) for i in range(d.ndim)]
        r = int(d.radius)
        shp = list(padded_shape(n, r))
        shp[0] += 2 * self.ghost if d.ndim > 1 else 0
        host = np.empty(shp, dtype=_NP[d.dtype])
        t = self.bufs()[0] if which is None else (self.a, self.b)[which]
        capi.check(capi.lib().sb200_download(ctypes.byref(d), self.ghost, t.data_ptr(),
                                             host.ctypes.data, self._stream()))
        self.torch.cuda.current_stream(self.device).synchronize()
        return host

    def sweep(self, steps: int, lo_boundary=1, hi_boundary=1) -> int:
        """Enqueue `steps` steps on the current torch stream; returns kernels launched."""
        src, dst = self.bufs()
        in_b = ctypes.c_int(0)
        nl = ctypes.c_uint32(0)
        capi.check(capi.lib().sb200_sweep_device(
            ctypes.byref(self.desc), self.ghost, lo_boundary, hi_boundary, src.data_ptr(),
            dst.data_ptr(), steps, self._stream(), ctypes.byref(in_b), ctypes.byref(nl)))
        if in_b.value:
            self.cur ^= 1
        return nl.value

    def checksum(self, z_offset=0, z_total=0) -> int:
        out = ctypes.c_uint64(0)
        capi.check(capi.lib().sb200_checksum_device(
            ctypes.byref(self.desc), self.ghost, self.bufs()[0].data_ptr(), z_offset, z_total,
            self._stream(), ctypes.byref(out)))
        return out.value
