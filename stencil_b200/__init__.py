"""stencil_b200 — B200 (sm_100a) backend for the iterative stencil sweep of Pluto-Zy/stencil.

Layout: csrc/ (CUDA kernels + the C ABI of include/stencil_b200.h), host/ (the C++17 host
side mirroring the reference's Stencil / ProgramOptions / main), capi.py (ctypes binding
used by tests and bench.py), sweep.py (numpy / torch conveniences over the C ABI),
slab.py (one-process-per-GPU slab decomposition with halo exchange).
"""
from . import capi  # noqa: F401
from .capi import (BOX, F32, F64, INIT_RANDOM, INIT_REF, KERNEL_AUTO, KERNEL_GENERIC,  # noqa: F401
                   KERNEL_TMA3D, KERNEL_WARP2D, STAR, StencilError, make_desc)
