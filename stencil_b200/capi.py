"""ctypes binding of the C ABI in include/stencil_b200.h (libstencil_b200.so).

This is the Python-side stub a maintainer would use to reach the same entry points the
C++ host (stencil_b200/host/) calls; tests and bench.py go through it.  There is NO
fallback: if the shared library is missing or no GPU is present, calls raise.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libstencil_b200.so")

STAR, BOX = 0, 1
F32, F64 = 0, 1
INIT_REF, INIT_RANDOM = 0, 1
KERNEL_AUTO, KERNEL_GENERIC, KERNEL_WARP2D, KERNEL_TMA3D, KERNEL_RESIDENT2D, KERNEL_LINE1D = 0, 1, 2, 3, 4, 5
OK, ERR_INVALID, ERR_CUDA, ERR_UNSUPPORTED, ERR_NO_DEVICE = 0, -1, -2, -3, -4

# every symbol include/stencil_b200.h declares
SYMBOLS = [
    "sb200_abi_version", "sb200_last_error", "sb200_device_count", "sb200_init",
    "sb200_finalize", "sb200_run_host", "sb200_layout_for", "sb200_sweep_device",
    "sb200_sweep_device_range", "sb200_fill_device", "sb200_upload", "sb200_download",
    "sb200_checksum_device", "sb200_host_alloc", "sb200_host_free", "sb200_max_time_block",
    "sb200_device_alloc", "sb200_device_free", "sb200_ipc_export", "sb200_ipc_open",
    "sb200_ipc_close", "sb200_sweep_slab_linked", "sb200_link_status",
    "sb200_host_register", "sb200_host_unregister", "sb200_plan", "sb200_pipeline_schedule",
    "sb200_init_multi", "sb200_group_size", "sb200_prepare", "sb200_slab_plan_for",
    "sb200_pipeline_schedule_slab", "sb200_link_reset", "sb200_host_alloc_for", "sb200_host_place",
    "sb200_checksum_device_noalloc", "sb200_group_sweep",
]
NOALLOC_MAX_POINTS = 8192 * 16384
MAX_GPUS = 16
FLAG_REASSOCIATE = 1
FLAG_BLOCK_TILE = 2
PIPE_UPLOAD, PIPE_UNPACK, PIPE_SWEEP, PIPE_DOWNLOAD = 0, 1, 2, 3
IPC_HANDLE_BYTES = 64
LINK_FLAG_WORDS = 4


class Desc(ctypes.Structure):
    _fields_ = [
        ("ndim", ctypes.c_int32), ("shape", ctypes.c_int32), ("dtype", ctypes.c_int32),
        ("radius", ctypes.c_uint32), ("n", ctypes.c_uint64 * 3),
        ("iterations", ctypes.c_uint32), ("time_block", ctypes.c_uint32),
        ("block_hint", ctypes.c_uint32), ("kernel", ctypes.c_int32),
        ("ngpus", ctypes.c_uint32), ("flags", ctypes.c_uint32),
    ]


class Timing(ctypes.Structure):
    _fields_ = [
        ("h2d_ms", ctypes.c_double), ("sweep_ms", ctypes.c_double), ("d2h_ms", ctypes.c_double),
        ("wall_ms", ctypes.c_double), ("launches", ctypes.c_uint32),
        ("time_block", ctypes.c_uint32), ("kernel", ctypes.c_int32),
        ("result_in_out", ctypes.c_int32), ("ngpus", ctypes.c_uint32), ("pipelined", ctypes.c_uint32),
    ]


class Layout(ctypes.Structure):
    _fields_ = [
        ("pitch_x", ctypes.c_uint64), ("pitch_y", ctypes.c_uint64), ("planes", ctypes.c_uint64),
        ("ox", ctypes.c_uint64), ("oy", ctypes.c_uint64), ("oz", ctypes.c_uint64),
        ("bytes", ctypes.c_uint64),
    ]


class SlabLink(ctypes.Structure):
    _fields_ = [
        ("peer_dst_lo", ctypes.c_void_p), ("peer_dst_hi", ctypes.c_void_p),
        ("flags_local", ctypes.c_void_p), ("flags_lo", ctypes.c_void_p),
        ("flags_hi", ctypes.c_void_p), ("wait_arrivals", ctypes.c_uint32),
    ]


class PlanInfo(ctypes.Structure):
    _fields_ = [
        ("kernel", ctypes.c_int32), ("time_block", ctypes.c_uint32), ("launches", ctypes.c_uint32),
        ("tiles_x", ctypes.c_uint32), ("tiles_y", ctypes.c_uint32), ("chunks", ctypes.c_uint32),
        ("chunk_extent", ctypes.c_uint32), ("pipeline_chunks", ctypes.c_uint32),
    ]


class _Slab(ctypes.Structure):
    _fields_ = [("sub_lo", ctypes.c_uint64), ("sub_hi", ctypes.c_uint64),
                ("own_lo", ctypes.c_uint64), ("own_hi", ctypes.c_uint64)]


class SlabPlan(ctypes.Structure):
    _fields_ = [("ngpus_used", ctypes.c_uint32), ("deep_halo", ctypes.c_uint32), ("slab", _Slab * MAX_GPUS)]


class GroupResult(ctypes.Structure):
    _fields_ = [("sweep_ms", ctypes.c_double), ("checksum", ctypes.c_uint64), ("ngpus", ctypes.c_uint32),
                ("time_block", ctypes.c_uint32), ("launches", ctypes.c_uint32), ("kernel", ctypes.c_int32),
                ("exchange", ctypes.c_int32)]


class PipeOp(ctypes.Structure):
    _fields_ = [
        ("kind", ctypes.c_int32), ("stream", ctypes.c_int32), ("parallelogram", ctypes.c_int32),
        ("block", ctypes.c_int32), ("steps", ctypes.c_uint32), ("src", ctypes.c_int32),
        ("lo", ctypes.c_int64), ("hi", ctypes.c_int64), ("wait", ctypes.c_int32 * 2),
    ]


class StencilError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"stencil_b200 error {code}: {msg}")
        self.code = code


_lib = None


def lib() -> ctypes.CDLL:
    """Load libstencil_b200.so (fails loudly if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build the CUDA extension first "
                "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        vp, u32, u64, i32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_int
        dp = ctypes.POINTER(Desc)
        L.sb200_abi_version.restype = i32
        L.sb200_last_error.restype = ctypes.c_char_p
        L.sb200_device_count.restype = i32
        L.sb200_init.argtypes = [i32]
        L.sb200_init_multi.argtypes = [i32, ctypes.POINTER(i32)]
        L.sb200_group_size.restype = i32
        L.sb200_prepare.argtypes = [dp]
        L.sb200_group_sweep.argtypes = [dp, i32, u64, u32, ctypes.POINTER(GroupResult)]
        L.sb200_slab_plan_for.argtypes = [dp, i32, ctypes.POINTER(SlabPlan)]
        L.sb200_pipeline_schedule_slab.argtypes = [dp, i32, i32, ctypes.POINTER(PipeOp), u32,
                                                   ctypes.POINTER(u32), ctypes.POINTER(u32)]
        L.sb200_finalize.restype = None
        L.sb200_run_host.argtypes = [dp, vp, vp, ctypes.POINTER(Timing)]
        L.sb200_layout_for.argtypes = [dp, u32, ctypes.POINTER(Layout)]
        L.sb200_sweep_device.argtypes = [dp, u32, i32, i32, vp, vp, u32, vp,
                                         ctypes.POINTER(i32), ctypes.POINTER(u32)]
        L.sb200_sweep_device_range.argtypes = [dp, u32, i32, i32, vp, vp, u32, u64, u64, vp,
                                               ctypes.POINTER(u32)]
        L.sb200_fill_device.argtypes = [dp, u32, vp, i32, u64, u64, u64, vp]
        L.sb200_upload.argtypes = [dp, u32, vp, vp, vp]
        L.sb200_download.argtypes = [dp, u32, vp, vp, vp]
        L.sb200_checksum_device.argtypes = [dp, u32, vp, u64, u64, vp, ctypes.POINTER(u64)]
        L.sb200_checksum_device_noalloc.argtypes = [dp, u32, vp, u64, u64, vp, ctypes.POINTER(u64)]
        L.sb200_host_alloc.argtypes = [ctypes.POINTER(vp), ctypes.c_size_t]
        L.sb200_host_free.argtypes = [vp]
        L.sb200_host_alloc_for.argtypes = [dp, ctypes.POINTER(vp), ctypes.POINTER(i32)]
        L.sb200_host_place.argtypes = [dp, vp, ctypes.POINTER(i32)]
        L.sb200_host_register.argtypes = [vp, ctypes.c_size_t]
        L.sb200_host_unregister.argtypes = [vp]
        L.sb200_plan.argtypes = [dp, u32, i32, ctypes.POINTER(PlanInfo)]
        L.sb200_max_time_block.argtypes = [dp]
        L.sb200_pipeline_schedule.argtypes = [dp, ctypes.POINTER(PipeOp), u32, ctypes.POINTER(u32)]
        L.sb200_device_alloc.argtypes = [ctypes.POINTER(vp), ctypes.c_size_t]
        L.sb200_device_free.argtypes = [vp]
        L.sb200_ipc_export.argtypes = [vp, ctypes.c_char_p]
        L.sb200_ipc_open.argtypes = [ctypes.c_char_p, ctypes.POINTER(vp)]
        L.sb200_ipc_close.argtypes = [vp]
        L.sb200_link_status.argtypes = [vp, ctypes.POINTER(u32)]
        L.sb200_link_reset.argtypes = [vp]
        L.sb200_sweep_slab_linked.argtypes = [dp, u32, i32, i32, vp, vp, u32,
                                              ctypes.POINTER(SlabLink), vp, ctypes.POINTER(u32)]
        _lib = L
    return _lib


def check(rc: int) -> int:
    if rc < 0:
        raise StencilError(rc, lib().sb200_last_error().decode())
    return rc


def make_desc(n, radius=1, shape=STAR, dtype=F32, iterations=1, time_block=0, block_hint=0,
              kernel=KERNEL_AUTO, ngpus=0, flags=0) -> Desc:
    """n = (nx,), (nx, ny) or (nx, ny, nz) interior extents, x fastest."""
    n = tuple(int(v) for v in n)
    d = Desc()
    d.ndim = len(n)
    d.shape, d.dtype, d.radius = shape, dtype, radius
    for a in range(3):
        d.n[a] = n[a] if a < len(n) else 1
    d.iterations, d.time_block, d.block_hint, d.kernel = iterations, time_block, block_hint, kernel
    d.ngpus, d.flags = ngpus, flags
    return d


def layout_for(desc: Desc, ghost: int = 0) -> Layout:
    lay = Layout()
    check(lib().sb200_layout_for(ctypes.byref(desc), ghost, ctypes.byref(lay)))
    return lay


def pipeline_schedule(desc: Desc) -> list:
    """The op list sb200_run_host would execute for this descriptor ([] if it does not
    pipeline the call); host logic only."""
    n = ctypes.c_uint32(0)
    check(lib().sb200_pipeline_schedule(ctypes.byref(desc), None, 0, ctypes.byref(n)))
    if n.value == 0:
        return []
    ops = (PipeOp * n.value)()
    check(lib().sb200_pipeline_schedule(ctypes.byref(desc), ops, n.value, ctypes.byref(n)))
    return list(ops)


def slab_plan(desc: Desc, ngpus: int) -> SlabPlan:
    """How sb200_run_host would spread `desc` over `ngpus` devices (host logic only)."""
    p = SlabPlan()
    check(lib().sb200_slab_plan_for(ctypes.byref(desc), ngpus, ctypes.byref(p)))
    return p


def pipeline_schedule_slab(desc: Desc, ngpus: int, slab: int):
    """(ops, zone_uploads) of one slab of a multi-GPU sb200_run_host; host logic only."""
    n, z = ctypes.c_uint32(0), ctypes.c_uint32(0)
    check(lib().sb200_pipeline_schedule_slab(ctypes.byref(desc), ngpus, slab, None, 0, ctypes.byref(n), ctypes.byref(z)))
    if n.value == 0:
        return [], 0
    ops = (PipeOp * n.value)()
    check(lib().sb200_pipeline_schedule_slab(ctypes.byref(desc), ngpus, slab, ops, n.value, ctypes.byref(n), ctypes.byref(z)))
    return list(ops), z.value


def plan(desc: Desc, steps: int, sm_count: int = 148) -> PlanInfo:
    """Kernel family / fused depth / geometry the library would use (host logic only)."""
    info = PlanInfo()
    check(lib().sb200_plan(ctypes.byref(desc), steps, sm_count, ctypes.byref(info)))
    return info
