"""Slab decomposition along the outermost axis, one process per GPU.

The reference decomposes its grid over an 8x8 core mesh and exchanges halos every step
(block_subview, boundary_matrix.hpp:190-218; RMA puts, stencil_rma.cpp:198-255).  Here
rank g owns a contiguous slab of planes plus `ghost = radius * k` ghost planes per
internal face, exchanged once per fused block of k time steps:

  exchange="p2p" (default for the radius-1 families, TMA3D in 3-D and WARP2D in 2-D): ONE launch
    per block; its face CTAs (warps in 2-D) run first and store their planes straight into the neighbours' ghost planes over
    NVLink (CUDA IPC peer mappings), signalling arrival counters the neighbours' next launch
    acquires — no NCCL call, no extra launch, no host sync on the data path
    (csrc/slablink.cuh: SlabLinkDev).
  exchange="nccl" (radius > 1, generic kernel, and the measured baseline):
    faces first (side stream)  ->  send/recv face planes (NCCL, NVLink)   \
    interior planes (main stream, concurrently)                             -> join

Per-point arithmetic is the single-GPU kernel's, so results are bitwise independent of
the number of ranks.  The compute engine is pluggable so the partition / exchange logic
is unit-tested on CPU with gloo (tests/test_slab_gloo.py injects a CPU engine); the
product engine is CudaEngine below and nothing else.
"""
from __future__ import annotations

import ctypes

from . import capi

_KERNEL_NAMES = {capi.KERNEL_GENERIC: "generic", capi.KERNEL_WARP2D: "warp2d",
                 capi.KERNEL_TMA3D: "tma3d", capi.KERNEL_LINE1D: "line1d"}


def resolve_kernel(desc: capi.Desc):
    """(family, fused depth) the library will use for this descriptor (mirrors choose()
    in csrc/capi.cu; max depth is asked from the library)."""
    w2d = desc.ndim == 2 and 1 <= desc.radius <= 4 and not (
        desc.shape == capi.BOX and desc.radius == 4 and desc.dtype == capi.F32)
    t3 = desc.ndim == 3 and (desc.radius <= 2 or (desc.shape == capi.STAR and 3 <= desc.radius <= 4
                                                  and desc.dtype == capi.F64))
    l1 = desc.ndim == 1 and desc.radius <= 4
    sep9 = desc.shape == capi.BOX and bool(desc.flags & capi.FLAG_REASSOCIATE)     # separable 9-point box
    kern = desc.kernel
    if kern == capi.KERNEL_AUTO:
        kern = (capi.KERNEL_WARP2D if w2d else capi.KERNEL_TMA3D if t3
                else capi.KERNEL_LINE1D if l1 else capi.KERNEL_GENERIC)
    d = capi.Desc.from_buffer_copy(desc)
    d.kernel = kern
    max_k = int(capi.lib().sb200_max_time_block(ctypes.byref(d)))
    if max_k < 1:
        raise capi.StencilError(max_k, capi.lib().sb200_last_error().decode())
    k = desc.time_block or {capi.KERNEL_GENERIC: 1, capi.KERNEL_TMA3D: 2, capi.KERNEL_LINE1D: min(16, 32 // max(1, desc.radius)),
                            capi.KERNEL_WARP2D: (6 if sep9 else 4) if desc.radius == 1 else max_k}[kern]
    return kern, min(k, max_k)


class CudaEngine:
    """Device buffers + kernels through the C ABI (the product path)."""

    def __init__(self, desc: capi.Desc, ghost: int, device_index: int, exportable: bool = False):
        import torch
        self.torch = torch
        self.desc, self.ghost = desc, ghost
        self.device = torch.device("cuda", device_index)
        self.lay = capi.layout_for(desc, ghost)
        self.np_dtype = torch.float64 if desc.dtype == capi.F64 else torch.float32
        es = 8 if desc.dtype == capi.F64 else 4
        self.exportable = exportable
        self.link = None
        if exportable:
            # cudaMalloc'ed by the library so the buffers can be IPC-exported to the neighbours
            self.raw = [_LibBuffer(self.lay.bytes) for _ in range(2)]
            self.flags = _LibBuffer(4 * capi.LINK_FLAG_WORDS)
            self.side = torch.cuda.Stream(self.device, priority=-1)
            return
        self.raw = [torch.zeros(self.lay.bytes, dtype=torch.uint8, device=self.device) for _ in range(2)]
        if desc.ndim == 3:
            shp = (self.lay.planes, self.lay.pitch_y, self.lay.pitch_x)
            self.off = int(self.lay.oz)
        else:
            shp = (self.lay.pitch_y, self.lay.pitch_x)
            self.off = int(self.lay.oy)
        self.view = [r.view(self.np_dtype).view(shp) for r in self.raw]
        assert es * self.view[0].numel() == self.lay.bytes
        self.side = torch.cuda.Stream(self.device, priority=-1)

    def planes(self, which: int, z0: int, z1: int):
        return self.view[which][z0 + self.off:z1 + self.off]

    def _st(self, stream=None):
        s = stream if stream is not None else self.torch.cuda.current_stream(self.device)
        return ctypes.c_void_p(s.cuda_stream)

    def fill(self, which, init, seed, z_offset, z_total):
        capi.check(capi.lib().sb200_fill_device(ctypes.byref(self.desc), self.ghost,
                                                self.raw[which].data_ptr(), init, seed, z_offset,
                                                z_total, self._st()))

    def sweep_range(self, src, dst, steps, zb, ze, lo_b, hi_b, stream=None) -> int:
        nl = ctypes.c_uint32(0)
        capi.check(capi.lib().sb200_sweep_device_range(
            ctypes.byref(self.desc), self.ghost, lo_b, hi_b, self.raw[src].data_ptr(),
            self.raw[dst].data_ptr(), steps, zb, ze, self._st(stream), ctypes.byref(nl)))
        return nl.value

    def sweep_all(self, src, dst, steps) -> tuple[int, int]:
        in_b, nl = ctypes.c_int(0), ctypes.c_uint32(0)
        capi.check(capi.lib().sb200_sweep_device(
            ctypes.byref(self.desc), 0, 1, 1, self.raw[src].data_ptr(), self.raw[dst].data_ptr(),
            steps, self._st(), ctypes.byref(in_b), ctypes.byref(nl)))
        return (dst if in_b.value else src), nl.value

    def checksum(self, which, z_offset, z_total) -> int:
        out = ctypes.c_uint64(0)
        capi.check(capi.lib().sb200_checksum_device(
            ctypes.byref(self.desc), self.ghost, self.raw[which].data_ptr(), z_offset, z_total,
            self._st(), ctypes.byref(out)))
        return out.value

    # -- peer link (exchange="p2p") ---------------------------------------------------
    def connect(self, rank: int, world: int, dist) -> None:
        """Exchange CUDA IPC handles of both grids + the flag block with all ranks and map
        the two neighbours' memory."""
        torch, L = self.torch, capi.lib()
        hb = capi.IPC_HANDLE_BYTES
        mine, export_err = bytearray(), None
        try:
            for buf in (self.raw[0], self.raw[1], self.flags):
                h = ctypes.create_string_buffer(hb)
                capi.check(L.sb200_ipc_export(buf.data_ptr(), h))
                mine += h.raw
        except capi.StencilError as e:
            export_err, mine = e, bytearray(3 * hb)
        # every rank takes part in the collective even if its export failed
        t = torch.tensor(list(mine), dtype=torch.uint8, device=self.device)
        allh = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allh, t)
        if export_err is not None:
            raise export_err

        def open_rank(r):
            if r < 0 or r >= world:
                return [None, None, None]
            raw = bytes(allh[r].cpu().tolist())
            ptrs = []
            for i in range(3):
                p = ctypes.c_void_p()
                capi.check(L.sb200_ipc_open(raw[i * hb:(i + 1) * hb], ctypes.byref(p)))
                ptrs.append(p.value)
            return ptrs
        self.peer_lo, self.peer_hi = open_rank(rank - 1), open_rank(rank + 1)
        self.arrivals = 0
        self.link = capi.SlabLink()

    def sweep_linked(self, src, dst, steps, lo_b, hi_b) -> int:
        lk = self.link
        lk.peer_dst_lo, lk.peer_dst_hi = self.peer_lo[dst], self.peer_hi[dst]
        lk.flags_local = self.flags.data_ptr()
        lk.flags_lo, lk.flags_hi = self.peer_lo[2], self.peer_hi[2]
        lk.wait_arrivals = self.arrivals & 0xFFFFFFFF
        arr = ctypes.c_uint32(0)
        capi.check(capi.lib().sb200_sweep_slab_linked(
            ctypes.byref(self.desc), self.ghost, lo_b, hi_b, self.raw[src].data_ptr(),
            self.raw[dst].data_ptr(), steps, ctypes.byref(lk), self._st(), ctypes.byref(arr)))
        self.arrivals += arr.value
        return 1

    def link_reset(self) -> None:
        """Zero the arrival counters and the time-out flag (all ranks, at a quiescent point)."""
        capi.check(capi.lib().sb200_link_reset(self.flags.data_ptr()))
        self.arrivals = 0

    def link_timed_out(self) -> bool:
        """True if any face CTA gave up waiting for a neighbour (results are then invalid)."""
        words = (ctypes.c_uint32 * capi.LINK_FLAG_WORDS)()
        capi.check(capi.lib().sb200_link_status(self.flags.data_ptr(), words))
        return words[2] != 0

    def close_peers(self):
        if self.exportable and self.link is not None:
            for ptrs in (self.peer_lo, self.peer_hi):
                for p in ptrs:
                    if p:
                        capi.check(capi.lib().sb200_ipc_close(p))
            self.peer_lo = self.peer_hi = [None, None, None]
            self.link = None

    def free(self):
        self.raw = []
        self.flags = None

    def sync(self):
        self.torch.cuda.synchronize(self.device)

    # stream plumbing used by SlabSweep
    def side_stream(self):
        return self.side

    def fork(self, stream):
        stream.wait_stream(self.torch.cuda.current_stream(self.device))

    def join(self, stream):
        self.torch.cuda.current_stream(self.device).wait_stream(stream)

    def stream_ctx(self, stream):
        return self.torch.cuda.stream(stream)


class _LibBuffer:
    """Zero-filled device memory owned by the library (cudaMalloc: IPC-exportable)."""

    def __init__(self, nbytes: int):
        p = ctypes.c_void_p()
        capi.check(capi.lib().sb200_device_alloc(ctypes.byref(p), nbytes))
        self.ptr, self.nbytes = p.value, nbytes

    def data_ptr(self) -> int:
        return self.ptr

    def __del__(self):
        try:
            if self.ptr:
                capi.lib().sb200_device_free(self.ptr)
                self.ptr = None
        except Exception:  # noqa: BLE001  (interpreter shutdown)
            pass


class SlabSweep:
    """T double-buffered steps of `desc` (the per-rank slab) across `world` ranks."""

    def __init__(self, desc: capi.Desc, rank: int = 0, world: int = 1, engine_factory=None,
                 dist=None, exchange: str = "auto"):
        self.desc, self.rank, self.world = desc, rank, world
        self.kernel, self.time_block = resolve_kernel(desc)
        self.kernel_name = _KERNEL_NAMES[self.kernel]
        self.r = int(desc.radius)
        self.n_outer = int(desc.n[desc.ndim - 1])
        if world > 1 and desc.ndim == 1:
            raise ValueError("1-D grids are not slab-decomposed")
        self.ghost = self.r * self.time_block if world > 1 else 0
        if world > 1 and self.n_outer < 2 * self.ghost:
            raise ValueError(f"slab of {self.n_outer} planes is thinner than two ghost zones "
                             f"({self.ghost} each); use fewer ranks or a smaller time block")
        auto = exchange == "auto"
        if auto:
            exchange = "p2p" if (world > 1 and self._linkable()) else "nccl"
        if exchange not in ("p2p", "nccl"):
            raise ValueError(f"exchange must be 'auto', 'p2p' or 'nccl' (got {exchange!r})")
        if exchange == "p2p" and not self._linkable():
            raise ValueError("exchange='p2p' needs the WARP2D family (2-D) or the TMA3D family (3-D)")
        self.exchange = exchange if world > 1 else "none"
        p2p = self.exchange == "p2p"
        factory = engine_factory or (lambda d, g: CudaEngine(d, g, _current_device(), exportable=p2p))
        self.eng = factory(desc, self.ghost)
        self.cur = 0
        self.lo_b = 1 if rank == 0 else 0
        self.hi_b = 1 if rank == world - 1 else 0
        self.kernels_per_launch_unit = 1 if (world == 1 or p2p) else (1 + (not self.lo_b) + (not self.hi_b))
        if world > 1:
            if dist is None:
                import torch.distributed as dist  # noqa: PLC0415
            self.dist = dist
            if p2p:
                # Peer mapping needs CUDA IPC between the ranks' processes.  If ANY rank cannot
                # map its neighbours (containers without shared IPC namespaces, GPUs without
                # peer access) every rank learns it here; exchange="auto" then drops to the
                # NCCL exchange on the GPU, an explicit exchange="p2p" fails loudly.
                err = None
                try:
                    self.eng.connect(rank, world, dist)
                except Exception as e:  # noqa: BLE001
                    err = e
                if not self._all_ok(err is None):
                    if not auto:
                        raise RuntimeError(f"exchange='p2p' unavailable on rank {rank}: {err or 'a peer rank failed'}")
                    if rank == 0:
                        print(f"stencil_b200: peer mapping unavailable ({err or 'on a peer rank'}); "
                              "using the NCCL halo exchange", flush=True)
                    self.exchange = "nccl"
                    self.kernels_per_launch_unit = 1 + (not self.lo_b) + (not self.hi_b)
                    self.eng = (engine_factory or (lambda d, g: CudaEngine(d, g, _current_device())))(desc, self.ghost)

    def _linkable(self) -> bool:
        """Kernel families with the fused peer-store exchange (sb200_sweep_slab_linked)."""
        return self.kernel in (capi.KERNEL_WARP2D, capi.KERNEL_TMA3D)

    def _all_ok(self, ok: bool) -> bool:
        """Logical AND of `ok` over all ranks (engines that are not CUDA-backed skip it)."""
        torch = getattr(self.eng, "torch", None)
        if torch is None:
            return ok
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=self.eng.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return bool(int(t.item()))

    # -- setup ----------------------------------------------------------------------
    def _quiesce(self):
        """p2p: neighbours write into this rank's ghost planes from inside their kernels, so
        host-side buffer writes (fill) must be fenced by device sync + barrier on both sides."""
        if self.exchange == "p2p":
            self.eng.sync()
            self.dist.barrier()

    def fill(self, init=capi.INIT_REF, seed=0):
        off, tot = self.rank * self.n_outer, self.n_outer * self.world
        if self.world == 1:
            off = tot = 0
        self._quiesce()
        if self.exchange == "p2p":
            # a time-out of an earlier run must not poison this one: counters restart at 0 on
            # every rank between the two barriers of the quiesce
            self.check_link()
            self.eng.link_reset()
            self.dist.barrier()
        self.eng.fill(0, init, seed, off, tot)
        self.eng.fill(1, capi.INIT_REF, 0, off, tot)
        self.cur = 0
        self._quiesce()

    def check_link(self) -> None:
        """p2p: raise if a face CTA ever gave up waiting for its neighbour (the ghost planes it
        then read were stale, so every result since is invalid).  Synchronises the device.
        Called at every host sync point of this class; callers that time a region call it
        right after the region."""
        if self.exchange == "p2p" and self.eng is not None and self.eng.link_timed_out():
            raise RuntimeError(f"rank {self.rank}: a fused halo exchange timed out waiting for a neighbour; "
                               "results are invalid")

    def close(self):
        """p2p: unmap the neighbours' memory on every rank BEFORE any rank frees its grids."""
        if self.exchange == "p2p" and self.eng is not None:
            self._quiesce()
            self.check_link()
            self.eng.close_peers()
            self.dist.barrier()
            self.eng.free()
        self.eng = None

    def checksum(self) -> int:
        off, tot = self.rank * self.n_outer, self.n_outer * self.world
        if self.world == 1:
            off = tot = 0
        self.check_link()
        return self.eng.checksum(self.cur, off, tot)

    def download_initial_into(self, ha, hb, seed):
        """Fill two host arrays (dense padded layout of ONE rank-local grid) with the same
        synthetic initial condition, generated on the device (bench e2e leg)."""
        import numpy as np  # noqa: PLC0415
        from . import sweep  # noqa: PLC0415
        g = sweep.DeviceGrids(self.desc, 0)
        g.fill(capi.INIT_RANDOM, seed)
        for which, h in ((0, ha), (1, hb)):
            capi.check(capi.lib().sb200_download(ctypes.byref(self.desc), 0,
                                                 (g.a, g.b)[which].data_ptr(), h.ctypes.data,
                                                 g._stream()))
        g.torch.cuda.synchronize()
        assert isinstance(ha, np.ndarray)

    # -- the sweep --------------------------------------------------------------------
    def sweep(self, steps: int) -> int:
        """Advance `steps` time steps; returns the number of kernels launched."""
        if self.world == 1:
            self.cur, nl = self.eng.sweep_all(self.cur, self.cur ^ 1, steps)
            return nl
        launches, done = 0, 0
        while done < steps:
            k = min(self.time_block, steps - done)
            launches += self._block(k)
            done += k
        return launches

    def _block(self, k: int) -> int:
        if self.exchange == "p2p":
            dst = self.cur ^ 1
            nl = self.eng.sweep_linked(self.cur, dst, k, self.lo_b, self.hi_b)
            self.cur = dst
            return nl
        # faces are always `ghost` deep (also for a shorter remainder block) so that every
        # ghost plane is refreshed to the new time level by each exchange
        eng, n, g = self.eng, self.n_outer, self.ghost
        src, dst = self.cur, self.cur ^ 1
        side = eng.side_stream()
        nl = 0
        # 1. faces on the side stream (they gate the exchange)
        eng.fork(side)
        with eng.stream_ctx(side):
            if not self.lo_b:
                nl += eng.sweep_range(src, dst, k, 0, g, self.lo_b, self.hi_b, side)
            if not self.hi_b:
                nl += eng.sweep_range(src, dst, k, n - g, n, self.lo_b, self.hi_b, side)
        # 2. interior on the main stream, concurrently
        zb = 0 if self.lo_b else g
        ze = n if self.hi_b else n - g
        nl += eng.sweep_range(src, dst, k, zb, ze, self.lo_b, self.hi_b)
        # 3. exchange the freshly computed face planes into the neighbours' ghost planes
        with eng.stream_ctx(side):
            ops = []
            P2P = self.dist.P2POp
            if not self.lo_b:
                ops.append(P2P(self.dist.isend, eng.planes(dst, 0, g), self.rank - 1))
                ops.append(P2P(self.dist.irecv, eng.planes(dst, -g, 0), self.rank - 1))
            if not self.hi_b:
                ops.append(P2P(self.dist.isend, eng.planes(dst, n - g, n), self.rank + 1))
                ops.append(P2P(self.dist.irecv, eng.planes(dst, n, n + g), self.rank + 1))
            for w in self.dist.batch_isend_irecv(ops):
                w.wait()
        # 4. join
        eng.join(side)
        self.cur = dst
        return nl


def _current_device() -> int:
    import torch  # noqa: PLC0415
    return torch.cuda.current_device()
