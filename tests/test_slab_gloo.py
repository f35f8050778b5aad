"""world_size-2/3 tests of the slab partition + halo-exchange logic on CPU (gloo).

The CUDA kernels cannot run here, so a CPU engine defined in THIS test file (it uses the
oracle for the arithmetic — allowed in tests/ only) is injected into
stencil_b200.slab.SlabSweep; ranges, ghost depths, exchange indices, buffer parity and the
remainder block are the product's own code.  The stitched result must equal the
single-grid oracle bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from stencil_b200 import capi, slab  # noqa: E402


class _NullCtx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class OracleEngine:
    """CPU stand-in for CudaEngine: dense padded numpy slabs, oracle arithmetic."""

    def __init__(self, desc, ghost):
        self.desc, self.ghost = desc, ghost
        self.nd = desc.ndim
        self.r = int(desc.radius)
        self.n = [int(desc.n[i]) for i in range(self.nd)]
        self.code = oracle.F64 if desc.dtype == capi.F64 else oracle.F32
        shp = list(oracle.padded_shape(self.n, self.r))
        shp[0] += 2 * ghost
        self.buf = [torch.zeros(shp, dtype=torch.float64 if self.code else torch.float32) for _ in range(2)]
        self.off = self.r + ghost

    def planes(self, which, z0, z1):
        return self.buf[which][z0 + self.off:z1 + self.off]

    def fill(self, which, init, seed, z_offset, z_total):
        # a slab with ghost planes == the slab of a taller grid starting `ghost` planes lower
        n = list(self.n)
        n[-1] += 2 * self.ghost
        tot = z_total if z_total else self.n[-1]
        full = oracle.fill(n, self.r, self.code, init, seed, z_offset=0, z_total=0)
        # generate from the global grid to get true neighbour data in the ghosts
        gn = list(self.n)
        gn[-1] = tot
        glob = oracle.fill(gn, self.r, self.code, init, seed)
        lo = z_offset - self.ghost
        for p in range(full.shape[0]):
            gp = lo + p
            if 0 <= gp < glob.shape[0]:
                full[p] = glob[gp]
            else:
                full[p] = 0
        self.buf[which].copy_(torch.from_numpy(full))

    def sweep_range(self, src, dst, steps, zb, ze, lo_b, hi_b, stream=None):
        if ze <= zb:
            return 0
        lo = 0 if lo_b else -self.ghost
        hi = self.n[-1] if hi_b else self.n[-1] + self.ghost
        sub_n = list(self.n)
        sub_n[-1] = hi - lo
        a = self.planes(src, lo - self.r, hi + self.r).numpy().copy()
        b = a.copy()
        shape = oracle.STAR if self.desc.shape == capi.STAR else oracle.BOX
        res = b if oracle.sweep(sub_n, self.r, shape, steps, a, b) else a
        out = self.planes(dst, zb, ze).numpy()
        r = self.r
        inner = (slice(None),) + tuple(slice(r, -r) for _ in range(self.nd - 1))
        out[inner] = res[zb - lo + r:ze - lo + r][inner]
        return 1

    # exchange="p2p" stand-in: one "launch" sweeps the whole slab, then the face planes land
    # in the neighbours' ghost planes (the CUDA engine does this with peer stores + flags)
    def connect(self, rank, world, dist_):
        self.rank, self.world, self.dist = rank, world, dist_

    def sync(self):
        pass

    def link_reset(self):
        self.resets = getattr(self, "resets", 0) + 1

    def link_timed_out(self):
        return False

    def sweep_linked(self, src, dst, steps, lo_b, hi_b):
        n, g = self.n[-1], self.ghost
        self.sweep_range(src, dst, steps, 0, n, lo_b, hi_b)
        ops = []
        if not lo_b:
            ops.append(self.dist.P2POp(self.dist.isend, self.planes(dst, 0, g).contiguous(), self.rank - 1))
            ops.append(self.dist.P2POp(self.dist.irecv, self.planes(dst, -g, 0), self.rank - 1))
        if not hi_b:
            ops.append(self.dist.P2POp(self.dist.isend, self.planes(dst, n - g, n).contiguous(), self.rank + 1))
            ops.append(self.dist.P2POp(self.dist.irecv, self.planes(dst, n, n + g), self.rank + 1))
        for w in self.dist.batch_isend_irecv(ops):
            w.wait()
        return 1

    def side_stream(self):
        return None

    def fork(self, s):
        pass

    def join(self, s):
        pass

    def stream_ctx(self, s):
        return _NullCtx()


def _worker(rank, world, port, n, shape, dtype, T, k, out, exchange="nccl", radius=1):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ln = list(n)
        ln[-1] //= world
        d = capi.make_desc(ln, radius, shape, dtype, T, time_block=k)
        sw = slab.SlabSweep(d, rank, world, engine_factory=OracleEngine, dist=dist, exchange=exchange)
        assert sw.exchange == exchange
        sw.fill(capi.INIT_RANDOM, seed=4)
        nl = sw.sweep(T)
        res = sw.eng.planes(sw.cur, 0, ln[-1]).clone()
        gathered = [torch.zeros_like(res) for _ in range(world)]
        dist.all_gather(gathered, res)
        if rank == 0:
            np.save(out, torch.cat(gathered).numpy())
            assert nl > 0
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,shape,dtype,T,k", [
    (2, (20, 12, 16), capi.STAR, capi.F64, 5, 2),     # 3-D, remainder block
    (2, (24, 40), capi.BOX, capi.F32, 9, 4),          # 2-D slabs along y
    (3, (10, 9, 18), capi.STAR, capi.F32, 4, 3),      # three ranks, middle rank has 2 faces
])
def test_slab_exchange_matches_single_grid(tmp_path, world, n, shape, dtype, T, k):
    out = str(tmp_path / "res.npy")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, n, shape, dtype, T, k, out), nprocs=world, join=True)
    got = np.load(out)
    code = oracle.F64 if dtype == capi.F64 else oracle.F32
    want = oracle.run(n, 1, oracle.STAR if shape == capi.STAR else oracle.BOX, T, code,
                      oracle.INIT_RANDOM, 4)
    assert np.array_equal(got.view(np.uint8), want[1:-1].view(np.uint8))


@pytest.mark.parametrize("world,n,T,k", [(2, (20, 12, 16), 5, 2), (3, (10, 9, 18), 7, 3),
                                         (2, (24, 40), 9, 4)])    # 2-D: the linked flow over row faces
def test_slab_linked_flow_matches_single_grid(tmp_path, world, n, T, k):
    """exchange="p2p": one linked launch per fused block, remainder block included."""
    out = str(tmp_path / "res.npy")
    port = 31500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, n, capi.STAR, capi.F64, T, k, out, "p2p"), nprocs=world, join=True)
    got = np.load(out)
    want = oracle.run(n, 1, oracle.STAR, T, oracle.F64, oracle.INIT_RANDOM, 4)
    assert np.array_equal(got.view(np.uint8), want[1:-1].view(np.uint8))


@pytest.mark.parametrize("world,n,r,shape,dtype,T", [
    (2, (14, 12, 24), 2, capi.BOX, capi.F32, 4),      # 125-point box (tma3db): faces are r planes, one step per launch
    (2, (16, 10, 20), 2, capi.STAR, capi.F64, 5),     # 13-point star (tma3dr)
    (3, (20, 36), 3, capi.BOX, capi.F64, 3),          # 49-point box (warp2dr), three ranks
])
def test_slab_linked_flow_radius_kernels(tmp_path, world, n, r, shape, dtype, T):
    """The linked flow at radius > 1 (ghost depth r * k, k = 1 for these families) against the
    single-grid oracle."""
    out = str(tmp_path / "res.npy")
    port = 33500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, n, shape, dtype, T, 1, out, "p2p", r), nprocs=world, join=True)
    got = np.load(out)
    code = oracle.F64 if dtype == capi.F64 else oracle.F32
    want = oracle.run(n, r, oracle.STAR if shape == capi.STAR else oracle.BOX, T, code, oracle.INIT_RANDOM, 4)
    assert np.array_equal(got.view(np.uint8), want[r:-r].view(np.uint8))


def test_exchange_mode_selection():
    d3 = capi.make_desc((8, 8, 8), time_block=2)
    d2 = capi.make_desc((8, 16), shape=capi.BOX)
    d2r = capi.make_desc((8, 32), radius=2)
    d3r = capi.make_desc((8, 8, 16), radius=2)
    d2g = capi.make_desc((8, 48), radius=5, shape=capi.BOX)
    assert slab.SlabSweep(d3, 0, 1, engine_factory=OracleEngine).exchange == "none"
    # the fused exchange exists for the WARP2D family (rows) and the TMA3D family (planes)
    assert slab.SlabSweep(d2, 0, 2, engine_factory=OracleEngine, dist=dist, exchange="p2p").exchange == "p2p"
    assert slab.SlabSweep(d2, 0, 2, engine_factory=OracleEngine, dist=dist).exchange == "p2p"
    assert slab.SlabSweep(d2r, 0, 2, engine_factory=OracleEngine, dist=dist).exchange == "p2p"
    assert slab.SlabSweep(d3r, 0, 2, engine_factory=OracleEngine, dist=dist).exchange == "p2p"    # 13-point: tma3dr
    assert slab.SlabSweep(d2g, 0, 2, engine_factory=OracleEngine, dist=dist).exchange == "nccl"   # generic kernel
    with pytest.raises(ValueError):
        slab.SlabSweep(d2g, 0, 2, engine_factory=OracleEngine, dist=dist, exchange="p2p")
    with pytest.raises(ValueError):
        slab.SlabSweep(d3, 0, 2, engine_factory=OracleEngine, dist=dist, exchange="bogus")


def test_resolve_kernel_defaults():
    assert slab.resolve_kernel(capi.make_desc((8, 8, 8))) == (capi.KERNEL_TMA3D, 2)
    assert slab.resolve_kernel(capi.make_desc((8, 8), shape=capi.BOX)) == (capi.KERNEL_WARP2D, 4)
    assert slab.resolve_kernel(capi.make_desc((8, 8, 8), shape=capi.BOX)) == (capi.KERNEL_TMA3D, 2)
    assert slab.resolve_kernel(capi.make_desc((8, 8, 8), radius=2)) == (capi.KERNEL_TMA3D, 1)
    assert slab.resolve_kernel(capi.make_desc((8, 8, 8), radius=2, shape=capi.BOX)) == (capi.KERNEL_TMA3D, 1)
    assert slab.resolve_kernel(capi.make_desc((8, 8, 8), radius=3, shape=capi.BOX)) == (capi.KERNEL_GENERIC, 1)
    assert slab.resolve_kernel(capi.make_desc((8, 8, 8), radius=3, dtype=capi.F64)) == (capi.KERNEL_TMA3D, 1)
    assert slab.resolve_kernel(capi.make_desc((8, 8, 8), radius=3)) == (capi.KERNEL_GENERIC, 1)
    assert slab.resolve_kernel(capi.make_desc((8, 8, 8), radius=5)) == (capi.KERNEL_GENERIC, 1)
    assert slab.resolve_kernel(capi.make_desc((8, 8), time_block=99)) == (capi.KERNEL_WARP2D, 6)
    assert slab.resolve_kernel(capi.make_desc((8, 8), radius=2)) == (capi.KERNEL_WARP2D, 4)
    assert slab.resolve_kernel(capi.make_desc((8, 8), radius=3, dtype=capi.F64, time_block=5)) == (capi.KERNEL_WARP2D, 2)
    assert slab.resolve_kernel(capi.make_desc((8, 8), radius=2, shape=capi.BOX)) == (capi.KERNEL_WARP2D, 2)
    assert slab.resolve_kernel(capi.make_desc((8, 8), radius=3, shape=capi.BOX)) == (capi.KERNEL_WARP2D, 1)
    assert slab.resolve_kernel(capi.make_desc((8, 8), radius=5, shape=capi.BOX)) == (capi.KERNEL_GENERIC, 1)
    assert slab.resolve_kernel(capi.make_desc((8, 8), radius=4, shape=capi.BOX)) == (capi.KERNEL_GENERIC, 1)
    assert slab.resolve_kernel(capi.make_desc((8, 8), radius=4, shape=capi.BOX, dtype=capi.F64)) == (capi.KERNEL_WARP2D, 1)
    assert slab.resolve_kernel(capi.make_desc((800,), radius=3)) == (capi.KERNEL_LINE1D, 10)
    assert slab.resolve_kernel(capi.make_desc((800,), radius=5)) == (capi.KERNEL_GENERIC, 1)


def test_thin_slab_rejected():
    with pytest.raises(ValueError):
        slab.SlabSweep(capi.make_desc((8, 8, 3), time_block=2), 0, 2, engine_factory=OracleEngine, dist=dist)
