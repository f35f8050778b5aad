"""GPU tests (-m gpu) of ONE C-ABI call driving several GPUs from one process
(sb200_init_multi + sb200_run_host with sb200_desc.ngpus): the counterpart of the reference's
single Stencil::run that spawns the whole core mesh (src/stencil/stencil.cpp:34-53).

Runs on a one-GPU box too: the device group then lists device 0 several times, so the slab
worker threads, the deep-halo sub-grids, the clipped schedules and the cross-slab ordering of
host reads / writes are all exercised; with >= 2 GPUs visible real devices are used."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

import oracle
from stencil_b200 import capi, sweep

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "stencil_b200", "host", "build", "bin", "stencil_main")


def _group(L, n):
    have = L.sb200_device_count()
    return [i % have for i in range(n)]


@pytest.fixture
def multi(gpu):
    """Re-binds the library to a device group for one test, then back to device 0."""
    L = gpu

    def bind(n):
        L.sb200_finalize()
        devs = (ctypes.c_int * n)(*_group(L, n))
        capi.check(L.sb200_init_multi(n, devs))
        assert L.sb200_group_size() == n
        return L
    yield bind
    L.sb200_finalize()
    capi.check(L.sb200_init(0))


def _run(n, r, shape, dtype, T, ngpus, seed=1, time_block=0, kernel=capi.KERNEL_AUTO):
    code = oracle.F64 if dtype == np.float64 else oracle.F32
    a = oracle.fill(n, r, code, oracle.INIT_RANDOM, seed)
    b = oracle.fill(n, r, code, oracle.INIT_REF)
    ga, gb = a.copy(), b.copy()
    a0, b0 = a.copy(), b.copy()
    want = b if oracle.sweep(n, r, shape, T, a, b, threads=oracle.max_threads()) else a
    got, t = sweep.run_host(n, ga, gb, T, radius=r, shape=shape, time_block=time_block, kernel=kernel, ngpus=ngpus)
    assert np.array_equal(want.view(np.uint8), got.view(np.uint8)), \
        f"max|d|={np.abs(want.astype(np.float64) - got.astype(np.float64)).max()} gpus={t.ngpus}"
    other, other0 = (ga, a0) if T & 1 else (gb, b0)
    assert np.array_equal(other.view(np.uint8), other0.view(np.uint8)), "the other buffer was touched"
    return t


@pytest.mark.parametrize("n,r,shape,dtype,T,G,chunks", [
    ((96, 80, 96), 1, capi.STAR, np.float64, 6, 2, 4),     # T even: results land in the array being uploaded
    ((96, 80, 96), 1, capi.STAR, np.float64, 7, 3, 3),
    ((70, 66, 128), 1, capi.BOX, np.float32, 4, 4, 3),
    ((300, 520), 1, capi.BOX, np.float32, 16, 4, 4),       # WARP2D k=4, y-slabs
    ((257, 400), 2, capi.STAR, np.float64, 6, 2, 5),       # warp2dr radius 2
    ((64, 48, 120), 2, capi.STAR, np.float64, 5, 2, 4),    # generic 3-D radius 2
])
def test_run_host_over_a_device_group(multi, monkeypatch, n, r, shape, dtype, T, G, chunks):
    multi(G)
    monkeypatch.setenv("SB200_PIPELINE_CHUNKS", str(chunks))
    d = capi.make_desc(n, r, shape, capi.F64 if dtype == np.float64 else capi.F32, T)
    assert capi.slab_plan(d, G).ngpus_used == G, "the case is meant to use every GPU of the group"
    t = _run(n, r, shape, dtype, T, ngpus=0)
    assert t.ngpus == G and t.pipelined == 1
    t1 = _run(n, r, shape, dtype, T, ngpus=1)              # same group, one GPU asked for
    assert t1.ngpus == 1


def test_default_size_slabs_are_pipelined_without_env(multi, monkeypatch):
    """Slabs of >= 64 MB take the pipelined schedule by themselves (no env override)."""
    monkeypatch.delenv("SB200_PIPELINE_CHUNKS", raising=False)
    multi(2)
    t = _run((256, 256, 512), 1, capi.STAR, np.float64, 20, ngpus=2)
    assert t.ngpus == 2 and t.pipelined == 1
    t = _run((256, 256, 512), 1, capi.STAR, np.float64, 21, ngpus=2)
    assert t.ngpus == 2


def test_thin_slabs_fall_back_to_fewer_gpus(multi, monkeypatch):
    monkeypatch.setenv("SB200_PIPELINE_CHUNKS", "3")
    multi(4)
    t = _run((64, 48, 64), 1, capi.STAR, np.float64, 12, ngpus=4)   # 16-plane slabs < 2 * 12: 2 GPUs
    assert t.ngpus == 2
    t = _run((64, 48, 40), 1, capi.STAR, np.float64, 12, ngpus=4)   # even 2 slabs are too thin
    assert t.ngpus == 1


def test_more_gpus_than_bound_is_an_error(multi):
    L = multi(2)
    a = oracle.fill((16, 16), 1, oracle.F32, oracle.INIT_REF)
    d = capi.make_desc((16, 16), iterations=2, ngpus=3)
    t = capi.Timing()
    assert L.sb200_run_host(ctypes.byref(d), a.ctypes.data, a.copy().ctypes.data, ctypes.byref(t)) == capi.ERR_INVALID
    assert b"device(s) are bound" in L.sb200_last_error()


def test_halo_precondition_is_checked(gpu):
    """`out` must carry the same halo as `in` (only `in` crosses PCIe): violated -> error."""
    a = oracle.fill((40, 30), 1, oracle.F32, oracle.INIT_RANDOM, 3)
    b = oracle.fill((40, 30), 1, oracle.F32, oracle.INIT_REF)
    b[:, 0] = 7.0                                            # a different x face
    d = capi.make_desc((40, 30), iterations=3)
    t = capi.Timing()
    assert gpu.sb200_run_host(ctypes.byref(d), a.ctypes.data, b.ctypes.data, ctypes.byref(t)) == capi.ERR_INVALID
    assert b"halo" in gpu.sb200_last_error()


def test_prepare_then_run(multi, monkeypatch):
    """sb200_prepare takes the one-time costs; the run after it gives the same bits."""
    monkeypatch.setenv("SB200_PIPELINE_CHUNKS", "3")
    L = multi(2)
    d = capi.make_desc((96, 80, 96), 1, capi.STAR, capi.F64, 7)
    capi.check(L.sb200_prepare(ctypes.byref(d)))
    _run((96, 80, 96), 1, capi.STAR, np.float64, 7, ngpus=0)
    for desc in (capi.make_desc((400, 400), iterations=40), capi.make_desc((5000,), dtype=capi.F64, iterations=20),
                 capi.make_desc((70, 60), 2, capi.BOX, iterations=3)):
        capi.check(L.sb200_prepare(ctypes.byref(desc)))


def test_reinit_on_another_binding_keeps_kernel_attributes(gpu):
    """finalize + init again (function attributes are per context: ADVICE r1)."""
    for _ in range(2):
        gpu.sb200_finalize()
        capi.check(gpu.sb200_init(0))
        _run((96, 80, 40), 1, capi.STAR, np.float64, 4, ngpus=0)     # TMA3D: > 48 KB dynamic shared memory
        _run((400, 400), 1, capi.STAR, np.float32, 40, ngpus=0)      # resident2d: cluster launch attribute


def test_cli_gpus_flag_self_check(gpu):
    """stencil_main --gpus N -m CUDA -c: the C++ host drives the device group with ONE call."""
    have = gpu.sb200_device_count()
    g = 2 if have >= 2 else 1
    env = dict(os.environ, SB200_PIPELINE_CHUNKS="4")
    r = subprocess.run([BIN, "-s", "1", "-i", "6", "-b", "1", "--shape", "160,120,96", "--dtype", "f64", "--init", "random",
                        "--gpus", str(g), "-m", "CUDA", "-c"], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "The results of method CUDA is correct." in r.stdout and "bit-exact vs naive loop: yes" in r.stdout
    assert re.search(rf", {g} GPU", r.stdout)
    r = subprocess.run([BIN, "-s", "64", "-i", "2", "-b", "1", "--gpus", str(have + 1), "-m", "CUDA"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 3, "asking for more GPUs than the box has is an error, not a silent single-GPU run"


@pytest.mark.parametrize("n,shape,dtype,T,hint", [
    ((300, 257), capi.BOX, np.float32, 9, 1), ((300, 257), capi.STAR, np.float64, 9, 7), ((300, 257), capi.BOX, np.float32, 9, 1000),
    ((200, 150), capi.STAR, np.float32, 9, 5),
    ((96, 80, 72), capi.STAR, np.float64, 5, 1), ((96, 80, 72), capi.STAR, np.float64, 5, 9), ((70, 66, 64), capi.BOX, np.float32, 4, 100),
])
def test_block_hint_tile_preference_keeps_the_bits(gpu, n, shape, dtype, T, hint):
    """-b as the CUDA tile preference (SB200_FLAG_BLOCK_TILE): any chunk size, same result."""
    code = oracle.F64 if dtype == np.float64 else oracle.F32
    a = oracle.fill(n, 1, code, oracle.INIT_RANDOM, 2)
    b = oracle.fill(n, 1, code, oracle.INIT_REF)
    ga, gb = a.copy(), b.copy()
    want = b if oracle.sweep(n, 1, shape, T, a, b, threads=oracle.max_threads()) else a
    d = capi.make_desc(n, 1, shape, capi.F64 if dtype == np.float64 else capi.F32, T, block_hint=hint,
                       flags=capi.FLAG_BLOCK_TILE)
    t = capi.Timing()
    capi.check(gpu.sb200_run_host(ctypes.byref(d), ga.ctypes.data, gb.ctypes.data, ctypes.byref(t)))
    got = gb if T & 1 else ga
    assert np.array_equal(want.view(np.uint8), got.view(np.uint8))


# ---- sb200_group_sweep: the device-resident multi-GPU run in one C-ABI call (no torch, no launcher) ----

def _single_gpu_digest(n, r, shape, dtype, T):
    d = capi.make_desc(n, r, shape, dtype, T)
    g = sweep.DeviceGrids(d)
    g.fill(capi.INIT_RANDOM, seed=6)
    g.sweep(T)
    return g.checksum()


@pytest.mark.parametrize("n,r,shape,dtype,T,G", [
    ((96, 80, 64), 1, capi.STAR, capi.F64, 7, 2),      # TMA3D, remainder block
    ((70, 66, 96), 1, capi.BOX, capi.F32, 6, 4),       # 27-point, four slabs
    ((300, 256), 1, capi.BOX, capi.F32, 13, 4),        # WARP2D rows
    ((257, 240), 2, capi.STAR, capi.F64, 5, 2),        # warp2dr radius 2
    ((200, 180), 2, capi.BOX, capi.F32, 6, 3),         # 25-point box, three slabs
    ((96, 80, 48), 2, capi.STAR, capi.F64, 5, 3),      # 13-point star (tma3dr), three slabs
    ((64, 48, 40), 2, capi.BOX, capi.F64, 3, 2),       # 125-point box (tma3db), two slabs
    ((72, 50, 36), 2, capi.BOX, capi.F32, 4, 3),       # the same in fp32, three slabs
    ((64, 48, 40), 3, capi.BOX, capi.F64, 3, 1),       # generic kernel: one GPU only
])
def test_group_sweep_digest_equals_one_gpu(multi, n, r, shape, dtype, T, G):
    L = multi(max(G, 2))
    want = _single_gpu_digest(n, r, shape, dtype, T)
    d = capi.make_desc(n, r, shape, dtype, T, ngpus=G)
    res = capi.GroupResult()
    capi.check(L.sb200_group_sweep(ctypes.byref(d), capi.INIT_RANDOM, 6, 2, ctypes.byref(res)))
    assert res.ngpus == G and res.exchange == (1 if G > 1 else 0)
    assert res.checksum == want, f"{res.checksum:#x} vs {want:#x}"
    assert res.sweep_ms > 0 and res.launches > 0
    # a second call (fresh buffers, counters from zero) gives the same digest
    capi.check(L.sb200_group_sweep(ctypes.byref(d), capi.INIT_RANDOM, 6, 0, ctypes.byref(res)))
    assert res.checksum == want


def test_group_sweep_rejects_what_it_cannot_link(multi):
    L = multi(2)
    res = capi.GroupResult()
    d = capi.make_desc((64, 48, 40), 3, capi.BOX, capi.F64, 3, ngpus=2)      # generic kernel has no fused exchange
    assert L.sb200_group_sweep(ctypes.byref(d), capi.INIT_REF, 0, 0, ctypes.byref(res)) == capi.ERR_UNSUPPORTED
    d = capi.make_desc((64, 48, 41), 1, capi.STAR, capi.F64, 3, ngpus=2)     # 41 planes over 2 GPUs
    assert L.sb200_group_sweep(ctypes.byref(d), capi.INIT_REF, 0, 0, ctypes.byref(res)) == capi.ERR_INVALID
    d = capi.make_desc((64, 48, 6), 1, capi.STAR, capi.F64, 3, ngpus=2)      # 3-plane slabs < 2 ghost zones
    assert L.sb200_group_sweep(ctypes.byref(d), capi.INIT_REF, 0, 0, ctypes.byref(res)) == capi.ERR_INVALID
    d = capi.make_desc((64, 48, 40), 1, capi.STAR, capi.F64, 3, ngpus=3)     # more than the group holds
    assert L.sb200_group_sweep(ctypes.byref(d), capi.INIT_REF, 0, 0, ctypes.byref(res)) == capi.ERR_INVALID


def test_cli_resident_mode_self_check(gpu):
    """stencil_main --resident [--gpus N] -c: the grid lives on the GPUs; -c compares the device digest
    with the digest of the host's naive loop."""
    have = gpu.sb200_device_count()
    g = 2 if have >= 2 else 1
    for extra in (["--shape", "160,120,96", "--dtype", "f64", "-i", "7"], ["-s", "512", "--stencil", "box", "-i", "21"]):
        r = subprocess.run([BIN, "-s", "1", "-b", "1", "--init", "random", "--resident", "--gpus", str(g), "-m", "CUDA", "-c"] + extra,
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout + r.stderr
        assert "The results of method CUDA is correct." in r.stdout and "bit-exact vs naive loop: yes" in r.stdout
        assert "device-resident" in r.stdout and re.search(rf", {g} GPU", r.stdout)
