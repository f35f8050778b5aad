"""GPU parity tests (-m gpu): every CUDA kernel family vs the CPU oracle, bit-exact, all
through the C ABI (ctypes -> libstencil_b200.so).  Edge cases follow what the reference
exposes: any radius (-r), tiny and ragged grids (block_subview clipping,
boundary_matrix.hpp:190-218), T parity of the result buffer (stencil.cpp:88-92)."""
import ctypes
import os

import numpy as np
import pytest

import oracle
from stencil_b200 import capi, sweep

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_2d_f32.npz"))


def _oracle_pair(n, r, dtype, seed):
    code = oracle.F64 if dtype == np.float64 else oracle.F32
    init = oracle.INIT_RANDOM if seed else oracle.INIT_REF
    return oracle.fill(n, r, code, init, seed), oracle.fill(n, r, code, oracle.INIT_REF)


def _check(n, r, shape, dtype, T, seed=1, time_block=0, kernel=capi.KERNEL_AUTO, flags=0):
    a, b = _oracle_pair(n, r, dtype, seed)
    ga, gb = a.copy(), b.copy()
    want = b if oracle.sweep(n, r, shape, T, a, b, threads=oracle.max_threads()) else a
    got, t = sweep.run_host(n, ga, gb, T, radius=r, shape=shape, time_block=time_block, kernel=kernel, flags=flags)
    assert t.result_in_out == (T & 1)
    assert got is (gb if T & 1 else ga)
    assert np.array_equal(want.view(np.uint8), got.view(np.uint8)), \
        f"max|d|={np.abs(want.astype(np.float64) - got.astype(np.float64)).max()} kernel={t.kernel} k={t.time_block}"
    # the other buffer keeps its halo
    other, other0 = (ga, a) if T & 1 else (gb, b)
    m = np.ones(other.shape, bool)
    m[tuple(slice(r, s - r) for s in other.shape)] = False
    assert np.array_equal(other[m], other0[m])
    return t


@pytest.mark.parametrize("key", sorted(GOLD.files))
def test_reference_golden_2d_f32(gpu, key):
    """CUDA path vs grids produced by the genuine reference loop (bit-exact)."""
    N, r, T = (int(p[1:]) for p in key.split("_"))
    a, b = _oracle_pair((N, N), r, np.float32, 0)
    got, _ = sweep.run_host((N, N), a, b, T, radius=r)
    assert np.array_equal(oracle.interior(got, r).view(np.uint32), GOLD[key].view(np.uint32))


GOLD64 = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_2d_f64.npz"))


@pytest.mark.parametrize("key", sorted(GOLD64.files))
def test_reference_golden_2d_f64(gpu, key):
    """CUDA path vs the reference's loop compiled at double precision (bit-exact)."""
    N, r, T = (int(p[1:]) for p in key.split("_"))
    a, b = _oracle_pair((N, N), r, np.float64, 0)
    got, _ = sweep.run_host((N, N), a, b, T, radius=r)
    assert np.array_equal(oracle.interior(got, r).view(np.uint64), GOLD64[key].view(np.uint64))


GOLD_EMU = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_emu_2d_f32.npz"))


@pytest.mark.parametrize("key", sorted(GOLD_EMU.files))
def test_reference_accelerator_kernels_2d_f32(gpu, key):
    """CUDA path vs grids produced by the reference's OWN accelerator kernels (DMA /
    DMASlavePack / RMA, src/stencil/slave/*.cpp) run on the athread emulator.  Those kernels
    add in another order than the reference's naive loop (stencil_dma.cpp:436-438), so this
    is a tolerance check: the reference's abs 1e-4 (stencil.cpp:137) and north_star's max
    relative error 1e-5 for fp32, exact zeros required to agree."""
    m, N, r, T = key.split("_")
    N, r, T = int(N[1:]), int(r[1:]), int(T[1:])
    a, b = _oracle_pair((N, N), r, np.float32, 0)
    got, _ = sweep.run_host((N, N), a, b, T, radius=r)
    g = oracle.interior(got, r).astype(np.float64)
    gold = GOLD_EMU[key].astype(np.float64)
    assert np.abs(g - gold).max() < 1e-4
    assert np.array_equal(g == 0, gold == 0)
    mx = np.maximum(np.abs(g), np.abs(gold))
    assert (np.abs(g - gold)[mx > 0] / mx[mx > 0]).max() <= 1e-5


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape", [capi.STAR, capi.BOX])
@pytest.mark.parametrize("n,r,T", [
    ((1,), 1, 1), ((37,), 1, 4), ((1000,), 3, 5),
    ((1, 1), 1, 2), ((2, 3), 1, 3), ((33, 17), 2, 4), ((130, 70), 4, 3), ((257, 64), 5, 2),
    ((1, 1, 1), 1, 1), ((5, 4, 3), 1, 3), ((40, 33, 21), 2, 3), ((70, 20, 9), 3, 2),
])
def test_generic_kernel(gpu, n, r, T, shape, dtype):
    t = _check(n, r, shape, dtype, T, kernel=capi.KERNEL_GENERIC)
    assert t.kernel == capi.KERNEL_GENERIC and t.launches == T


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape", [capi.STAR, capi.BOX])
@pytest.mark.parametrize("n,T,k", [
    ((1, 1), 3, 2), ((3, 2), 5, 4), ((120, 50), 4, 4), ((121, 97), 9, 4), ((250, 31), 7, 3),
    ((500, 300), 17, 8), ((1000, 129), 16, 8), ((777, 555), 13, 5), ((4096, 64), 6, 6),
    ((119, 200), 7, 7), ((241, 40), 2, 1), ((64, 2000), 10, 2),
])
def test_warp2d_kernel(gpu, n, T, k, shape, dtype):
    t = _check(n, 1, shape, dtype, T, time_block=k, kernel=capi.KERNEL_WARP2D)
    k = min(k, 6)   # deeper requests are clamped to the family's maximum
    assert t.kernel == capi.KERNEL_WARP2D and t.time_block == k and t.launches == -(-T // k)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,r,T,k", [
    ((1, 1), 2, 3, 2), ((3, 2), 3, 5, 2), ((120, 50), 2, 4, 4), ((121, 97), 2, 9, 3), ((250, 31), 3, 7, 2),
    ((500, 300), 4, 6, 2), ((1000, 129), 2, 11, 4), ((777, 555), 3, 5, 1), ((2100, 64), 4, 3, 1),
    ((119, 200), 2, 7, 2), ((241, 40), 3, 2, 2), ((64, 2000), 4, 5, 2), ((113, 113), 2, 8, 1),
])
def test_warp2d_radius_kernel(gpu, n, r, T, k, dtype):
    """General-radius star (reference -r 2..4) through the fused warp kernel, bit-exact."""
    kmax = int(capi.lib().sb200_max_time_block(ctypes.byref(
        capi.make_desc(n, r, capi.STAR, capi.F64 if dtype == np.float64 else capi.F32, T,
                       kernel=capi.KERNEL_WARP2D))))
    assert kmax == {(2, 4): 4, (3, 4): 2, (4, 4): 2, (2, 8): 2, (3, 8): 2, (4, 8): 1}[(r, np.dtype(dtype).itemsize)]
    t = _check(n, r, capi.STAR, dtype, T, time_block=k, kernel=capi.KERNEL_WARP2D)
    ke = min(k, kmax)
    assert t.kernel == capi.KERNEL_WARP2D and t.time_block == ke and t.launches == -(-T // ke)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape", [capi.STAR, capi.BOX])
@pytest.mark.parametrize("n,r,T,k", [
    (1, 1, 3, 2), (3, 2, 5, 8), (37, 1, 4, 0), (509, 1, 20, 8), (512, 4, 9, 4), (1000, 3, 5, 5),
    (4099, 2, 33, 16), (100003, 1, 24, 0), (65536, 4, 16, 16), (20000, 3, 7, 1), (497, 1, 8, 8),
])
def test_line1d_kernel(gpu, n, r, T, k, shape, dtype):
    """1-D grids: K steps fused in registers per warp segment, bit-exact for any K."""
    t = _check((n,), r, shape, dtype, T, time_block=k)
    ke = min(k or min(16, 32 // r), 16)
    assert t.kernel == capi.KERNEL_LINE1D and t.time_block == ke and t.launches == -(-T // ke)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,T,k", [
    ((1, 1), 3, 2), ((3, 2), 5, 1), ((120, 50), 4, 2), ((121, 97), 9, 2), ((250, 31), 7, 1),
    ((1000, 129), 11, 2), ((113, 113), 8, 1), ((119, 200), 7, 2), ((64, 2000), 5, 2),
])
def test_warp2d_radius2_box_kernel(gpu, n, T, k, dtype):
    """25-point box (radius 2) through the fused warp kernel: 2R+1 accumulator chains per level,
    each in the oracle's order; bit-exact."""
    kmax = 2 if dtype == np.float32 else 1
    t = _check(n, 2, capi.BOX, dtype, T, time_block=k, kernel=capi.KERNEL_WARP2D)
    ke = min(k, kmax)
    assert t.kernel == capi.KERNEL_WARP2D and t.time_block == ke and t.launches == -(-T // ke)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("r", [3, 4])
@pytest.mark.parametrize("n,T,seed", [
    ((1, 1), 3, 1), ((3, 2), 5, 1), ((120, 50), 4, 1), ((121, 97), 6, 2), ((250, 31), 5, 1),
    ((1000, 129), 4, 3), ((113, 113), 8, 1), ((64, 700), 3, 1), ((300, 300), 12, 0),
])
def test_warp2d_radius34_box_kernel(gpu, n, T, r, seed, dtype):
    """49- / 81-point box (2-D radius 3, 4) through the warp kernel (2R+1 chains, one level per
    launch), bit-exact; AUTO picks it; radius 5 is the generic kernel's."""
    t = _check(n, r, capi.BOX, dtype, T, seed=seed, kernel=capi.KERNEL_WARP2D)
    assert t.kernel == capi.KERNEL_WARP2D and t.time_block == 1 and t.launches == T
    # AUTO picks it except for the 81-point box in fp32, where the generic kernel is faster
    auto = capi.KERNEL_GENERIC if (r == 4 and dtype == np.float32) else capi.KERNEL_WARP2D
    assert _check(n, r, capi.BOX, dtype, T, seed=seed).kernel == auto


def test_warp2d_radius5_box_is_unsupported(gpu):
    d = capi.make_desc((64, 64), radius=5, shape=capi.BOX, kernel=capi.KERNEL_WARP2D)
    a = np.zeros((74, 74), np.float32)
    assert capi.lib().sb200_run_host(ctypes.byref(d), a.ctypes.data, a.ctypes.data, None) == capi.ERR_UNSUPPORTED


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,T,k", [
    ((1, 1, 1), 3, 2), ((5, 4, 3), 4, 3), ((60, 60, 20), 4, 2), ((61, 59, 33), 5, 2),
    ((130, 70, 40), 6, 3), ((200, 130, 17), 3, 1), ((123, 125, 64), 7, 3), ((300, 64, 9), 2, 2),
])
def test_tma3d_kernel(gpu, n, T, k, dtype):
    t = _check(n, 1, capi.STAR, dtype, T, time_block=k, kernel=capi.KERNEL_TMA3D)
    k = min(k, 2)   # deeper requests are clamped to the family's maximum (K = 3 is not instantiated)
    assert t.kernel == capi.KERNEL_TMA3D and t.time_block == k and t.launches == -(-T // k)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,T,k", [
    ((1, 1, 1), 3, 2), ((5, 4, 3), 3, 1), ((60, 62, 20), 4, 2), ((61, 59, 33), 3, 1),
    ((130, 70, 40), 5, 2), ((250, 130, 17), 2, 1), ((123, 125, 64), 3, 2),
])
def test_tma3d_box27_kernel(gpu, n, T, k, dtype):
    t = _check(n, 1, capi.BOX, dtype, T, time_block=k, kernel=capi.KERNEL_TMA3D)
    assert t.kernel == capi.KERNEL_TMA3D and t.launches == -(-T // k)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,T", [
    ((1, 1, 1), 3), ((5, 4, 3), 4), ((60, 60, 20), 3), ((61, 59, 33), 5), ((130, 70, 40), 2),
    ((200, 130, 17), 3), ((123, 125, 64), 4), ((300, 64, 9), 2), ((57, 121, 5), 1),
])
@pytest.mark.parametrize("r", [2, 3, 4])
def test_tma3d_radius2_star_kernel(gpu, n, T, r, dtype):
    """13- / 19- / 25-point star (3-D radius 2..4) through the TMA kernel (z history in registers),
    bit-exact."""
    t = _check(n, r, capi.STAR, dtype, T, kernel=capi.KERNEL_TMA3D)
    assert t.kernel == capi.KERNEL_TMA3D and t.time_block == 1 and t.launches == T
    # AUTO picks it at radius 2, and at radius 3..4 in fp64 (where it beats the generic kernel)
    auto = capi.KERNEL_TMA3D if (r == 2 or dtype == np.float64) else capi.KERNEL_GENERIC
    assert _check(n, r, capi.STAR, dtype, T).kernel == auto


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,T,seed", [
    ((1, 1, 1), 3, 1), ((5, 4, 3), 4, 1), ((60, 60, 20), 3, 1), ((61, 59, 33), 5, 2), ((130, 70, 40), 2, 1),
    ((200, 130, 17), 3, 3), ((123, 125, 64), 4, 1), ((300, 64, 9), 2, 1), ((57, 121, 5), 1, 1), ((126, 60, 30), 6, 0),
    ((240, 50, 12), 3, 1), ((120, 130, 10), 2, 2), ((359, 20, 8), 2, 1),    # fp32: the 128-wide tile variant (fewer columns wasted)
])
def test_tma3d_radius2_box_kernel(gpu, n, T, seed, dtype):
    """125-point box (3-D radius 2) through the TMA kernel of tma3db.cuh (2R+1 running sums per
    point, each in the oracle's order), bit-exact; AUTO picks it; radius 3 stays generic."""
    t = _check(n, 2, capi.BOX, dtype, T, seed=seed, kernel=capi.KERNEL_TMA3D)
    assert t.kernel == capi.KERNEL_TMA3D and t.time_block == 1 and t.launches == T
    assert _check(n, 2, capi.BOX, dtype, T, seed=seed).kernel == capi.KERNEL_TMA3D
    a, b = _oracle_pair(n, 3, dtype, seed)
    with pytest.raises(capi.StencilError):
        sweep.run_host(n, a, b, T, radius=3, shape=capi.BOX, kernel=capi.KERNEL_TMA3D)


@pytest.mark.parametrize("dtype,tol", [(np.float32, 1e-5), (np.float64, 1e-12)])
@pytest.mark.parametrize("n,T,k,seed", [
    ((1, 1, 1), 3, 2, 1), ((5, 4, 3), 3, 1, 1), ((60, 62, 20), 4, 2, 1), ((61, 59, 33), 3, 1, 2),
    ((130, 70, 40), 9, 2, 3), ((123, 125, 64), 50, 2, 1), ((70, 66, 30), 6, 2, 0),
])
def test_tma3d_box27_reassociated_within_tolerance(gpu, n, T, k, seed, dtype, tol):
    """SB200_FLAG_REASSOCIATE: the separable 27-point kernel (x, y, z partial sums) against the
    reference-order oracle: max relative error <= 1e-12 (fp64) / 1e-5 (fp32) after T steps, the
    north star's tolerance, with exact zeros where the oracle has exact zeros (seed 0 = the
    reference init: zeros ahead of the diffusion front)."""
    a, b = _oracle_pair(n, 1, dtype, seed)
    ga, gb = a.copy(), b.copy()
    want = b if oracle.sweep(n, 1, capi.BOX, T, a, b, threads=oracle.max_threads()) else a
    got, t = sweep.run_host(n, ga, gb, T, radius=1, shape=capi.BOX, time_block=k, kernel=capi.KERNEL_TMA3D,
                            flags=capi.FLAG_REASSOCIATE)
    assert t.kernel == capi.KERNEL_TMA3D
    w, g = want.astype(np.float64), got.astype(np.float64)
    assert np.array_equal(w == 0, g == 0), "zero sets differ"
    scale = np.maximum(np.abs(w), np.abs(g))
    rel = np.abs(w - g)[scale > 0] / scale[scale > 0]
    assert rel.size == 0 or rel.max() <= tol, f"max relative error {rel.max():.3e}"
    # without the flag the same call is bit-exact
    ga2, gb2 = _oracle_pair(n, 1, dtype, seed)
    got2, _ = sweep.run_host(n, ga2, gb2, T, radius=1, shape=capi.BOX, time_block=k, kernel=capi.KERNEL_TMA3D)
    assert np.array_equal(want.view(np.uint8), got2.view(np.uint8))


@pytest.mark.parametrize("dtype,tol", [(np.float32, 1e-5), (np.float64, 1e-12)])
@pytest.mark.parametrize("n,T,k,seed", [
    ((1, 1), 3, 2, 1), ((5, 4), 3, 1, 1), ((131, 77), 7, 3, 1), ((300, 200), 12, 4, 2), ((97, 403), 30, 6, 3),
    ((1000, 300), 60, 4, 1), ((260, 240), 24, 6, 0),
])
def test_warp2d_box9_reassociated_within_tolerance(gpu, n, T, k, seed, dtype, tol):
    """SB200_FLAG_REASSOCIATE in 2-D: the separable 9-point level (row sums kept across rows) against
    the reference-order oracle, same bar as the 27-point test above; without the flag bit-exact."""
    a, b = _oracle_pair(n, 1, dtype, seed)
    ga, gb = a.copy(), b.copy()
    want = b if oracle.sweep(n, 1, capi.BOX, T, a, b, threads=oracle.max_threads()) else a
    got, t = sweep.run_host(n, ga, gb, T, radius=1, shape=capi.BOX, time_block=k, flags=capi.FLAG_REASSOCIATE)
    assert t.kernel == capi.KERNEL_WARP2D
    w, g = want.astype(np.float64), got.astype(np.float64)
    assert np.array_equal(w == 0, g == 0), "zero sets differ"
    scale = np.maximum(np.abs(w), np.abs(g))
    rel = np.abs(w - g)[scale > 0] / scale[scale > 0]
    assert rel.size == 0 or rel.max() <= tol, f"max relative error {rel.max():.3e}"
    if n[0] > 100:
        assert not np.array_equal(want.view(np.uint8), got.view(np.uint8)), "the flag changed nothing: separable level not used?"
    # the star has no separable form: the flag leaves it bit-exact
    _check(n, 1, capi.STAR, dtype, T, seed=seed, time_block=k, flags=capi.FLAG_REASSOCIATE)
    ga2, gb2 = _oracle_pair(n, 1, dtype, seed)
    got2, _ = sweep.run_host(n, ga2, gb2, T, radius=1, shape=capi.BOX, time_block=k)
    assert np.array_equal(want.view(np.uint8), got2.view(np.uint8))


def test_auto_dispatch(gpu):
    assert _check((200, 100), 1, capi.BOX, np.float32, 8).kernel == capi.KERNEL_WARP2D
    assert _check((80, 70, 60), 1, capi.STAR, np.float64, 4).kernel == capi.KERNEL_TMA3D
    assert _check((80, 70, 20), 1, capi.BOX, np.float64, 2).kernel == capi.KERNEL_TMA3D
    assert _check((80, 70, 20), 2, capi.BOX, np.float64, 2).kernel == capi.KERNEL_TMA3D
    assert _check((80, 70, 20), 3, capi.BOX, np.float64, 2).kernel == capi.KERNEL_GENERIC
    assert _check((80, 70, 20), 2, capi.STAR, np.float64, 2).kernel == capi.KERNEL_TMA3D
    assert _check((80, 70, 20), 3, capi.STAR, np.float64, 2).kernel == capi.KERNEL_TMA3D
    assert _check((80, 70, 20), 5, capi.STAR, np.float64, 2).kernel == capi.KERNEL_GENERIC
    assert _check((80, 70), 2, capi.STAR, np.float32, 2).kernel == capi.KERNEL_WARP2D
    assert _check((80, 70), 2, capi.BOX, np.float32, 2).kernel == capi.KERNEL_WARP2D
    assert _check((80, 70), 3, capi.BOX, np.float32, 2).kernel == capi.KERNEL_WARP2D
    assert _check((80, 70), 4, capi.BOX, np.float64, 2).kernel == capi.KERNEL_WARP2D
    assert _check((80, 70), 5, capi.BOX, np.float32, 2).kernel == capi.KERNEL_GENERIC
    assert _check((80, 70), 5, capi.STAR, np.float32, 2).kernel == capi.KERNEL_GENERIC
    assert _check((3000,), 2, capi.BOX, np.float64, 9).kernel == capi.KERNEL_LINE1D
    assert _check((3000,), 6, capi.STAR, np.float32, 3).kernel == capi.KERNEL_GENERIC


def test_ref_init_denormals_preserved(gpu):
    """fp32 reference init creates denormals ahead of the diffusion front; no flush-to-zero."""
    _check((300, 300), 1, capi.BOX, np.float32, 60, seed=0, time_block=4)
    _check((64, 64, 64), 1, capi.STAR, np.float32, 40, seed=0, time_block=2)


def test_zero_iterations_is_identity(gpu):
    a, b = _oracle_pair((20, 10), 1, np.float32, 3)
    a0 = a.copy()
    got, t = sweep.run_host((20, 10), a, b, 0)
    assert got is a and np.array_equal(a, a0) and t.launches == 0


def test_unsupported_kernel_request_is_an_error(gpu):
    d = capi.make_desc((8, 8, 8), radius=3, shape=capi.BOX, kernel=capi.KERNEL_TMA3D)   # 343-point: generic only
    a = np.zeros((14, 14, 14), np.float32)
    rc = capi.lib().sb200_run_host(ctypes.byref(d), a.ctypes.data, a.ctypes.data, None)
    assert rc == capi.ERR_UNSUPPORTED


@pytest.mark.parametrize("n,shape,dtype,k", [
    ((640, 500), capi.BOX, capi.F32, 4), ((200, 180, 150), capi.STAR, capi.F64, 2),
])
def test_device_api_fill_checksum_matches_oracle(gpu, n, shape, dtype, k):
    """Device-side init + sweep + checksum == oracle init + sweep + checksum."""
    T = 6
    d = capi.make_desc(n, 1, shape, dtype, T, time_block=k)
    g = sweep.DeviceGrids(d)
    g.fill(capi.INIT_RANDOM, seed=5)
    host0 = g.download()
    code = oracle.F64 if dtype == capi.F64 else oracle.F32
    a = oracle.fill(n, 1, code, oracle.INIT_RANDOM, 5)
    assert np.array_equal(host0.view(np.uint8), a.view(np.uint8))
    b = oracle.fill(n, 1, code, oracle.INIT_REF)
    want = b if oracle.sweep(n, 1, shape, T, a, b, threads=oracle.max_threads()) else a
    g.sweep(T)
    assert g.checksum() == oracle.checksum(n, 1, want)
    assert np.array_equal(g.download().view(np.uint8), want.view(np.uint8))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape", [capi.STAR, capi.BOX])
@pytest.mark.parametrize("n,T", [
    ((1, 1), 3), ((3, 2), 5), ((64, 64), 16), ((64, 64), 101), ((130, 70), 33), ((129, 5), 7),
    ((300, 200), 24), ((400, 400), 40), ((512, 512), 18), ((1000, 60), 9), ((127, 511), 12),
])
def test_resident2d_kernel(gpu, n, T, shape, dtype):
    """Register-resident cluster kernel (1..16 CTAs, DSMEM edge exchange): one launch for all
    T steps, bit-exact against the oracle; T parity still selects the host buffer."""
    t = _check(n, 1, shape, dtype, T, kernel=capi.KERNEL_RESIDENT2D)
    assert t.kernel == capi.KERNEL_RESIDENT2D and t.launches == 1


def test_resident2d_auto_and_limits(gpu):
    # AUTO keeps small many-step 2-D grids on chip, and falls back to warp2d otherwise
    t = _check((400, 400), 1, capi.STAR, np.float32, 64)
    assert t.kernel == capi.KERNEL_RESIDENT2D and t.launches == 1
    t = _check((400, 400), 1, capi.STAR, np.float32, 8)
    assert t.kernel == capi.KERNEL_WARP2D
    t = _check((576, 576), 1, capi.BOX, np.float32, 20)
    assert t.kernel == capi.KERNEL_RESIDENT2D
    t = _check((640, 640), 1, capi.BOX, np.float32, 20)
    assert t.kernel == capi.KERNEL_WARP2D
    t = _check((1000, 900), 1, capi.STAR, np.float32, 20)
    assert t.kernel == capi.KERNEL_WARP2D
    a, b = _oracle_pair((1000, 900), 1, np.float32, 1)
    with pytest.raises(capi.StencilError):
        sweep.run_host((1000, 900), a, b, 4, kernel=capi.KERNEL_RESIDENT2D)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape", [capi.STAR, capi.BOX])
@pytest.mark.parametrize("n,r,T,k", [
    ((70, 66, 12), 1, 4, 2), ((130, 61, 9), 1, 5, 2), ((64, 64, 40), 1, 3, 1),            # 3-D: TMA3D, faces are planes
    ((300, 40), 1, 9, 4), ((131, 24), 1, 7, 2), ((700, 300), 1, 13, 6), ((64, 17), 1, 5, 1),  # 2-D: WARP2D, faces are rows
    ((300, 64), 2, 7, 2), ((131, 40), 2, 5, 1), ((260, 90), 3, 5, 2), ((200, 70), 4, 3, 1),   # 2-D radius > 1 (warp2dr)
    ((70, 66, 16), 2, 4, 1), ((130, 61, 11), 2, 3, 1), ((70, 66, 18), 3, 3, 1), ((90, 61, 16), 4, 2, 1),  # 3-D radius 2..4 (tma3dr)
])
def test_slab_link_self_periodic(gpu, n, r, T, k, shape, dtype):
    """The fused halo exchange (sb200_sweep_slab_linked: peer stores + arrival counters) on
    ONE GPU: the slab is its own lower and upper neighbour, i.e. periodic along the outermost
    axis.  For T <= n_outer that equals the middle copy of three stacked copies swept with the
    ordinary Dirichlet oracle, bit for bit."""
    import torch
    if shape == capi.BOX and r > 2 and len(n) == 3:
        pytest.skip("no fused box kernel at this radius")
    if len(n) == 3 and r >= 3 and dtype == np.float32:
        pytest.skip("AUTO keeps the generic kernel for 3-D radius 3..4 in fp32")
    if len(n) == 2 and r == 4 and shape == capi.BOX and dtype == np.float32:
        pytest.skip("AUTO keeps the generic kernel for the 81-point box in fp32")
    d = capi.make_desc(n, r, shape, capi.F64 if dtype == np.float64 else capi.F32, T, time_block=k)
    k = min(k, int(capi.lib().sb200_max_time_block(ctypes.byref(d))))
    no = n[-1]
    g = r * k  # ghost planes / rows = radius * fused depth
    assert T * r <= no
    code = oracle.F64 if dtype == np.float64 else oracle.F32
    P = oracle.fill(n, r, code, oracle.INIT_RANDOM, 11)
    n3 = tuple(n[:-1]) + (3 * no,)
    A3, B3 = oracle.fill(n3, r, code, oracle.INIT_REF), oracle.fill(n3, r, code, oracle.INIT_REF)
    for c in range(3):
        A3[r + c * no:r + (c + 1) * no] = P[r:-r]
    want3 = B3 if oracle.sweep(n3, r, shape, T, A3, B3, threads=oracle.max_threads()) else A3
    inner = (slice(None),) + (slice(r, -r),) * (len(n) - 1)
    want = want3[r + no:r + 2 * no][inner]

    dg = sweep.DeviceGrids(d, ghost=g)
    S = np.empty((no + 2 * (r + g),) + P.shape[1:], dtype)
    for p in range(S.shape[0]):
        S[p] = P[r + (p - (r + g)) % no]
    dg.upload(S, 0)
    dg.upload(S, 1)
    flags = torch.zeros(capi.LINK_FLAG_WORDS, dtype=torch.int32, device="cuda")
    L, arrivals, cur, done = capi.lib(), 0, 0, 0
    bufs = (dg.a, dg.b)
    while done < T:
        kk = min(k, T - done)
        lk = capi.SlabLink()
        lk.peer_dst_lo = lk.peer_dst_hi = bufs[cur ^ 1].data_ptr()
        lk.flags_local = lk.flags_lo = lk.flags_hi = flags.data_ptr()
        lk.wait_arrivals = arrivals
        arr = ctypes.c_uint32(0)
        capi.check(L.sb200_sweep_slab_linked(ctypes.byref(d), g, 0, 0, bufs[cur].data_ptr(),
                                             bufs[cur ^ 1].data_ptr(), kk, ctypes.byref(lk),
                                             dg._stream(), ctypes.byref(arr)))
        assert arr.value > 0
        arrivals += arr.value
        cur ^= 1
        done += kk
    words = (ctypes.c_uint32 * capi.LINK_FLAG_WORDS)()
    capi.check(L.sb200_link_status(flags.data_ptr(), words))
    assert words[2] == 0, "a face owner timed out waiting for its neighbour"
    assert words[0] == arrivals and words[1] == arrivals
    got = dg.download(cur)[r + g:r + g + no][inner]
    assert np.array_equal(got.view(np.uint8), np.ascontiguousarray(want).view(np.uint8))


@pytest.mark.parametrize("chunks", [2, 3, 5])
@pytest.mark.parametrize("n,r,shape,dtype,T,k,kernel", [
    ((96, 80, 72), 1, capi.STAR, np.float64, 7, 2, capi.KERNEL_TMA3D),    # remainder block first
    ((70, 66, 64), 1, capi.BOX, np.float32, 6, 2, capi.KERNEL_TMA3D),
    ((130, 61, 90), 1, capi.STAR, np.float32, 9, 3, capi.KERNEL_TMA3D),
    ((300, 400), 1, capi.BOX, np.float32, 21, 4, capi.KERNEL_WARP2D),
    ((257, 300), 1, capi.STAR, np.float64, 16, 8, capi.KERNEL_WARP2D),
    ((64, 48, 80), 2, capi.STAR, np.float64, 5, 1, capi.KERNEL_GENERIC),
    ((200, 150), 3, capi.BOX, np.float32, 4, 1, capi.KERNEL_GENERIC),
])
def test_pipelined_host_path(gpu, monkeypatch, chunks, n, r, shape, dtype, T, k, kernel):
    """sb200_run_host with the grid cut into (z, t) parallelograms (upload, sweep and download
    overlapped on three streams): forced on small grids here; same bits as the oracle."""
    monkeypatch.setenv("SB200_PIPELINE_CHUNKS", str(chunks))
    t = _check(n, r, shape, dtype, T, time_block=k, kernel=kernel)
    k = t.time_block   # requests beyond the family's maximum are clamped
    blocks = -(-T // k)
    wmin = 4 * k * r
    last = max(T * r, wmin)
    if n[-1] > last:
        assert t.launches > blocks, "the pipelined path was not taken"
    monkeypatch.setenv("SB200_PIPELINE_CHUNKS", "1")
    t1 = _check(n, r, shape, dtype, T, time_block=k, kernel=kernel)
    assert t1.launches == blocks


@pytest.mark.parametrize("n,r,shape,dtype,T,k,kernel", [
    ((96, 80, 40), 1, capi.STAR, np.float64, 7, 2, capi.KERNEL_TMA3D),
    ((70, 66, 33), 1, capi.BOX, np.float32, 5, 2, capi.KERNEL_TMA3D),
    ((300, 90), 1, capi.BOX, np.float32, 13, 4, capi.KERNEL_WARP2D),
    ((257, 120), 1, capi.STAR, np.float64, 16, 6, capi.KERNEL_WARP2D),
    ((200, 150), 2, capi.STAR, np.float32, 9, 4, capi.KERNEL_WARP2D),   # radius kernels: ghost = r*k rows
    ((130, 140), 3, capi.STAR, np.float64, 5, 2, capi.KERNEL_WARP2D),
    ((150, 101), 4, capi.STAR, np.float32, 4, 2, capi.KERNEL_WARP2D),
    ((64, 48, 30), 2, capi.STAR, np.float64, 3, 1, capi.KERNEL_GENERIC),
    ((120, 90), 3, capi.BOX, np.float32, 4, 1, capi.KERNEL_GENERIC),
])
def test_two_slabs_on_one_gpu(gpu, n, r, shape, dtype, T, k, kernel):
    """The slab path's kernels (ghost planes/rows that are ordinary updatable points, update
    domain open on the internal face) without a second GPU: the grid is cut in two along the
    outermost axis, each slab is swept k steps at a time through sb200_sweep_device with
    ghost = r*k, and the ghosts are refreshed from the other slab on the host in between.
    Same bits as the single-grid oracle."""
    code = oracle.F64 if dtype == np.float64 else oracle.F32
    a = oracle.fill(n, r, code, oracle.INIT_RANDOM, 9)
    b = oracle.fill(n, r, code, oracle.INIT_REF)
    G = a.copy()
    want = b if oracle.sweep(n, r, shape, T, a, b, threads=oracle.max_threads()) else a
    nz = n[-1]
    cuts = [0, nz // 2 + 1, nz]
    g = r * k
    slabs = []
    for s in range(2):
        ns = tuple(n[:-1]) + (cuts[s + 1] - cuts[s],)
        d = capi.make_desc(ns, r, shape, capi.F64 if dtype == np.float64 else capi.F32, T,
                           time_block=k, kernel=kernel)
        slabs.append((d, sweep.DeviceGrids(d, ghost=g), ns))

    def scatter(s):
        """Host image of slab s (halo + ghost + interior + ghost + halo) cut out of G."""
        d, dg, ns = slabs[s]
        S = np.zeros((ns[-1] + 2 * (r + g),) + G.shape[1:], dtype)
        for p in range(S.shape[0]):
            zg = r + cuts[s] + (p - (r + g))  # padded index into G
            if 0 <= zg < G.shape[0]:
                S[p] = G[zg]
        return S

    done = 0
    while done < T:
        kk = min(k, T - done)
        for s in range(2):
            d, dg, ns = slabs[s]
            S = scatter(s)
            dg.upload(S, 0)
            dg.upload(S, 1)
            dg.cur = 0
            nl = dg.sweep(kk, lo_boundary=int(s == 0), hi_boundary=int(s == 1))
            assert nl == 1
        for s in range(2):
            d, dg, ns = slabs[s]
            R = dg.download()
            G[r + cuts[s]:r + cuts[s + 1]] = R[r + g:r + g + ns[-1]]
        done += kk
    assert np.array_equal(G.view(np.uint8), want.view(np.uint8))
