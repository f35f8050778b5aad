"""CPU tests of the oracle (oracle/stencil_oracle.c) against the reference's golden data.

Golden grids in tests/golden/ref_2d_f32.npz were produced by the GENUINE reference naive
loop (Stencil::check_result, /root/reference/src/stencil/stencil.cpp:75-131) via
tests/golden/make_golden.py.  Known answers are the ones SURVEY.md §8(c) lists.
tests/golden/ref_2d_f64.npz is the same reference source compiled with float -> double
(oracle/Makefile: _ref/libref64.so): the reference's loop nest and add order at fp64.
"""
import os

import numpy as np
import pytest

import oracle

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_2d_f32.npz"))


def _case(key):
    n, r, t = key.split("_")
    return int(n[1:]), int(r[1:]), int(t[1:])


@pytest.mark.parametrize("key", sorted(GOLD.files))
def test_oracle_matches_reference_golden_bitwise(key):
    N, r, T = _case(key)
    got = oracle.interior(oracle.run((N, N), r, oracle.STAR, T), r)
    assert np.array_equal(got.view(np.uint32), GOLD[key].view(np.uint32))


GOLD64 = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_2d_f64.npz"))


@pytest.mark.parametrize("key", sorted(GOLD64.files))
def test_oracle_matches_reference_golden_f64_bitwise(key):
    N, r, T = _case(key)
    got = oracle.interior(oracle.run((N, N), r, oracle.STAR, T, dtype=oracle.F64), r)
    assert got.dtype == np.float64
    assert np.array_equal(got.view(np.uint64), GOLD64[key].view(np.uint64))


def test_f64_golden_is_not_the_f32_grid():
    """The fp64 fixture really is double arithmetic: it differs from the fp32 grid where
    avg = 1/(4r) is not a dyadic rational."""
    assert np.abs(GOLD64["N32_r3_T5"] - GOLD["N32_r3_T5"].astype(np.float64)).max() > 0


KNOWN = [  # (N, r, T, element (0,0)) from the reference loop, SURVEY.md §8(c)
    (64, 1, 1, 0.25), (64, 1, 2, 0.3125), (64, 1, 3, 0.359375),
    (64, 1, 10, 0.444484710693359), (64, 1, 100, 0.493727654218674),
    (64, 2, 1, 0.25), (64, 2, 3, 0.373046875), (64, 3, 5, 0.421959400177002),
    (400, 1, 1000, 0.499364346265793),
]


@pytest.mark.parametrize("N,r,T,want", KNOWN)
def test_known_answers(N, r, T, want):
    g = oracle.run((N, N), r, oracle.STAR, T)
    assert abs(float(oracle.interior(g, r)[0, 0]) - want) < 5e-16


def test_known_answer_4096():
    # N=4096,r=1,T=10 -> 0.444484710693359 ; r=2 -> 0.455086767673492
    for r, want in ((1, 0.444484710693359), (2, 0.455086767673492)):
        g = oracle.run((4096, 4096), r, oracle.STAR, 10, threads=oracle.max_threads())
        assert abs(float(oracle.interior(g, r)[0, 0]) - want) < 5e-16


@pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref/libref.so not built here")
@pytest.mark.parametrize("N,r,T", [(7, 1, 3), (12, 2, 5), (20, 1, 17)])
def test_oracle_vs_live_reference(N, r, T):
    ref = oracle.ref_naive_2d_f32(N, r, T)
    got = oracle.interior(oracle.run((N, N), r, oracle.STAR, T), r)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_halo_never_written_and_parity_rule():
    n, r = (9, 7), 2
    a = oracle.fill(n, r, oracle.F32, oracle.INIT_RANDOM, seed=3)
    b = oracle.fill(n, r, oracle.F32, oracle.INIT_REF)
    halo_a, halo_b = a.copy(), b.copy()
    assert oracle.sweep(n, r, oracle.STAR, 3, a, b) == 1      # odd T -> result in b
    assert oracle.sweep(n, r, oracle.STAR, 2, a, b) == 0      # even T -> back in a
    for g, h in ((a, halo_a), (b, halo_b)):
        m = np.ones(g.shape, bool)
        m[r:-r, r:-r] = False
        assert np.array_equal(g[m], h[m])


def test_ref_init_layout():
    # left/right halo columns 1 over the full padded height (corners included), top/bottom
    # halo 0, interior 0: Stencil::generate_initialized_matrix stencil.cpp:199-204
    g = oracle.fill((5, 4), 2, oracle.F32)
    assert g.shape == (8, 9)
    assert (g[:, :2] == 1).all() and (g[:, -2:] == 1).all() and (g[:, 2:-2] == 0).all()
    g3 = oracle.fill((3, 4, 5), 1, oracle.F64)
    assert g3.shape == (7, 6, 5)
    assert (g3[:, :, 0] == 1).all() and (g3[:, :, -1] == 1).all() and (g3[:, :, 1:-1] == 0).all()


@pytest.mark.parametrize("dtype", [oracle.F32, oracle.F64])
@pytest.mark.parametrize("shape", [oracle.STAR, oracle.BOX])
@pytest.mark.parametrize("init", [oracle.INIT_RANDOM, oracle.INIT_REF])
@pytest.mark.parametrize("n,r", [((17,), 1), ((6, 5), 1), ((5, 4, 3), 1), ((23,), 3), ((11, 9), 2), ((9, 12), 4),
                                 ((7, 6, 5), 2), ((9, 8, 7), 3)])
def test_oracle_matches_numpy_restatement(n, r, init, shape, dtype):
    """Independent numpy restatement (same add order) for every ndim / shape / dtype, radius
    1..4, and both initial conditions — the reference one puts 1 on the x faces over the full
    padded extent of the other axes, so box corners see it."""
    T = 3
    a = oracle.fill(n, r, dtype, init, seed=11)
    b = oracle.fill(n, r, dtype, oracle.INIT_REF)
    x, y = a.copy(), b.copy()
    nd = len(n)
    dt = a.dtype.type
    inner = tuple(slice(r, s - r) for s in a.shape)

    def sh(arr, off):  # off in (z,y,x)-minor order matching numpy axes
        return arr[tuple(slice(r + o, s - r + o) for s, o in zip(arr.shape, off))]

    for _ in range(T):
        s = np.zeros(x[inner].shape, dt)
        if shape == oracle.STAR:
            for ax in range(nd):               # x first = last numpy axis
                npax = nd - 1 - ax
                # low side far -> near, then high side near -> far (stencil.cpp:103-123)
                for k in list(range(-r, 0)) + list(range(1, r + 1)):
                    off = [0] * nd
                    off[npax] = k
                    s = (s + sh(x, off)).astype(dt)
            avg = dt(1) / dt(2 * nd * r)
        else:
            import itertools
            for off in itertools.product(*[range(-r, r + 1)] * nd):   # z outer ... x inner, ascending
                s = (s + sh(x, off)).astype(dt)
            avg = dt(1) / dt((2 * r + 1) ** nd)
        y[inner] = s * avg
        x, y = y, x
    want = b if oracle.sweep(n, r, shape, T, a, b) else a
    assert np.array_equal(want.view(np.uint8), x.view(np.uint8))


def test_openmp_is_bit_identical():
    n = (40, 30, 20)
    one = oracle.run(n, 1, oracle.BOX, 4, oracle.F64, oracle.INIT_RANDOM, 5, threads=1)
    many = oracle.run(n, 1, oracle.BOX, 4, oracle.F64, oracle.INIT_RANDOM, 5, threads=4)
    assert np.array_equal(one.view(np.uint8), many.view(np.uint8))


def test_slab_fill_and_checksum_compose():
    """Slabs generated independently equal slices of the global grid; checksums add up."""
    n, r = (6, 5, 8), 1
    whole = oracle.fill(n, r, oracle.F64, oracle.INIT_RANDOM, seed=9)
    total = oracle.checksum(n, r, whole)
    acc = 0
    for off in (0, 4):
        slab = oracle.fill((6, 5, 4), r, oracle.F64, oracle.INIT_RANDOM, seed=9, z_offset=off, z_total=8)
        assert np.array_equal(slab, whole[off:off + 4 + 2 * r])
        acc = (acc + oracle.checksum((6, 5, 4), r, slab, z_offset=off, z_total=8)) % 2 ** 64
    assert acc == total


# ---- the reference's ACCELERATOR kernels (DMA / DMASlavePack / RMA) on the athread emulator ----
# tests/golden/ref_emu_2d_f32.npz: grids produced by the unmodified src/stencil/slave/*.cpp
# kernels through Stencil::run on 64 emulated CPEs (oracle/ref_shim/emu/).  Their add order
# (0.25f*(up+left+right+down), stencil_dma.cpp:436-438; row sum + column sum - 2*centre for
# r > 1, :636-651) is not the naive loop's, so the reference's own kernels agree with its own
# check only within its tolerance (abs 1e-4, stencil.cpp:137) — measured: <= 2 ulp of 0.5.
GOLD_EMU = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_emu_2d_f32.npz"))
REF_ABS_TOL = 1e-4   # stencil.cpp:137
NORTH_STAR_REL_F32 = 1e-5


def _emu_case(key):
    m, n, r, t = key.split("_")
    return m, int(n[1:]), int(r[1:]), int(t[1:])


def rel_err(a, b):
    """max |a-b| / max(|a|,|b|) with exact zeros required to agree."""
    a = a.astype(np.float64)
    b = b.astype(np.float64)
    assert np.array_equal(a == 0, b == 0)
    m = np.maximum(np.abs(a), np.abs(b))
    return float((np.abs(a - b)[m > 0] / m[m > 0]).max()) if (m > 0).any() else 0.0


@pytest.mark.parametrize("key", sorted(GOLD_EMU.files))
def test_oracle_vs_reference_accelerator_kernels(key):
    m, N, r, T = _emu_case(key)
    got = oracle.interior(oracle.run((N, N), r, oracle.STAR, T), r)
    gold = GOLD_EMU[key]
    assert np.abs(got.astype(np.float64) - gold).max() < REF_ABS_TOL
    assert np.abs(got.astype(np.float64) - gold).max() <= 4 * 2.0 ** -24  # measured: 2 ulp of 0.5
    assert rel_err(got, gold) <= NORTH_STAR_REL_F32


needs_emu = pytest.mark.skipif(oracle.ref_emu_lib() is None,
                               reason="oracle/_ref/libref_emu.so needs /root/reference")


@needs_emu
@pytest.mark.parametrize("key", sorted(GOLD_EMU.files))
def test_emulated_reference_kernels_reproduce_golden(key):
    """Determinism of the emulation (64 threads, barriers, RMA reply counters) and the
    reference's own verdict on its kernels."""
    m, N, r, T = _emu_case(key)
    g, ok = oracle.ref_emu_run_2d_f32(m, N, N // 8, r, T)
    assert ok
    assert np.array_equal(g.view(np.uint32), GOLD_EMU[key].view(np.uint32))


@needs_emu
def test_emulated_transports_agree_bitwise():
    """DMA, DMASlavePack and RMA only differ in how halos travel: same bits."""
    a, _ = oracle.ref_emu_run_2d_f32("DMA", 128, 16, 1, 77)
    b, _ = oracle.ref_emu_run_2d_f32("DMASlavePack", 128, 16, 1, 77)
    c, _ = oracle.ref_emu_run_2d_f32("RMA", 128, 16, 1, 77)
    assert np.array_equal(a, b) and np.array_equal(a, c)


@needs_emu
def test_emulator_rejects_what_the_kernels_cannot_run():
    with pytest.raises(ValueError):
        oracle.ref_emu_run_2d_f32("RMA", 64, 8, 2, 3)       # radius 1 only
    with pytest.raises(ValueError):
        oracle.ref_emu_run_2d_f32("DMA", 60, 8, 1, 3)       # N must be 8 * block
    with pytest.raises(ValueError):
        oracle.ref_emu_run_2d_f32("DMAStaticUnroll", 64, 8, 1, 3)
