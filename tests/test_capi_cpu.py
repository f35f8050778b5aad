"""CPU tests of the C-ABI boundary: the library loads, exports every symbol
include/stencil_b200.h declares, validates descriptors, and FAILS LOUDLY without a GPU
(no compute is attempted here)."""
import ctypes
import os
import re

import numpy as np
import pytest

from stencil_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_all_exported(built):
    hdr = open(os.path.join(ROOT, "include", "stencil_b200.h")).read()
    declared = set(re.findall(r"\b(sb200_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(capi.SYMBOLS)
    L = capi.lib()
    for s in declared:
        assert getattr(L, s) is not None
    assert L.sb200_abi_version() == 4


def test_struct_sizes_match_header(built):
    assert ctypes.sizeof(capi.Desc) == 64
    assert ctypes.sizeof(capi.Timing) == 56
    assert ctypes.sizeof(capi.SlabPlan) == 8 + 32 * capi.MAX_GPUS
    assert ctypes.sizeof(capi.Layout) == 56
    assert ctypes.sizeof(capi.SlabLink) == 48   # 5 pointers + uint32 (+ padding)
    assert ctypes.sizeof(capi.PlanInfo) == 32


def test_layout_is_aligned(built):
    for dtype, es in ((capi.F32, 4), (capi.F64, 8)):
        for n in ((5,), (100, 7), (512, 512, 512), (1023, 3, 2)):
            lay = capi.layout_for(capi.make_desc(n, dtype=dtype), 0)
            assert (lay.ox * es) % 128 == 0 and (lay.pitch_x * es) % 128 == 0
            assert lay.pitch_x >= lay.ox + n[0] + 1
            assert lay.bytes == lay.pitch_x * lay.pitch_y * lay.planes * es
    lay = capi.layout_for(capi.make_desc((8, 8, 8)), ghost=3)
    assert lay.planes == 8 + 2 * (1 + 3) and lay.oz == 4


@pytest.mark.parametrize("kw", [dict(n=(0, 4)), dict(n=(4, 4), radius=0), dict(n=(4, 4), radius=17),
                                dict(n=(4, 4), shape=7), dict(n=(4, 4), dtype=5)])
def test_invalid_descriptors_rejected(built, kw):
    d = capi.make_desc(**kw)
    lay = capi.Layout()
    assert capi.lib().sb200_layout_for(ctypes.byref(d), 0, ctypes.byref(lay)) == capi.ERR_INVALID
    assert capi.lib().sb200_last_error()


def test_no_gpu_means_loud_failure(built):
    L = capi.lib()
    if L.sb200_device_count() > 0:
        pytest.skip("a GPU is visible")
    a = np.zeros((6, 6), np.float32)
    b = np.zeros((6, 6), np.float32)
    d = capi.make_desc((4, 4))
    rc = L.sb200_run_host(ctypes.byref(d), a.ctypes.data, b.ctypes.data, None)
    assert rc == capi.ERR_NO_DEVICE
    assert b"no CPU fallback" in L.sb200_last_error()
    with pytest.raises(capi.StencilError):
        capi.check(rc)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "stencil_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "liboracle" not in src, f


def test_plan_host_logic(monkeypatch):
    """sb200_plan: kernel choice, fused depth, tiling and pipelining decisions need no GPU."""
    monkeypatch.delenv("SB200_PIPELINE_CHUNKS", raising=False)
    # config 2: 512^3 fp64 7-point, T=100 -> TMA3D k=2, 9x9 tiles of 60x60, 9 z chunks, 11 parallelograms
    p = capi.plan(capi.make_desc((512, 512, 512), 1, capi.STAR, capi.F64, 100), 100)
    assert (p.kernel, p.time_block, p.launches) == (capi.KERNEL_TMA3D, 2, 50)
    assert (p.tiles_x, p.tiles_y) == (9, 9) and p.chunks * p.chunk_extent >= 512
    assert p.chunks * p.tiles_x * p.tiles_y >= 148 and p.pipeline_chunks == 11
    # config 4: 16384^2 fp32 9-point, T=1000 -> WARP2D k=4, one strip per ~120 columns
    p = capi.plan(capi.make_desc((16384, 16384), 1, capi.BOX, capi.F32, 1000), 1000)
    assert (p.kernel, p.time_block, p.launches) == (capi.KERNEL_WARP2D, 4, 250)
    assert p.tiles_x == -(-16384 // 120) and p.chunks * p.chunk_extent >= 16384 and p.pipeline_chunks == 11
    # the reference's own run.sh grid: resident in one cluster, one launch, not pipelined
    p = capi.plan(capi.make_desc((400, 400), 1, capi.STAR, capi.F32, 1000), 1000)
    assert (p.kernel, p.launches, p.pipeline_chunks) == (capi.KERNEL_RESIDENT2D, 1, 1)
    assert p.tiles_x == 4 and p.tiles_x * p.tiles_y <= 16 * p.chunks and p.tiles_y * p.chunk_extent >= 400
    # too few steps for the resident kernel, too large for it, radius 2, 1-D
    assert capi.plan(capi.make_desc((400, 400), 1, capi.STAR, capi.F32, 8), 8).kernel == capi.KERNEL_WARP2D
    p = capi.plan(capi.make_desc((960, 960), 1, capi.STAR, capi.F32, 1000), 1000)
    assert p.kernel == capi.KERNEL_WARP2D and p.chunk_extent < 64  # small-grid planner: short chunks
    p = capi.plan(capi.make_desc((400, 400), 2, capi.STAR, capi.F32, 100), 100)
    assert (p.kernel, p.time_block, p.launches) == (capi.KERNEL_WARP2D, 4, 25)  # radius-2 star: fused warp kernel
    p = capi.plan(capi.make_desc((400, 400), 4, capi.STAR, capi.F64, 100), 100)
    assert (p.kernel, p.time_block, p.launches) == (capi.KERNEL_WARP2D, 1, 100)
    p = capi.plan(capi.make_desc((400, 400), 2, capi.BOX, capi.F32, 100), 100)
    assert (p.kernel, p.time_block, p.launches) == (capi.KERNEL_WARP2D, 2, 50)   # 25-point box: fused warp kernel
    p = capi.plan(capi.make_desc((400, 400), 2, capi.BOX, capi.F64, 100), 100)
    assert (p.kernel, p.time_block) == (capi.KERNEL_WARP2D, 1)
    p = capi.plan(capi.make_desc((400, 400), 3, capi.BOX, capi.F32, 100), 100)
    assert (p.kernel, p.time_block, p.launches) == (capi.KERNEL_WARP2D, 1, 100)  # 49-point box: warp kernel, one level
    assert capi.plan(capi.make_desc((400, 400), 5, capi.BOX, capi.F32, 100), 100).kernel == capi.KERNEL_GENERIC
    # config 1 in fp64 (k=4: 160 registers, 12 resident warps per SM): 64-row chunks would make 2240 warp items for
    # 1776 places — a full round and a nearly empty one; the planner lengthens the chunks to fit one round.  With 16
    # resident warps (2368 places) the items already fit, and config 4's many rounds keep their short chunks.
    c1 = capi.make_desc((4096, 4096), 1, capi.STAR, capi.F64, 100)
    base = capi.plan(c1, 100)
    assert (base.tiles_x, base.chunk_extent, base.chunks) == (35, 64, 64)
    monkeypatch.setenv("SB200_PLAN_RESIDENT_WARPS", "12")
    fit = capi.plan(c1, 100)
    assert (fit.tiles_x, fit.chunk_extent, fit.chunks) == (35, 82, 50) and fit.tiles_x * fit.chunks <= 148 * 12
    c4 = capi.make_desc((16384, 16384), 1, capi.BOX, capi.F32, 100)
    assert capi.plan(c4, 100).chunk_extent == 160
    monkeypatch.setenv("SB200_PLAN_RESIDENT_WARPS", "16")
    assert capi.plan(c1, 100).chunk_extent == 64 and capi.plan(c4, 100).chunk_extent == 160
    monkeypatch.delenv("SB200_PLAN_RESIDENT_WARPS")
    # 125-point box (3-D radius 2): TMA kernel, one step per launch; fp64 tiles store 60 x 60, fp32 tiles
    # 120 x 60 or 56 x 124, whichever computes fewer columns for this row length
    p = capi.plan(capi.make_desc((512, 512, 512), 2, capi.BOX, capi.F64, 4), 4)
    assert (p.kernel, p.time_block, p.launches, p.tiles_x, p.tiles_y) == (capi.KERNEL_TMA3D, 1, 4, 9, 9)
    p = capi.plan(capi.make_desc((512, 512, 512), 2, capi.BOX, capi.F32, 4), 4)
    assert (p.kernel, p.tiles_x, p.tiles_y) == (capi.KERNEL_TMA3D, 5, 9)             # 5 x 128 = 10 x 64: the wide tile
    p = capi.plan(capi.make_desc((130, 70, 40), 2, capi.BOX, capi.F32, 4), 4)
    assert (p.kernel, p.tiles_x, p.tiles_y) == (capi.KERNEL_TMA3D, 3, 1)             # 3 x 64 < 2 x 128: the narrow tile
    assert capi.plan(capi.make_desc((64, 64, 64), 3, capi.BOX, capi.F32, 4), 4).kernel == capi.KERNEL_GENERIC
    assert capi.plan(capi.make_desc((400, 400), 5, capi.STAR, capi.F32, 100), 100).kernel == capi.KERNEL_GENERIC
    # 1-D: warp segments of 512 points; 16 fused steps by default at r=1 (10 here: one launch,
    # 512 - 12 - 10 -> 488 stored per segment), 8 at r=4
    p = capi.plan(capi.make_desc((1 << 20,), 1, capi.STAR, capi.F64, 10), 10)
    assert (p.kernel, p.time_block, p.launches, p.pipeline_chunks) == (capi.KERNEL_LINE1D, 16, 1, 1)
    assert p.chunk_extent == 488 and p.tiles_x == -(-(1 << 20) // 488)
    p = capi.plan(capi.make_desc((1 << 20,), 4, capi.BOX, capi.F32, 100), 100)
    assert (p.kernel, p.time_block, p.launches, p.chunk_extent) == (capi.KERNEL_LINE1D, 8, 13, 448)
    p = capi.plan(capi.make_desc((1 << 20,), 5, capi.STAR, capi.F64, 10), 10)
    assert (p.kernel, p.launches) == (capi.KERNEL_GENERIC, 10)
    # T*r too long for the grid: the last parallelogram cannot start inside it -> not pipelined
    p = capi.plan(capi.make_desc((512, 512, 512), 1, capi.STAR, capi.F64, 600), 600)
    assert p.pipeline_chunks == 1
    monkeypatch.setenv("SB200_PIPELINE_CHUNKS", "3")
    assert capi.plan(capi.make_desc((96, 80, 72), 1, capi.STAR, capi.F64, 7), 7).pipeline_chunks == 3
    with pytest.raises(capi.StencilError):
        capi.plan(capi.make_desc((2000, 2000), 1, capi.STAR, capi.F32, 10, kernel=capi.KERNEL_RESIDENT2D), 10)


def test_block_hint_is_a_tile_preference_when_asked(built):
    """-b/--block-size (reference program_options.cpp:16-18) steers the chunk the planners give
    one warp / CTA when SB200_FLAG_BLOCK_TILE is set, and is inert otherwise."""
    base = capi.plan(capi.make_desc((4096, 4096), dtype=capi.F64, iterations=100, block_hint=64), 100)
    assert base.kernel == capi.KERNEL_WARP2D
    for b in (8, 64, 500):
        p = capi.plan(capi.make_desc((4096, 4096), dtype=capi.F64, iterations=100, block_hint=b,
                                     flags=capi.FLAG_BLOCK_TILE), 100)
        assert p.chunk_extent == b and p.chunks == -(-4096 // b) and p.tiles_x == base.tiles_x
    inert = capi.plan(capi.make_desc((4096, 4096), dtype=capi.F64, iterations=100, block_hint=8), 100)
    assert (inert.chunk_extent, inert.chunks) == (base.chunk_extent, base.chunks)
    p3 = capi.plan(capi.make_desc((512, 512, 512), dtype=capi.F64, iterations=100, block_hint=32,
                                  flags=capi.FLAG_BLOCK_TILE), 100)
    assert p3.kernel == capi.KERNEL_TMA3D and p3.chunk_extent == 32 and p3.chunks == 16
    big = capi.plan(capi.make_desc((512, 512, 512), dtype=capi.F64, iterations=100, block_hint=100000,
                                   flags=capi.FLAG_BLOCK_TILE), 100)
    assert big.chunk_extent == 512 and big.chunks == 1
    d = capi.make_desc((64, 64), flags=8)
    lay = capi.Layout()
    assert capi.lib().sb200_layout_for(ctypes.byref(d), 0, ctypes.byref(lay)) == capi.ERR_INVALID
