"""GPU test (-m gpu) of sb200_checksum_device_noalloc, the entry point for callers that cannot
allocate (no counterpart in the reference).

Its scratch is a file-scope `static __device__` array shared by every call, so the checks are:
the digest equals sb200_checksum_device's bit for bit (integer work: the tolerance of this
suite is equality) on two calls in a row over DIFFERENT contents (stale scratch would show),
for both dtypes, 2-D / 3-D and a slab offset; grids beyond the stated limit are rejected.

Collected last on purpose (file name).  Runs alone in a fresh
process as well as after the other tests (it only needs the `gpu` fixture)."""
import ctypes

import pytest

from stencil_b200 import capi, sweep

pytestmark = pytest.mark.gpu


def _both(L, g, z_offset=0, z_total=0):
    """(existing entry point's digest, no-allocation entry point's digest) of g's current buffer."""
    want = g.checksum(z_offset, z_total)
    out = ctypes.c_uint64(0)
    capi.check(L.sb200_checksum_device_noalloc(ctypes.byref(g.desc), g.ghost, g.bufs()[0].data_ptr(),
                                               z_offset, z_total, g._stream(), ctypes.byref(out)))
    return want, out.value


@pytest.mark.parametrize("n,shape,dtype", [
    ((200, 180, 150), capi.STAR, capi.F64),     # 5.4 M points: 330 partial sums
    ((640, 500), capi.BOX, capi.F32),           # 2-D, fp32
    ((37, 5, 3), capi.STAR, capi.F64),          # one block, ragged
    ((512, 512, 512), capi.STAR, capi.F64),     # exactly the limit: 8192 partial sums
])
def test_noalloc_checksum_matches_on_two_calls_in_a_row(gpu, n, shape, dtype):
    d = capi.make_desc(n, 1, shape, dtype, 4)
    g = sweep.DeviceGrids(d)
    g.fill(capi.INIT_RANDOM, seed=5)
    want1, got1 = _both(gpu, g)
    assert got1 == want1
    g.sweep(3)                                   # other contents, other buffer
    want2, got2 = _both(gpu, g)
    assert got2 == want2 and want2 != want1
    # and once more over the first contents' digest path: the ticket was put back to zero
    out = ctypes.c_uint64(0)
    capi.check(gpu.sb200_checksum_device_noalloc(ctypes.byref(d), 0, g.bufs()[0].data_ptr(), 0, 0,
                                                 g._stream(), ctypes.byref(out)))
    assert out.value == want2


def test_noalloc_checksum_of_a_slab(gpu):
    """ghost planes + position inside a taller grid (z_offset / z_total), like the slab drivers use."""
    d = capi.make_desc((96, 80, 40), 1, capi.STAR, capi.F64, 2)
    g = sweep.DeviceGrids(d, ghost=2)
    g.fill(capi.INIT_RANDOM, seed=9, z_offset=40, z_total=120)
    for _ in range(2):
        want, got = _both(gpu, g, z_offset=40, z_total=120)
        assert got == want


def test_noalloc_rejects_grids_beyond_the_limit(gpu):
    small = capi.make_desc((64, 64, 64), dtype=capi.F64)
    g = sweep.DeviceGrids(small)
    g.fill(capi.INIT_REF)
    big = capi.make_desc((512, 512, 513), dtype=capi.F64)
    assert 512 * 512 * 513 > capi.NOALLOC_MAX_POINTS
    out = ctypes.c_uint64(0)
    rc = gpu.sb200_checksum_device_noalloc(ctypes.byref(big), 0, g.a.data_ptr(), 0, 0, g._stream(), ctypes.byref(out))
    assert rc == capi.ERR_INVALID
    assert b"exceeds the no-allocation limit" in gpu.sb200_last_error()
    # the refusal touched nothing: the next call is right
    want, got = _both(gpu, g)
    assert got == want
