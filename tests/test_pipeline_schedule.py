"""CPU proof obligations of the pipelined drop-in call (sb200_run_host on large grids).

sb200_pipeline_schedule hands out the exact op list run_host_pipelined executes (capi.cu:
build_pipeline): uploads, re-pitch kernels, range sweeps and downloads on 2 + NS streams with
cross-stream event edges.  Two things are checked here without a GPU:

 1. RACE FREEDOM — for every pair of ops not ordered by happens-before (stream order + event
    edges, transitively), their read / write sets (resource, plane range) must not conflict.
 2. FUNCTION — executing the list with the CPU oracle standing in for the kernels, on
    NaN-poisoned device grids, in list order and in random happens-before-respecting orders,
    reproduces the plain T-step sweep bit for bit (reference semantics: stencil.cpp:88-134).
"""
import os
import random

import numpy as np
import pytest

import oracle
from stencil_b200 import capi

U, P, S, D = capi.PIPE_UPLOAD, capi.PIPE_UNPACK, capi.PIPE_SWEEP, capi.PIPE_DOWNLOAD


def _schedule(monkeypatch, n, r, shape, dtype, T, C, NS, kernel=capi.KERNEL_AUTO):
    monkeypatch.setenv("SB200_PIPELINE_CHUNKS", str(C))
    monkeypatch.setenv("SB200_PIPELINE_STREAMS", str(NS))
    d = capi.make_desc(n, radius=r, shape=shape, dtype=dtype, iterations=T, kernel=kernel)
    return capi.pipeline_schedule(d)


def _preds(ops):
    """Direct predecessors: the previous op of the same stream + the wait edges."""
    last = {}
    out = []
    for j, o in enumerate(ops):
        p = [w for w in o.wait if w >= 0]
        assert all(w < j for w in p), "wait edges must point backwards (list order is an issue order)"
        if o.stream in last:
            p.append(last[o.stream])
        last[o.stream] = j
        out.append(p)
    return out


def _reach(ops):
    """reach[j] = bitset of ops that happen before op j."""
    reach = []
    for p in _preds(ops):
        m = 0
        for q in p:
            m |= reach[q] | (1 << q)
        reach.append(m)
    return reach


def _accesses(o, n_outer, r, T):
    """[(resource, lo, hi, is_write)] in PADDED outermost indices."""
    host_dst = "host_out" if T & 1 else "host_in"
    if o.kind == U:
        return [("host_in", o.lo, o.hi, False), ("staging", o.lo, o.hi, True)]
    if o.kind == P:
        return [("staging", o.lo, o.hi, False), ("g0", o.lo, o.hi, True), ("g1", o.lo, o.hi, True)]
    if o.kind == S:
        k = o.steps
        lo = max(o.lo - k * r, -r) + r
        hi = min(o.hi + k * r, n_outer + r) + r
        return [(f"g{o.src}", lo, hi, False), (f"g{o.src ^ 1}", o.lo + r, o.hi + r, True)]
    return [(f"g{o.src}", o.lo + r, o.hi + r, False), (host_dst, o.lo + r, o.hi + r, True)]


def _check_race_free(ops, n_outer, r, T):
    reach = _reach(ops)
    acc = [_accesses(o, n_outer, r, T) for o in ops]
    checked = 0
    for b in range(len(ops)):
        for a in range(b):
            if (reach[b] >> a) & 1:
                continue
            checked += 1
            for ra, la, ha, wa in acc[a]:
                for rb, lb, hb, wb in acc[b]:
                    if ra == rb and (wa or wb) and la < hb and lb < ha:
                        raise AssertionError(
                            f"unordered ops {a} and {b} conflict on {ra}: [{la},{ha}) vs [{lb},{hb}) "
                            f"kinds {ops[a].kind}/{ops[b].kind} parallelograms "
                            f"{ops[a].parallelogram}/{ops[b].parallelogram} blocks {ops[a].block}/{ops[b].block}")
    return checked


def _check_structure(ops, n_outer, r, T):
    """Coverage: every padded index uploaded and re-pitched exactly once, every interior index
    downloaded exactly once, every sweep inside the grid."""
    padded = n_outer + 2 * r
    for kind, extent in ((U, padded), (P, padded), (D, n_outer)):
        seen = np.zeros(extent, int)
        for o in ops:
            if o.kind == kind:
                assert 0 <= o.lo < o.hi <= extent
                seen[o.lo:o.hi] += 1
        assert (seen == 1).all()
    for o in ops:
        if o.kind == S:
            assert 0 <= o.lo < o.hi <= n_outer and o.steps >= 1 and o.src in (0, 1)
    assert {o.stream for o in ops if o.kind == U} == {0}
    assert {o.stream for o in ops if o.kind == D} == {1}
    assert all(o.stream >= 2 for o in ops if o.kind in (P, S))


def _range_pass(src, dst, n, r, shape, lo, hi, k):
    """What a fused range launch computes: dst[lo:hi) = k steps from src (outermost axis)."""
    n_outer = n[-1] if len(n) > 1 else n[0]
    a, b = max(lo - k * r, 0), min(hi + k * r, n_outer)
    sub = src[a:b + 2 * r].copy()
    sub2 = sub.copy()
    sub_n = tuple(n[:-1]) + (b - a,)
    res = sub2 if oracle.sweep(sub_n, r, shape, k, sub, sub2) else sub
    inner = (slice(None),) + tuple(slice(r, s - r) for s in src.shape[1:])
    dst[lo + r:hi + r][inner] = res[lo - a + r:hi - a + r][inner]


def _execute(ops, order, n, r, shape, dtype, T, seed=3):
    code = oracle.F64 if dtype == capi.F64 else oracle.F32
    host_in = oracle.fill(n, r, code, oracle.INIT_RANDOM, seed)
    host_out = oracle.fill(n, r, code, oracle.INIT_REF)
    want_a, want_b = host_in.copy(), host_out.copy()
    want = want_b if oracle.sweep(n, r, shape, T, want_a, want_b) else want_a
    in0, out0 = host_in.copy(), host_out.copy()
    staging = np.full_like(host_in, np.nan)
    g = [np.full_like(host_in, np.nan), np.full_like(host_in, np.nan)]
    host_dst = host_out if T & 1 else host_in
    inner = (slice(None),) + tuple(slice(r, s - r) for s in host_in.shape[1:])
    for j in order:
        o = ops[j]
        if o.kind == U:
            staging[o.lo:o.hi] = host_in[o.lo:o.hi]
        elif o.kind == P:
            g[0][o.lo:o.hi] = staging[o.lo:o.hi]
            g[1][o.lo:o.hi] = staging[o.lo:o.hi]
        elif o.kind == S:
            _range_pass(g[o.src], g[o.src ^ 1], n, r, shape, o.lo, o.hi, o.steps)
        else:
            host_dst[o.lo + r:o.hi + r][inner] = g[o.src][o.lo + r:o.hi + r][inner]
    got = host_dst
    assert not np.isnan(oracle.interior(got, r)).any(), "a launch consumed planes nobody had produced"
    assert np.array_equal(got.view(np.uint8), want.view(np.uint8))
    other, other0 = (host_in, in0) if T & 1 else (host_out, out0)
    assert np.array_equal(other.view(np.uint8), other0.view(np.uint8))  # untouched


def _random_order(ops, rng):
    preds = _preds(ops)
    indeg = [len(set(p)) for p in preds]
    succ = [[] for _ in ops]
    for j, p in enumerate(preds):
        for q in set(p):
            succ[q].append(j)
    ready = [j for j, d in enumerate(indeg) if d == 0]
    order = []
    while ready:
        j = ready.pop(rng.randrange(len(ready)))
        order.append(j)
        for s in succ[j]:
            indeg[s] -= 1
            if indeg[s] == 0:
                ready.append(s)
    assert len(order) == len(ops)
    return order


CASES = [
    # n (x,y[,z]), r, shape, dtype, T, C, NS
    ((12, 10, 64), 1, capi.STAR, capi.F64, 7, 5, 2),     # TMA3D k=2, remainder block first
    ((12, 10, 64), 1, capi.STAR, capi.F64, 8, 9, 2),
    ((12, 10, 37), 1, capi.STAR, capi.F64, 33, 3, 2),    # drift T*r > chunk width: early parallelograms download nothing
    ((9, 7, 80), 1, capi.BOX, capi.F32, 5, 6, 3),
    ((9, 7, 80), 1, capi.STAR, capi.F32, 1, 11, 4),      # four sweep streams
    ((9, 7, 80), 1, capi.STAR, capi.F32, 12, 4, 1),      # one sweep stream
    ((8, 6, 90), 2, capi.STAR, capi.F64, 6, 5, 2),       # generic 3-D radius 2 (k=1)
    ((40, 200), 1, capi.BOX, capi.F32, 9, 7, 2),         # WARP2D k=4: outermost axis is y
    ((40, 300), 2, capi.STAR, capi.F32, 10, 6, 2),       # WARP2D radius 2 (k=4)
    ((33, 150), 3, capi.BOX, capi.F64, 4, 5, 2),         # generic 2-D radius 3
]


@pytest.mark.parametrize("n,r,shape,dtype,T,C,NS", CASES)
def test_schedule_is_race_free_and_computes_the_sweep(built, monkeypatch, n, r, shape, dtype, T, C, NS):
    ops = _schedule(monkeypatch, n, r, shape, dtype, T, C, NS)
    assert ops, "the case is meant to be pipelined"
    n_outer = n[-1]
    _check_structure(ops, n_outer, r, T)
    assert _check_race_free(ops, n_outer, r, T) > 0
    _execute(ops, range(len(ops)), n, r, shape, dtype, T)
    rng = random.Random(1234)
    for _ in range(3):
        _execute(ops, _random_order(ops, rng), n, r, shape, dtype, T)


@pytest.mark.parametrize("T", [100, 99, 50])
def test_config2_schedule_is_race_free(built, monkeypatch, T):
    """BASELINE configs[1] exactly as sb200_run_host pipelines it (default C and streams)."""
    monkeypatch.delenv("SB200_PIPELINE_CHUNKS", raising=False)
    monkeypatch.delenv("SB200_PIPELINE_STREAMS", raising=False)
    d = capi.make_desc((512, 512, 512), radius=1, shape=capi.STAR, dtype=capi.F64, iterations=T)
    ops = capi.pipeline_schedule(d)
    assert capi.plan(d, T).pipeline_chunks == 11
    assert len({o.parallelogram for o in ops}) == 11
    _check_structure(ops, 512, 1, T)
    _check_race_free(ops, 512, 1, T)


@pytest.mark.parametrize("n,T", [((4096, 4096), 100), ((16384, 16384), 40), ((1024, 1024, 2048), 7)])
def test_other_large_grids_race_free(built, monkeypatch, n, T):
    monkeypatch.delenv("SB200_PIPELINE_CHUNKS", raising=False)
    monkeypatch.delenv("SB200_PIPELINE_STREAMS", raising=False)
    d = capi.make_desc(n, radius=1, shape=capi.STAR, dtype=capi.F64, iterations=T)
    ops = capi.pipeline_schedule(d)
    assert ops
    _check_structure(ops, n[-1], 1, T)
    _check_race_free(ops, n[-1], 1, T)


def test_small_grids_are_not_pipelined(built, monkeypatch):
    monkeypatch.delenv("SB200_PIPELINE_CHUNKS", raising=False)
    d = capi.make_desc((64, 64, 64), dtype=capi.F64, iterations=10)
    assert capi.pipeline_schedule(d) == []
    assert capi.pipeline_schedule(capi.make_desc((100000,), iterations=10)) == []   # 1-D never
    assert capi.pipeline_schedule(capi.make_desc((512, 512, 512), dtype=capi.F64, iterations=0)) == []


def test_checker_catches_a_missing_edge(built, monkeypatch):
    """The race check is not vacuous: dropping one cross-stream wait edge is detected."""
    ops = _schedule(monkeypatch, (12, 10, 64), 1, capi.STAR, capi.F64, 8, 5, 2)
    victim = next(j for j, o in enumerate(ops) if o.kind == S and o.parallelogram == 2 and o.wait[0] >= 0)
    ops[victim].wait[0] = -1
    with pytest.raises(AssertionError, match="conflict"):
        _check_race_free(ops, 64, 1, 8)


def test_random_shapes_race_free_and_correct(built, monkeypatch):
    """Seeded sweep over grid length, radius, steps, parallelogram and stream counts."""
    rng = random.Random(20261105)
    pipelined = 0
    for _ in range(80):
        ndim = rng.choice((2, 3))
        r = rng.choice((1, 1, 2, 3))
        n = ((rng.randrange(4, 12),) if ndim == 3 else ()) + (rng.randrange(2 * r, 14),)
        n = (rng.randrange(4, 40),) + n[1:] + (rng.randrange(8, 140),) if ndim == 3 else \
            (rng.randrange(4, 60), rng.randrange(8, 140))
        shape = rng.choice((capi.STAR, capi.BOX))
        dtype = rng.choice((capi.F32, capi.F64))
        T = rng.randrange(1, 41)
        ops = _schedule(monkeypatch, n, r, shape, dtype, T, rng.randrange(2, 13), rng.randrange(1, 5))
        if not ops:   # the grid is too short for its drift: plain upload -> sweep -> download
            continue
        pipelined += 1
        _check_structure(ops, n[-1], r, T)
        _check_race_free(ops, n[-1], r, T)
        _execute(ops, _random_order(ops, rng), n, r, shape, dtype, T, seed=rng.randrange(1, 1000))
    assert pipelined >= 30


# ---- one call, N GPUs: slab schedules (sb200_slab_plan_for / sb200_pipeline_schedule_slab) ----

def _slab_setup(monkeypatch, n, r, shape, dtype, T, C, NS, G):
    monkeypatch.setenv("SB200_PIPELINE_CHUNKS", str(C))
    monkeypatch.setenv("SB200_PIPELINE_STREAMS", str(NS))
    d = capi.make_desc(n, radius=r, shape=shape, dtype=dtype, iterations=T)
    plan = capi.slab_plan(d, G)
    slabs = []
    for g in range(plan.ngpus_used):
        ops, zones = capi.pipeline_schedule_slab(d, G, g)
        slabs.append((plan.slab[g], ops, zones))
    return plan, slabs


def _check_slab_structure(sl, ops, zones, r):
    sub_n = sl.sub_hi - sl.sub_lo
    padded = sub_n + 2 * r
    for kind in (U, P):
        seen = np.zeros(padded, int)
        for o in ops:
            if o.kind == kind:
                assert 0 <= o.lo < o.hi <= padded
                seen[o.lo:o.hi] += 1
        assert (seen == 1).all()
    seen = np.zeros(sub_n, int)
    for o in ops:
        if o.kind == D:
            seen[o.lo:o.hi] += 1
    want = np.zeros(sub_n, int)
    want[sl.own_lo - sl.sub_lo:sl.own_hi - sl.sub_lo] = 1
    assert (seen == want).all(), "a slab downloads exactly the planes it owns"
    ups = [o for o in ops if o.kind == U]
    assert [o for o in ops[:len(ups)] if o.kind != U] == [], "uploads lead the list (they are issued first)"
    # the leading `zones` uploads cover both deep-halo zones
    cov = np.zeros(padded, int)
    for o in ups[:zones]:
        cov[o.lo:o.hi] = 1
    lo_zone, hi_zone = sl.own_lo - sl.sub_lo, sl.own_hi - sl.sub_lo
    if lo_zone > 0:
        assert cov[:lo_zone + r].all()
    if hi_zone < sub_n:
        assert cov[hi_zone + r:].all()


def _execute_slabs(plan, slabs, n, r, shape, dtype, T, rng, seed=5):
    """All slabs on ONE pair of host arrays, ops interleaved in a random order that respects each
    slab's happens-before plus the cross-device rule (a slab's downloads start after its
    neighbours' zone uploads); the result must be the plain T-step sweep of the whole grid."""
    code = oracle.F64 if dtype == capi.F64 else oracle.F32
    host_in = oracle.fill(n, r, code, oracle.INIT_RANDOM, seed)
    host_out = oracle.fill(n, r, code, oracle.INIT_REF)
    want_a, want_b = host_in.copy(), host_out.copy()
    want = want_b if oracle.sweep(n, r, shape, T, want_a, want_b) else want_a
    host_dst = host_out if T & 1 else host_in
    inner = (slice(None),) + tuple(slice(r, s - r) for s in host_in.shape[1:])
    state = []
    nodes, preds = [], {}
    for g, (sl, ops, zones) in enumerate(slabs):
        sub_n = sl.sub_hi - sl.sub_lo
        shp = (sub_n + 2 * r,) + host_in.shape[1:]
        state.append(dict(staging=np.full(shp, np.nan, host_in.dtype),
                          g=[np.full(shp, np.nan, host_in.dtype), np.full(shp, np.nan, host_in.dtype)],
                          n=tuple(n[:-1]) + (sub_n,)))
        p = _preds(ops)
        ups = [j for j, o in enumerate(ops) if o.kind == U]
        for j in range(len(ops)):
            nodes.append((g, j))
            preds[(g, j)] = [(g, q) for q in set(p[j])]
    for g, (sl, ops, zones) in enumerate(slabs):
        for j, o in enumerate(ops):
            if o.kind == D:
                for nb in (g - 1, g + 1):
                    if 0 <= nb < len(slabs):
                        nb_ups = [q for q, oo in enumerate(slabs[nb][1]) if oo.kind == U]
                        if slabs[nb][2]:
                            preds[(g, j)].append((nb, nb_ups[slabs[nb][2] - 1]))
    indeg = {v: len(set(p)) for v, p in preds.items()}
    succ = {v: [] for v in nodes}
    for v, p in preds.items():
        for q in set(p):
            succ[q].append(v)
    ready = [v for v in nodes if indeg[v] == 0]
    done = 0
    while ready:
        g, j = ready.pop(rng.randrange(len(ready)))
        done += 1
        sl, ops, _ = slabs[g]
        o, st, base = ops[j], state[g], sl.sub_lo
        if o.kind == U:
            st["staging"][o.lo:o.hi] = host_in[base + o.lo:base + o.hi]
        elif o.kind == P:
            st["g"][0][o.lo:o.hi] = st["staging"][o.lo:o.hi]
            st["g"][1][o.lo:o.hi] = st["staging"][o.lo:o.hi]
        elif o.kind == S:
            _range_pass(st["g"][o.src], st["g"][o.src ^ 1], st["n"], r, shape, o.lo, o.hi, o.steps)
        else:
            host_dst[base + o.lo + r:base + o.hi + r][inner] = st["g"][o.src][o.lo + r:o.hi + r][inner]
        for s in succ[(g, j)]:
            indeg[s] -= 1
            if indeg[s] == 0:
                ready.append(s)
    assert done == len(nodes)
    assert not np.isnan(oracle.interior(host_dst, r)).any()
    assert np.array_equal(host_dst.view(np.uint8), want.view(np.uint8))


def _check_cross_slab_host_hazards(slabs, r, T):
    """T even: downloads write the array uploads read.  Every (download of slab a, upload of slab
    b) pair that touches the same host planes must be ordered by the zone rule."""
    if T & 1:
        return 0
    found = 0
    for a, (sla, opsa, _) in enumerate(slabs):
        for b, (slb, opsb, zb) in enumerate(slabs):
            if a == b:
                continue
            ups = [o for o in opsb if o.kind == U]
            for d in (o for o in opsa if o.kind == D):
                dlo, dhi = sla.sub_lo + d.lo + r, sla.sub_lo + d.hi + r
                for q, u in enumerate(ups):
                    if slb.sub_lo + u.lo < dhi and dlo < slb.sub_lo + u.hi:
                        found += 1
                        assert abs(a - b) == 1 and q < zb, "an unordered download/upload pair shares host planes"
    return found


SLAB_CASES = [
    # n, r, shape, dtype, T, C, NS, G
    ((12, 10, 96), 1, capi.STAR, capi.F64, 6, 4, 2, 3),     # T even: downloads land in the array others upload from
    ((12, 10, 96), 1, capi.STAR, capi.F64, 7, 5, 2, 2),
    ((9, 7, 128), 1, capi.BOX, capi.F32, 4, 3, 3, 4),
    ((40, 260), 1, capi.BOX, capi.F32, 8, 4, 2, 4),         # WARP2D k=4, y-slabs
    ((33, 200), 2, capi.STAR, capi.F64, 6, 5, 2, 2),        # radius 2: deep halo 12 rows
    ((8, 6, 120), 2, capi.STAR, capi.F64, 5, 4, 1, 2),      # generic 3-D radius 2
]


@pytest.mark.parametrize("n,r,shape,dtype,T,C,NS,G", SLAB_CASES)
def test_slab_schedules_race_free_and_exact(built, monkeypatch, n, r, shape, dtype, T, C, NS, G):
    plan, slabs = _slab_setup(monkeypatch, n, r, shape, dtype, T, C, NS, G)
    assert plan.ngpus_used == G and plan.deep_halo == T * r
    assert slabs[0][0].own_lo == 0 and slabs[-1][0].own_hi == n[-1]
    for g, (sl, ops, zones) in enumerate(slabs):
        assert sl.sub_lo == max(0, sl.own_lo - T * r) and sl.sub_hi == min(n[-1], sl.own_hi + T * r)
        if g:
            assert sl.own_lo == slabs[g - 1][0].own_hi
        _check_slab_structure(sl, ops, zones, r)
        _check_race_free(ops, sl.sub_hi - sl.sub_lo, r, T)
    if T % 2 == 0:
        assert _check_cross_slab_host_hazards(slabs, r, T) > 0
    rng = random.Random(99)
    for _ in range(3):
        _execute_slabs(plan, slabs, n, r, shape, dtype, T, rng)


def test_slab_plan_uses_fewer_gpus_when_slabs_get_thin(built, monkeypatch):
    monkeypatch.delenv("SB200_PIPELINE_CHUNKS", raising=False)
    big = capi.make_desc((1024, 1024, 2048), dtype=capi.F64, iterations=40)       # config 5
    p = capi.slab_plan(big, 8)
    assert p.ngpus_used == 8 and p.slab[3].own_hi - p.slab[3].own_lo == 256
    assert p.slab[3].sub_lo == p.slab[3].own_lo - 40 and p.slab[7].sub_hi == 2048
    c2 = capi.make_desc((512, 512, 512), dtype=capi.F64, iterations=100)          # config 2
    assert capi.slab_plan(c2, 8).ngpus_used == 2                                   # 256-plane slabs >= 2 * 100
    assert capi.slab_plan(c2, 1).ngpus_used == 1
    long_run = capi.make_desc((16384, 16384), shape=capi.BOX, iterations=1000)    # config 4
    assert capi.slab_plan(long_run, 8).ngpus_used == 8                             # 2048 rows >= 2 * 1000
    tiny = capi.make_desc((64, 64), iterations=1000)
    assert capi.slab_plan(tiny, 8).ngpus_used == 1
    assert capi.slab_plan(capi.make_desc((1 << 20,), iterations=4), 4).ngpus_used == 1   # 1-D never
