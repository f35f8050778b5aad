// warp2d.cuh — 2-D radius-1 star / box sweep with K time steps fused per HBM round trip.
//
// B200 design (nothing like the reference's Sunway tiles, but the same idea as its
// "tile stays on-chip for all T steps", stencil_dma.cpp:418,567): a WARP owns a strip
// of 128 consecutive x and streams along y (the slowest axis of a 2-D grid).  Each lane
// holds 4 consecutive x of the current row in registers (one 128-bit load for f32, two
// for f64); x-neighbours come from the adjacent lanes by warp shuffle; the rows y-1 /
// y+1 needed by each of the K fused time levels live in per-lane registers as running
// partial sums.  No shared memory, no block barrier: a level's output row is the next
// level's arriving row in the same iteration.  The strip carries a halo of K columns
// per side which is recomputed redundantly (trapezoid), and a chunk of rows carries K
// extra rows per side.
//
// Summation order per point is the oracle's (reference stencil.cpp:103-125):
//   star: ((((0 + xl) + xr) + yl) + yr) * avg      box: rows y-1,y,y+1 each x-1,x,x+1
// so results are bit-identical for any K.  Points outside the update domain (Dirichlet
// halo: x<0, x>=nx, rows outside [dom.lo, dom.hi)) pass through unchanged at every level.
#pragma once
#include "common.cuh"
#include "slablink.cuh"

namespace sb200 {

constexpr int W2D_PX = 4;                 // x points per lane
constexpr int W2D_WIDTH = 32 * W2D_PX;    // strip width incl. halo
constexpr int W2D_WARPS = 4;              // warps (independent strips) per block
constexpr int W2D_DEPTH = 8;              // ring slots per warp (DEPTH-1 rows in flight)
constexpr int W2D_MAXK = 6;  // the curve turns at 6 (fp64 star) / 4 (others); 7 and 8 only spilled at 255 registers (profiles/r01c_probe.log)

__host__ __device__ constexpr int w2d_halo_left(int K) { return (K + 3) & ~3; }
__host__ __device__ constexpr int w2d_valid(int K) { return (W2D_WIDTH - w2d_halo_left(K) - K) & ~3; }

template <class T> struct Row4 { T v[4]; };

// cp.async (LDGSTS) staging: each lane copies the 4 x it owns (16 B for f32, 2 x 16 B for
// f64) from global straight into a warp-private shared-memory ring, bypassing registers and
// L1; the lane later reads back exactly the bytes it copied, so cp.async.wait_group alone
// orders the accesses (no barrier).
template <class T>
__device__ __forceinline__ void w2d_async_row(T* sdst, const T* __restrict__ gsrc) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(sdst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gsrc) : "memory");
    if constexpr (sizeof(T) == 8)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa + 16), "l"(gsrc + 2) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void w2d_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
template <class T>
__device__ __forceinline__ Row4<T> w2d_read_row(const T* sp) {
    Row4<T> r;
    if constexpr (sizeof(T) == 4) {
        const float4 f = *reinterpret_cast<const float4*>(sp);
        r.v[0] = f.x; r.v[1] = f.y; r.v[2] = f.z; r.v[3] = f.w;
    } else {
        const double2 a = reinterpret_cast<const double2*>(sp)[0];
        const double2 b = reinterpret_cast<const double2*>(sp)[1];
        r.v[0] = a.x; r.v[1] = a.y; r.v[2] = b.x; r.v[3] = b.y;
    }
    return r;
}

template <class T>
__device__ __forceinline__ void w2d_store(T* __restrict__ p, const Row4<T>& r) {
    if constexpr (sizeof(T) == 4) {
        *reinterpret_cast<float4*>(p) = make_float4(r.v[0], r.v[1], r.v[2], r.v[3]);
    } else {
        reinterpret_cast<double2*>(p)[0] = make_double2(r.v[0], r.v[1]);
        reinterpret_cast<double2*>(p)[1] = make_double2(r.v[2], r.v[3]);
    }
}

// Per-lane state of one time level.
template <class T, int SHAPE> struct W2dLevel;

template <class T> struct W2dLevel<T, STAR> {
    T t[4];    // ((0+xl)+xr)+yl of the previous row, waiting for its yr
    T cm1[4];  // previous row (centre values; also the pass-through value)
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int i = 0; i < 4; ++i) t[i] = cm1[i] = T(0);
    }
    // Row `c` (with left/right neighbour values l, rg) arrives; emits the finished
    // previous row into `o`.  `upd[i]`: point i of the emitted row is updatable.
    template <bool EDGE>
    __device__ __forceinline__ void step(const Row4<T>& c, T l, T rg, T avg, unsigned upd,
                                         Row4<T>& o) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const T xl = i == 0 ? l : c.v[i - 1];
            const T xr = i == 3 ? rg : c.v[i + 1];
            const T h = add_rn(add_rn(T(0), xl), xr);
            const T f = mul_rn(add_rn(t[i], c.v[i]), avg);
            o.v[i] = (!EDGE || ((upd >> i) & 1u)) ? f : cm1[i];
            t[i] = add_rn(h, cm1[i]);
            cm1[i] = c.v[i];
        }
    }
};

template <class T> struct W2dLevel<T, BOX> {
    T a[4];    // row-sum chain of the output row whose FIRST row (y-1) has been added
    T b[4];    // chain of the output row whose first TWO rows have been added
    T cm1[4];  // previous row, pass-through value
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = b[i] = cm1[i] = T(0);
    }
    template <bool EDGE>
    __device__ __forceinline__ void step(const Row4<T>& c, T l, T rg, T avg, unsigned upd,
                                         Row4<T>& o) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const T xl = i == 0 ? l : c.v[i - 1];
            const T xr = i == 3 ? rg : c.v[i + 1];
            const T f = mul_rn(add_rn(add_rn(add_rn(b[i], xl), c.v[i]), xr), avg);
            o.v[i] = (!EDGE || ((upd >> i) & 1u)) ? f : cm1[i];
            b[i] = add_rn(add_rn(add_rn(a[i], xl), c.v[i]), xr);
            a[i] = add_rn(add_rn(add_rn(T(0), xl), c.v[i]), xr);
            if (EDGE) cm1[i] = c.v[i];  // only the pass-through needs it
        }
    }
};

// 9-point box with separable sums (SB200_FLAG_REASSOCIATE): row sum rs = (xl + c) + xr, output row
// y-1 = ((rs(y-2) + rs(y-1)) + rs(y)) * avg — 4 adds + 1 multiply per point and level instead of
// 9 + 1, in an order the reference does not use: equal within 1e-5 (fp32) / 1e-12 (fp64) relative.
template <class T> struct W2dLevel<T, BOXSEP> {
    T p2[4];   // row sum of row y-2
    T p1[4];   // row sum of row y-1
    T cm1[4];  // previous row, pass-through value
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int i = 0; i < 4; ++i) p2[i] = p1[i] = cm1[i] = T(0);
    }
    template <bool EDGE>
    __device__ __forceinline__ void step(const Row4<T>& c, T l, T rg, T avg, unsigned upd,
                                         Row4<T>& o) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const T xl = i == 0 ? l : c.v[i - 1];
            const T xr = i == 3 ? rg : c.v[i + 1];
            const T rs = add_rn(add_rn(xl, c.v[i]), xr);
            const T f = mul_rn(add_rn(add_rn(p2[i], p1[i]), rs), avg);
            o.v[i] = (!EDGE || ((upd >> i) & 1u)) ? f : cm1[i];
            p2[i] = p1[i];
            p1[i] = rs;
            if (EDGE) cm1[i] = c.v[i];
        }
    }
};

template <class T>
__device__ __forceinline__ T shfl_up1(T v) { return __shfl_up_sync(0xffffffffu, v, 1); }
template <class T>
__device__ __forceinline__ T shfl_down1(T v) { return __shfl_down_sync(0xffffffffu, v, 1); }

// Stream one (strip, chunk) item through the K fused levels.  EDGE = false is the
// select-free variant for items that lie entirely inside the update domain (every streamed
// row and every column of the strip, halo columns included, is an updatable point), which
// is the case for all but the outermost strips / chunks.
template <class T, int SHAPE, int K, bool EDGE>
__device__ __forceinline__ void w2d_stream(const T* __restrict__ src, T* __restrict__ dst,
                                           long long pitch, int nsteps, int j_dom_lo, int j_dom_hi,
                                           int j_row_lo, int j_row_hi, unsigned mxmask, T* ring,
                                           bool lane_valid, int ncols_valid, T avg) {
    W2dLevel<T, SHAPE> lv[K];
#pragma unroll
    for (int l = 0; l < K; ++l) lv[l].init();
    // Row index j counts streamed rows: global row = y_first + j; src/dst point at row
    // y_first.  Loads are UNCONDITIONAL (no predicated zero-fill that would serialise the
    // register prefetch queue): rows past the end of the item or outside the allocation are
    // clamped to a valid row — whatever they return lies outside the dependence cone of
    // every valid output (beyond the Dirichlet halo, which passes through).
    const int j_ld_lo = EDGE ? (j_row_lo > 0 ? j_row_lo : 0) : 0;
    const int j_ld_hi = (EDGE && j_row_hi < nsteps ? j_row_hi : nsteps) - 1;
    auto row_ptr = [&](int jn) -> const T* {
        int jc = jn < j_ld_hi ? jn : j_ld_hi;
        if (EDGE) jc = jc > j_ld_lo ? jc : j_ld_lo;
        return src + (long long)jc * pitch;
    };
    // prologue: rows 0 .. DEPTH-2 in flight
#pragma unroll
    for (int d = 0; d < W2D_DEPTH - 1; ++d) w2d_async_row(ring + d * W2D_WIDTH, row_ptr(d));
    T* __restrict__ sp = dst - (long long)K * pitch;  // row emitted by the last level at j
    for (int j0 = 0; j0 < nsteps; j0 += W2D_DEPTH) {
#pragma unroll
        for (int d = 0; d < W2D_DEPTH; ++d) {
            // steps past nsteps (loop rounded up to the ring depth) compute on clamped
            // rows and store nothing
            const int j = j0 + d;
            {
                // refill the slot that was consumed in the previous step, then wait until
                // all but the newest DEPTH-1 copies (i.e. row j) have landed
                w2d_async_row(ring + ((d + W2D_DEPTH - 1) % W2D_DEPTH) * W2D_WIDTH, row_ptr(j + W2D_DEPTH - 1));
                w2d_async_wait<W2D_DEPTH - 1>();
                Row4<T> cur = w2d_read_row(ring + d * W2D_WIDTH);
#pragma unroll
                for (int l = 0; l < K; ++l) {
                    // level l+1: row j - l arrives, row j - l - 1 is emitted
                    unsigned upd = 0;
                    if (EDGE) {
                        const int jo = j - l - 1;
                        upd = (jo >= j_dom_lo && jo < j_dom_hi) ? mxmask : 0u;
                    }
                    const T lft = shfl_up1(cur.v[3]);
                    const T rgt = shfl_down1(cur.v[0]);
                    Row4<T> o;
                    lv[l].template step<EDGE>(cur, lft, rgt, avg, upd, o);
                    cur = o;
                }
                if (lane_valid && j >= 2 * K && j < nsteps) {  // emitted row j - K is an output row
                    if (ncols_valid >= W2D_PX) {
                        w2d_store(sp, cur);
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (i < ncols_valid) sp[i] = cur.v[i];
                    }
                }
            }
            sp += pitch;
        }
    }
}

// ---- slab link at warp granularity (shared with warp2dr.cuh) ------------------------------------
// A face-owning warp waits until the neighbour's previous launch has delivered the ghost rows it
// is about to read: lane 0 acquires the arrival counter, __syncwarp orders the other lanes'
// loads behind it (cp.async reads through the generic proxy: no proxy fence is needed).
__device__ __forceinline__ void w2d_link_acquire(const SlabLinkDev& link, bool is_lo, bool is_hi, int lane) {
    if ((is_lo || is_hi) && link.wait_value != 0) {
        if (lane == 0) {
            if (is_lo) link_wait(link.wait_lo, link.wait_value, link.err_flag);
            if (is_hi) link_wait(link.wait_hi, link.wait_value, link.err_flag);
        }
        __syncwarp();
    }
}
// After its stream a face-owning warp copies the face-row points its lanes stored (own writes:
// no barrier before the read-back) into the ghost rows of the neighbour's destination buffer
// with 128-bit stores over NVLink and counts itself in with one release-add.
template <class T>
__device__ __forceinline__ void w2d_link_push(T* __restrict__ out, const Grid& g, long long gx, bool lane_valid,
                                              int ncols_valid, const SlabLinkDev& link, bool is_lo, bool is_hi,
                                              int lane) {
    if (!(is_lo || is_hi)) return;
    T* home = out + g.idx(gx, 0, 0);
#pragma unroll 1
    for (int side = 0; side < 2; ++side) {
        if (side == 0 ? !is_lo : !is_hi) continue;
        const long long y0 = side == 0 ? 0 : g.n[1] - link.face_planes;
        const long long rdelta = (static_cast<T*>(side == 0 ? link.peer_out_lo : link.peer_out_hi) - out) +
                                 (side == 0 ? link.shift_lo : link.shift_hi);
        if (lane_valid) {
#pragma unroll 1
            for (int p = 0; p < link.face_planes; ++p) {
                T* srow = home + (y0 + p) * g.pitch_x;
                if (ncols_valid >= W2D_PX) {
                    w2d_store(srow + rdelta, w2d_read_row(srow));
                } else {
                    for (int i = 0; i < ncols_valid; ++i) srow[rdelta + i] = srow[i];
                }
            }
        }
    }
    __syncwarp();
    if (lane == 0) {
        __threadfence_system();
        if (is_lo) red_release_sys_add_u32(link.signal_lo, 1u);
        if (is_hi) red_release_sys_add_u32(link.signal_hi, 1u);
    }
}

// in -> out, K fused steps, output rows [ob, oe) (slab-interior coordinates).
// LINKED (slablink.cuh): the launch covers a whole slab; the WARPS of its first / last row chunk
// also own the low / high face (g rows each).  They are dispatched first (last chunk, then
// first), acquire the neighbours' arrival counter before their first load — cp.async reads
// through the generic proxy, so the acquire + __syncwarp is all the ordering needed — and after
// their stream copy the face rows they stored into the ghost rows of the neighbour's destination
// buffer with 128-bit stores over NVLink, then count themselves in with one release-add per warp.
// The row loop carries no link code.
template <class T, int SHAPE, int K, bool LINKED = false>
__global__ void __launch_bounds__(32 * W2D_WARPS)
sweep_warp2d_kernel(const T* __restrict__ in, T* __restrict__ out, Grid g, Domain dom,
                    long long ob, long long oe, int nstrips, int chunk_rows, long long nitems,
                    T avg, const SlabLinkDev link) {
    constexpr int HL = w2d_halo_left(K);
    constexpr int V = w2d_valid(K);
    const int lane = threadIdx.x & 31;
    const long long item = (long long)blockIdx.x * W2D_WARPS + (threadIdx.x >> 5);
    if (item >= nitems) return;
    const int strip = (int)(item % nstrips);
    long long chunk = item / nstrips;
    if (LINKED && link.link_hi) chunk = chunk == 0 ? link.nchunks - 1 : chunk - 1;  // the last chunk first
    const bool is_lo = LINKED && link.link_lo && chunk == 0;
    const bool is_hi = LINKED && link.link_hi && chunk == link.nchunks - 1;
    // linked launches balance the chunks: every chunk holds at least g rows, so the first / last
    // chunk contains the whole low / high face
    const long long yb = LINKED ? ob + (oe - ob) * chunk / link.nchunks : ob + chunk * chunk_rows;
    const long long ye = LINKED ? ob + (oe - ob) * (chunk + 1) / link.nchunks
                                : ((yb + chunk_rows < oe) ? yb + chunk_rows : oe);
    if (LINKED) w2d_link_acquire(link, is_lo, is_hi, lane);

    const long long gx = (long long)strip * V - HL + lane * W2D_PX;  // first x of this lane
    // loads stay inside the allocation: lanes beyond it re-read the nearest valid column
    // group (their values never reach a valid output)
    long long gx_ld = gx < -(long long)g.ox ? -(long long)g.ox : gx;
    if (gx_ld + W2D_PX > g.pitch_x - g.ox) gx_ld = g.pitch_x - g.ox - W2D_PX;
    unsigned mxmask = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (gx + i >= 0 && gx + i < g.n[0]) mxmask |= 1u << i;
    const bool lane_valid = lane * W2D_PX >= HL && lane * W2D_PX < HL + V && gx < g.n[0];
    const long long rem = g.n[0] - gx;
    const int ncols_valid = rem >= W2D_PX ? W2D_PX : (rem > 0 ? (int)rem : 0);

    const long long y_first = yb - K;           // first row to arrive at level 1
    const int nsteps = (int)(ye - yb) + 2 * K;  // rows streamed through
    // bounds as streamed-row indices j (global row = y_first + j), clamped to int range
    auto rel = [&](long long y) -> int {
        const long long v = y - y_first;
        return v < -8 ? -8 : (v > nsteps + 8 ? nsteps + 8 : (int)v);
    };
    const int j_dom_lo = rel(dom.lo), j_dom_hi = rel(dom.hi);
    const int j_row_lo = rel(-(long long)g.oy), j_row_hi = rel(g.pitch_y - g.oy);  // allocated rows
    const T* __restrict__ src = in + g.idx(gx_ld, y_first, 0);
    T* __restrict__ dst = out + g.idx(gx, y_first, 0);

    __shared__ __align__(16) T ring_all[W2D_WARPS][W2D_DEPTH][W2D_WIDTH];
    T* ring = &ring_all[threadIdx.x >> 5][0][lane * W2D_PX];
    const bool interior = __all_sync(0xffffffffu, mxmask == 0xFu) && j_dom_lo <= 0 && j_dom_hi >= nsteps;
    if (interior) {
        w2d_stream<T, SHAPE, K, false>(src, dst, g.pitch_x, nsteps, j_dom_lo, j_dom_hi, j_row_lo, j_row_hi,
                                       mxmask, ring, lane_valid, ncols_valid, avg);
    } else {
        w2d_stream<T, SHAPE, K, true>(src, dst, g.pitch_x, nsteps, j_dom_lo, j_dom_hi, j_row_lo, j_row_hi,
                                      mxmask, ring, lane_valid, ncols_valid, avg);
    }
    if (LINKED) w2d_link_push<T>(out, g, gx, lane_valid, ncols_valid, link, is_lo, is_hi, lane);
}

struct W2dPlan {
    int nstrips;
    int chunk_rows;
    long long nchunks, nitems;
};

// Rows per chunk.  Large grids: ~2 rounds of 16 warps per SM with chunks of at least 16K
// rows, so the 2K redundant rows per chunk stay a small fraction (HBM-bound regime).  Small
// grids — the reference's own N <= 960 regime (run_expr.py:6-14) — cannot fill the machine
// that way and a launch costs the latency of ONE warp walking its chunk row by row, so they
// are cut into short chunks (>= 4 rows) until every SM sub-partition has a few warps.
// R > 1: the radius-R kernel of warp2dr.cuh, whose halo is K*R per side.
// resident_warps (0 = unknown): warps of this kernel one SM holds at once.  A grid that needs between
// one and two rounds of them (config 1: 2240 items for 1776 places) runs a nearly empty second round;
// its chunks are lengthened until the items fit ONE round (config 1 fp64: +10 % at k=4, +16 % at k=6,
// profiles/r02t_c1_tile_probe.log).  Grids of many rounds keep their short chunks (config 4).
static inline W2dPlan w2d_plan(long long nx, long long rows, int K, int sm_count, int R = 1, int prefer_rows = 0,
                               int resident_warps = 0) {
    W2dPlan p;
    const int V = (W2D_WIDTH - ((K * R + 3) & ~3) - K * R) & ~3;
    p.nstrips = (int)((nx + V - 1) / V);
    const long long want_items = (long long)sm_count * 16 * 2;
    long long nchunks = (want_items + p.nstrips - 1) / p.nstrips;
    long long chunk_rows = (rows + nchunks - 1) / nchunks;
    const long long min_rows = 16LL * K * R;
    // Large grids: more, shorter chunks balance the SMs better than few long ones, up to the point
    // where the 2KR redundant rows per chunk cost more than the balance gains — measured on config 4
    // (16384^2, profiles/r02t_c4_tile_probe.log): 150-300 rows beat the 469 of the rule above by 8-14 %.
    const long long max_rows = 40LL * K * R;  // 5 % redundant rows
    if (chunk_rows > max_rows) chunk_rows = max_rows;
    const bool small = ((rows + min_rows - 1) / min_rows) * p.nstrips < (long long)sm_count * 8;
    if (chunk_rows < min_rows && !small) chunk_rows = min_rows;
    if (chunk_rows < min_rows && small) {
        // latency-bound regime: warps per sub-partition that still overlap (fewer for deep K,
        // whose row step is issue-heavy)
        const int per_smsp = K <= 2 ? 4 : (K <= 4 ? 2 : 1);
        const long long target = (long long)sm_count * 4 * per_smsp;
        nchunks = (target + p.nstrips - 1) / p.nstrips;
        chunk_rows = (rows + nchunks - 1) / nchunks;
        if (chunk_rows < 4) chunk_rows = 4;
    }
    if (resident_warps > 0 && !small) {
        const long long places = (long long)sm_count * resident_warps;
        const long long items = ((rows + chunk_rows - 1) / chunk_rows) * p.nstrips;
        const long long fit = places / p.nstrips;  // chunks per strip that fill one round
        if (items > places && items < 2 * places && fit >= 1) chunk_rows = (rows + fit - 1) / fit;
    }
    if (prefer_rows > 0) chunk_rows = prefer_rows;  // the caller's tile preference wins
    if (chunk_rows > rows) chunk_rows = rows;
    if (chunk_rows < 1) chunk_rows = 1;
    p.chunk_rows = (int)chunk_rows;
    p.nchunks = (rows + chunk_rows - 1) / chunk_rows;
    p.nitems = p.nchunks * p.nstrips;
    return p;
}

// Warps of sweep_warp2d_kernel<T, SHAPE, K> one SM holds (occupancy API; 0 without a device).
template <class T, int SHAPE, int K>
static int w2d_resident_warps() {
    static const int resident = [] {
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, sweep_warp2d_kernel<T, SHAPE, K, false>, 32 * W2D_WARPS, 0) !=
            cudaSuccess) {
            (void)cudaGetLastError();
            return 0;
        }
        return nb * W2D_WARPS;
    }();
    return resident;
}

template <class T, int SHAPE, int K>
static cudaError_t launch_warp2d_k(const T* in, T* out, const Grid& g, Domain dom, long long ob,
                                   long long oe, const Tuning& tn, cudaStream_t st, const SlabLinkDev* link_in) {
    if (oe <= ob) return cudaSuccess;
    const W2dPlan p = w2d_plan(g.n[0], oe - ob, K, tn.sm_count, 1, tn.prefer_chunk, w2d_resident_warps<T, SHAPE, K>());
    const T avg = Avg<T>::value(SHAPE == BOXSEP ? (int)BOX : SHAPE, 2, 1);
    if (link_in && (link_in->link_lo || link_in->link_hi)) {
        SlabLinkDev link = *link_in;
        long long nc = p.nchunks;
        if (nc > (oe - ob) / link.face_planes) nc = (oe - ob) / link.face_planes;
        if (nc < 1) nc = 1;
        link.nchunks = (int)nc;
        const long long nitems = nc * p.nstrips;
        const unsigned blocks = (unsigned)((nitems + W2D_WARPS - 1) / W2D_WARPS);
        sweep_warp2d_kernel<T, SHAPE, K, true><<<blocks, 32 * W2D_WARPS, 0, st>>>(
            in, out, g, dom, ob, oe, p.nstrips, p.chunk_rows, nitems, avg, link);
        return cudaGetLastError();
    }
    const unsigned blocks = (unsigned)((p.nitems + W2D_WARPS - 1) / W2D_WARPS);
    sweep_warp2d_kernel<T, SHAPE, K, false><<<blocks, 32 * W2D_WARPS, 0, st>>>(
        in, out, g, dom, ob, oe, p.nstrips, p.chunk_rows, p.nitems, avg, SlabLinkDev{});
    return cudaGetLastError();
}

template <class T, int SHAPE>
static cudaError_t launch_warp2d_s(const T* in, T* out, const Grid& g, Domain dom, int K,
                                   long long ob, long long oe, const Tuning& tn, cudaStream_t st,
                                   const SlabLinkDev* link) {
    switch (K) {
#define SB200_W2D(KK) case KK: return launch_warp2d_k<T, SHAPE, KK>(in, out, g, dom, ob, oe, tn, st, link)
        SB200_W2D(1); SB200_W2D(2); SB200_W2D(3); SB200_W2D(4);
        SB200_W2D(5); SB200_W2D(6);
#undef SB200_W2D
        default: return cudaErrorInvalidValue;
    }
}

template <class T>
static cudaError_t launch_warp2d(const T* in, T* out, const Grid& g, int shape, Domain dom, int K,
                                 long long ob, long long oe, const Tuning& tn, cudaStream_t st,
                                 const SlabLinkDev* link = nullptr) {
    return shape == STAR  ? launch_warp2d_s<T, STAR>(in, out, g, dom, K, ob, oe, tn, st, link)
           : shape == BOX ? launch_warp2d_s<T, BOX>(in, out, g, dom, K, ob, oe, tn, st, link)
                          : launch_warp2d_s<T, BOXSEP>(in, out, g, dom, K, ob, oe, tn, st, link);
}

}  // namespace sb200
