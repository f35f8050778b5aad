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

namespace sb200 {

constexpr int W2D_PX = 4;                 // x points per lane
constexpr int W2D_WIDTH = 32 * W2D_PX;    // strip width incl. halo
constexpr int W2D_WARPS = 4;              // warps (independent strips) per block
constexpr int W2D_DEPTH = 4;              // rows prefetched per lane
constexpr int W2D_MAXK = 8;

__host__ __device__ constexpr int w2d_halo_left(int K) { return (K + 3) & ~3; }
__host__ __device__ constexpr int w2d_valid(int K) { return (W2D_WIDTH - w2d_halo_left(K) - K) & ~3; }

template <class T> struct Row4 { T v[4]; };

template <class T>
__device__ __forceinline__ Row4<T> w2d_load(const T* __restrict__ p, bool ok) {
    Row4<T> r;
    if (!ok) {
        r.v[0] = r.v[1] = r.v[2] = r.v[3] = T(0);
        return r;
    }
    if constexpr (sizeof(T) == 4) {
        const float4 f = __ldg(reinterpret_cast<const float4*>(p));
        r.v[0] = f.x; r.v[1] = f.y; r.v[2] = f.z; r.v[3] = f.w;
    } else {
        const double2 a = __ldg(reinterpret_cast<const double2*>(p));
        const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
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
    __device__ __forceinline__ void step(const Row4<T>& c, T l, T rg, T avg, const bool* upd,
                                         Row4<T>& o) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const T xl = i == 0 ? l : c.v[i - 1];
            const T xr = i == 3 ? rg : c.v[i + 1];
            const T h = add_rn(add_rn(T(0), xl), xr);
            const T f = mul_rn(add_rn(t[i], c.v[i]), avg);
            o.v[i] = upd[i] ? f : cm1[i];
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
    __device__ __forceinline__ void step(const Row4<T>& c, T l, T rg, T avg, const bool* upd,
                                         Row4<T>& o) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const T xl = i == 0 ? l : c.v[i - 1];
            const T xr = i == 3 ? rg : c.v[i + 1];
            const T f = mul_rn(add_rn(add_rn(add_rn(b[i], xl), c.v[i]), xr), avg);
            o.v[i] = upd[i] ? f : cm1[i];
            b[i] = add_rn(add_rn(add_rn(a[i], xl), c.v[i]), xr);
            a[i] = add_rn(add_rn(add_rn(T(0), xl), c.v[i]), xr);
            cm1[i] = c.v[i];
        }
    }
};

template <class T>
__device__ __forceinline__ T shfl_up1(T v) { return __shfl_up_sync(0xffffffffu, v, 1); }
template <class T>
__device__ __forceinline__ T shfl_down1(T v) { return __shfl_down_sync(0xffffffffu, v, 1); }

// in -> out, K fused steps, output rows [ob, oe) (slab-interior coordinates).
template <class T, int SHAPE, int K>
__global__ void __launch_bounds__(32 * W2D_WARPS)
sweep_warp2d_kernel(const T* __restrict__ in, T* __restrict__ out, Grid g, Domain dom,
                    long long ob, long long oe, int nstrips, int chunk_rows, long long nitems,
                    T avg) {
    constexpr int HL = w2d_halo_left(K);
    constexpr int V = w2d_valid(K);
    const int lane = threadIdx.x & 31;
    const long long item = (long long)blockIdx.x * W2D_WARPS + (threadIdx.x >> 5);
    if (item >= nitems) return;
    const int strip = (int)(item % nstrips);
    const long long chunk = item / nstrips;
    const long long yb = ob + chunk * chunk_rows;
    const long long ye = (yb + chunk_rows < oe) ? yb + chunk_rows : oe;

    const long long gx = (long long)strip * V - HL + lane * W2D_PX;  // first x of this lane
    // loads stay inside the allocation (columns beyond it are zero)
    const bool col_ok = gx >= -(long long)g.ox && gx + W2D_PX <= g.pitch_x - g.ox;
    bool mx[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) mx[i] = gx + i >= 0 && gx + i < g.n[0];
    const bool lane_valid = lane * W2D_PX >= HL && lane * W2D_PX < HL + V && gx < g.n[0];
    const long long row_lo = -(long long)g.oy, row_hi = g.pitch_y - g.oy;  // allocated rows

    W2dLevel<T, SHAPE> lv[K];
#pragma unroll
    for (int l = 0; l < K; ++l) lv[l].init();

    const long long y_first = yb - K;                  // first row to arrive at level 1
    const int nsteps = (int)(ye - yb) + 2 * K;         // rows streamed through
    const T* __restrict__ src = in + g.idx(gx, 0, 0);
    T* __restrict__ dst = out + g.idx(gx, 0, 0);

    Row4<T> q[W2D_DEPTH];
#pragma unroll
    for (int d = 0; d < W2D_DEPTH; ++d) {
        const long long y = y_first + d;
        q[d] = w2d_load(src + y * g.pitch_x, col_ok && d < nsteps && y >= row_lo && y < row_hi);
    }

    for (int j0 = 0; j0 < nsteps; j0 += W2D_DEPTH) {
#pragma unroll
        for (int d = 0; d < W2D_DEPTH; ++d) {
            const int j = j0 + d;
            if (j < nsteps) {
                Row4<T> cur = q[d];
                {
                    const long long y = y_first + j + W2D_DEPTH;
                    q[d] = w2d_load(src + y * g.pitch_x, col_ok && j + W2D_DEPTH < nsteps &&
                                                             y >= row_lo && y < row_hi);
                }
#pragma unroll
                for (int l = 0; l < K; ++l) {
                    // level l+1: row (y_first + j - l) arrives, row (y_first + j - l - 1) is emitted
                    const long long yo = y_first + j - l - 1;
                    const bool row_upd = yo >= dom.lo && yo < dom.hi;
                    bool upd[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) upd[i] = mx[i] && row_upd;
                    const T lft = shfl_up1(cur.v[3]);
                    const T rgt = shfl_down1(cur.v[0]);
                    Row4<T> o;
                    lv[l].step(cur, lft, rgt, avg, upd, o);
                    cur = o;
                }
                const long long yo = y_first + j - K;  // row emitted by the last level
                if (lane_valid && yo >= yb && yo < ye) {
                    T* p = dst + yo * g.pitch_x;
                    if (gx + W2D_PX <= g.n[0]) {
                        w2d_store(p, cur);
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (gx + i < g.n[0]) p[i] = cur.v[i];
                    }
                }
            }
        }
    }
}

struct W2dPlan {
    int nstrips;
    int chunk_rows;
    long long nchunks, nitems;
};

static inline W2dPlan w2d_plan(long long nx, long long rows, int K, int sm_count) {
    W2dPlan p;
    const int V = (W2D_WIDTH - ((K + 3) & ~3) - K) & ~3;
    p.nstrips = (int)((nx + V - 1) / V);
    // enough warps to fill the machine (~16 resident warps per SM, 2 rounds), but chunks
    // of at least 16*K rows so the 2K redundant rows per chunk stay a small fraction
    const long long want_items = (long long)sm_count * 16 * 2;
    long long nchunks = (want_items + p.nstrips - 1) / p.nstrips;
    long long min_rows = 16LL * K;
    long long chunk_rows = (rows + nchunks - 1) / nchunks;
    if (chunk_rows < min_rows) chunk_rows = min_rows;
    if (chunk_rows > rows) chunk_rows = rows;
    if (chunk_rows < 1) chunk_rows = 1;
    p.chunk_rows = (int)chunk_rows;
    p.nchunks = (rows + chunk_rows - 1) / chunk_rows;
    p.nitems = p.nchunks * p.nstrips;
    return p;
}

template <class T, int SHAPE, int K>
static cudaError_t launch_warp2d_k(const T* in, T* out, const Grid& g, Domain dom, long long ob,
                                   long long oe, int sm_count, cudaStream_t st) {
    if (oe <= ob) return cudaSuccess;
    const W2dPlan p = w2d_plan(g.n[0], oe - ob, K, sm_count);
    const T avg = Avg<T>::value(SHAPE, 2, 1);
    const unsigned blocks = (unsigned)((p.nitems + W2D_WARPS - 1) / W2D_WARPS);
    sweep_warp2d_kernel<T, SHAPE, K><<<blocks, 32 * W2D_WARPS, 0, st>>>(
        in, out, g, dom, ob, oe, p.nstrips, p.chunk_rows, p.nitems, avg);
    return cudaGetLastError();
}

template <class T, int SHAPE>
static cudaError_t launch_warp2d_s(const T* in, T* out, const Grid& g, Domain dom, int K,
                                   long long ob, long long oe, int sm_count, cudaStream_t st) {
    switch (K) {
#define SB200_W2D(KK) case KK: return launch_warp2d_k<T, SHAPE, KK>(in, out, g, dom, ob, oe, sm_count, st)
        SB200_W2D(1); SB200_W2D(2); SB200_W2D(3); SB200_W2D(4);
        SB200_W2D(5); SB200_W2D(6); SB200_W2D(7); SB200_W2D(8);
#undef SB200_W2D
        default: return cudaErrorInvalidValue;
    }
}

template <class T>
static cudaError_t launch_warp2d(const T* in, T* out, const Grid& g, int shape, Domain dom, int K,
                                 long long ob, long long oe, int sm_count, cudaStream_t st) {
    return shape == STAR ? launch_warp2d_s<T, STAR>(in, out, g, dom, K, ob, oe, sm_count, st)
                         : launch_warp2d_s<T, BOX>(in, out, g, dom, K, ob, oe, sm_count, st);
}

}  // namespace sb200
