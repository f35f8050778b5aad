// warp2dr.cuh — 2-D star stencils of radius R = 2..4 with K time steps fused per HBM round
// trip: the general-radius sibling of warp2d.cuh (what the reference's `-r` flag selects,
// program_options.cpp:22-23; its general-radius tile kernel is neighbors_impl,
// stencil_dma.cpp:571-1099, and its compile-time-radius one StencilImpl<T,BW,BH>,
// stencil_dma_static_unroll.cpp:502).
//
// Same design: a warp owns a 128-wide x strip and streams along y, 4 x per lane in
// registers, rows staged by cp.async into a warp-private ring, no block barrier.  What
// radius R adds: the x window takes R values from each neighbouring lane (2R shuffles per
// row and level), and a level keeps the last R input rows plus R pending partial sums —
// output row y is finished when row y+R arrives, so every level lags R rows and the
// trapezoid halo is K*R columns / rows per side.  The R-deep rings are indexed by the row's
// phase (row mod R), a compile-time constant because the row loop is unrolled by a multiple
// of R: no register rotation.
//
// Per-point order is the oracle's (reference stencil.cpp:103-125): x low side far->near,
// x high side near->far, y low side far->near, y high side near->far (one add per arriving
// row), one multiply by avg = 1/(4R): bit-identical for any K.
#pragma once
#include "warp2d.cuh"

namespace sb200 {

// ring slots: a multiple of the level's phase count (star: R, box: 2R)
__host__ __device__ constexpr int w2r_depth(int R, int shape = STAR) { return R == 3 ? (shape == BOX ? 6 : 9) : 8; }
__host__ __device__ constexpr int w2r_halo_left(int K, int R) { return (K * R + 3) & ~3; }
__host__ __device__ constexpr int w2r_valid(int K, int R) { return (W2D_WIDTH - w2r_halo_left(K, R) - K * R) & ~3; }
// deepest fusion per radius and element size (register budget: 8R state registers per level
// for fp32, 16R for fp64)
__host__ __device__ constexpr int w2r_max_k(int R, int elem_size) {
    return elem_size == 4 ? (R == 2 ? 4 : 2) : (R == 2 ? 2 : (R == 3 ? 2 : 1));
}

template <class T, int R> struct W2rLevel {
    T prev[R][4];  // slot s: the newest input row whose index is congruent to s (mod R)
    T pend[R][4];  // its partial sum: x part + rows above it + the rows below that have arrived
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int s = 0; s < R; ++s)
#pragma unroll
            for (int i = 0; i < 4; ++i) prev[s][i] = pend[s][i] = T(0);
    }
    // Input row y (phase P = y mod R) arrives with its x window w[0 .. 4+2R) (w[R+i] = own
    // point i); emits the finished row y-R into o.  upd: bit i = point i of the emitted row is
    // updatable (else it passes its value through).
    template <int P, bool EDGE>
    __device__ __forceinline__ void step(const T (&w)[4 + 2 * R], T avg, unsigned upd, Row4<T>& o) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            T xs = T(0);
#pragma unroll
            for (int k = R; k >= 1; --k) xs = add_rn(xs, w[R + i - k]);
#pragma unroll
            for (int k = 1; k <= R; ++k) xs = add_rn(xs, w[R + i + k]);
            // rows y-R .. y-1 (far -> near) complete the low side of row y
            T tn = xs;
#pragma unroll
            for (int k = R; k >= 1; --k) tn = add_rn(tn, prev[(P + R - k) % R][i]);
            // row y is the +k neighbour of row y-k
#pragma unroll
            for (int k = 1; k <= R; ++k) pend[(P + R - k) % R][i] = add_rn(pend[(P + R - k) % R][i], w[R + i]);
            // row y-R lives in slot P (about to be replaced by row y) and is complete now
            const T f = mul_rn(pend[P][i], avg);
            o.v[i] = (!EDGE || ((upd >> i) & 1u)) ? f : prev[P][i];
            pend[P][i] = tn;
            prev[P][i] = w[R + i];
        }
    }
};

// ---- box of radius R (25-point for R = 2) ------------------------------------------------------
// Oracle order: one chain from 0 over dy ascending (outer) and dx ascending
// (inner), then one multiply.  An arriving row y is row dy = y - y' of the 2R+1 open outputs
// y' = y-R .. y+R, so 2R+1 accumulator chains advance by its 2R+1 window values each; output
// y-R completes and output y+R opens in the same slot (slots = row index mod 2R).  Same lag of
// R rows per level as the star, so the streaming code is shared.
// Radius 3 and 4 (49- and 81-point): 2R accumulator chains + R rows of state per point leave room for one level.
__host__ __device__ constexpr int w2r_box_max_k(int R, int elem_size) {
    return R == 2 ? (elem_size == 4 ? 2 : 1) : (R == 3 || R == 4 ? 1 : 0);
}

template <class T, int R> struct W2rBoxLevel {
    static constexpr int PHASES = 2 * R;
    T prev[R][4];       // the last R input rows (pass-through values of the rows still to be emitted)
    T pend[2 * R][4];   // open accumulator chains
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int s = 0; s < R; ++s) prev[s][i] = T(0);
#pragma unroll
            for (int s = 0; s < 2 * R; ++s) pend[s][i] = T(0);
        }
    }
    template <int P, bool EDGE>
    __device__ __forceinline__ void step(const T (&w)[4 + 2 * R], T avg, unsigned upd, Row4<T>& o) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            auto chain = [&](T acc) {
#pragma unroll
                for (int k = 0; k <= 2 * R; ++k) acc = add_rn(acc, w[i + k]);
                return acc;
            };
            constexpr int SL = (P + R) % (2 * R);  // slot shared by the completing and the opening output
            const T f = mul_rn(chain(pend[SL][i]), avg);
            o.v[i] = (!EDGE || ((upd >> i) & 1u)) ? f : prev[P % R][i];
            pend[SL][i] = chain(T(0));
#pragma unroll
            for (int d = -R + 1; d <= R - 1; ++d) {
                const int sl = (P + d + 2 * R) % (2 * R);
                pend[sl][i] = chain(pend[sl][i]);
            }
            prev[P % R][i] = w[R + i];
        }
    }
};

template <class T, int R, int SHAPE> struct W2rLevelSel { using type = W2rLevel<T, R>; static constexpr int PHASES = R; };
template <class T, int R> struct W2rLevelSel<T, R, BOX> { using type = W2rBoxLevel<T, R>; static constexpr int PHASES = 2 * R; };

template <class T, int R, int K, bool EDGE, int SHAPE = STAR>
__device__ __forceinline__ void w2r_stream(const T* __restrict__ src, T* __restrict__ dst, long long pitch,
                                           int nsteps, int j_dom_lo, int j_dom_hi, int j_row_lo,
                                           int j_row_hi, unsigned mxmask, T* ring, bool lane_valid,
                                           int ncols_valid, T avg) {
    constexpr int D = w2r_depth(R, SHAPE);
    constexpr int PH = W2rLevelSel<T, R, SHAPE>::PHASES;
    static_assert(D % PH == 0 && PH <= 8, "the row loop is unrolled by a multiple of the phase count");
    typename W2rLevelSel<T, R, SHAPE>::type lv[K];
#pragma unroll
    for (int l = 0; l < K; ++l) lv[l].init();
    // unconditional, clamped loads (see warp2d.cuh): rows outside the allocation or past the
    // item re-read a valid row; nothing they return reaches a valid output
    const int j_ld_lo = EDGE ? (j_row_lo > 0 ? j_row_lo : 0) : 0;
    const int j_ld_hi = (EDGE && j_row_hi < nsteps ? j_row_hi : nsteps) - 1;
    auto row_ptr = [&](int jn) -> const T* {
        int jc = jn < j_ld_hi ? jn : j_ld_hi;
        if (EDGE) jc = jc > j_ld_lo ? jc : j_ld_lo;
        return src + (long long)jc * pitch;
    };
#pragma unroll
    for (int d = 0; d < D - 1; ++d) w2d_async_row(ring + d * W2D_WIDTH, row_ptr(d));
    T* __restrict__ sp = dst - (long long)K * R * pitch;  // row emitted by the last level at step j
    for (int j0 = 0; j0 < nsteps; j0 += D) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const int j = j0 + d;
            constexpr int dummy = 0;
            (void)dummy;
            w2d_async_row(ring + ((d + D - 1) % D) * W2D_WIDTH, row_ptr(j + D - 1));
            w2d_async_wait<D - 1>();
            Row4<T> cur = w2d_read_row(ring + d * W2D_WIDTH);
#pragma unroll
            for (int l = 0; l < K; ++l) {
                // level l+1: row j - l*R arrives (phase d mod R, D being a multiple of R), row
                // j - (l+1)*R is emitted
                unsigned upd = 0;
                if (EDGE) {
                    const int jo = j - (l + 1) * R;
                    upd = (jo >= j_dom_lo && jo < j_dom_hi) ? mxmask : 0u;
                }
                T w[4 + 2 * R];
#pragma unroll
                for (int k = 1; k <= R; ++k) {
                    w[R - k] = shfl_up1(cur.v[4 - k]);
                    w[R + 3 + k] = shfl_down1(cur.v[k - 1]);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) w[R + i] = cur.v[i];
                Row4<T> o;
                switch (d % PH) {  // compile-time after unrolling
                    case 0: lv[l].template step<0, EDGE>(w, avg, upd, o); break;
                    case 1: lv[l].template step<1 % PH, EDGE>(w, avg, upd, o); break;
                    case 2: lv[l].template step<2 % PH, EDGE>(w, avg, upd, o); break;
                    case 3: lv[l].template step<3 % PH, EDGE>(w, avg, upd, o); break;
                    case 4: lv[l].template step<4 % PH, EDGE>(w, avg, upd, o); break;
                    case 5: lv[l].template step<5 % PH, EDGE>(w, avg, upd, o); break;
                    case 6: lv[l].template step<6 % PH, EDGE>(w, avg, upd, o); break;
                    default: lv[l].template step<7 % PH, EDGE>(w, avg, upd, o); break;
                }
                cur = o;
            }
            if (lane_valid && j >= 2 * K * R && j < nsteps) {  // emitted row j - K*R is an output row
                if (ncols_valid >= W2D_PX) {
                    w2d_store(sp, cur);
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (i < ncols_valid) sp[i] = cur.v[i];
                }
            }
            sp += pitch;
        }
    }
}

// in -> out, K fused steps of the radius-R star, output rows [ob, oe) (slab-interior
// coordinates).  Work item = (x strip, y chunk), one per warp, as in sweep_warp2d_kernel.
template <class T, int R, int K, int SHAPE = STAR, bool LINKED = false>
__global__ void __launch_bounds__(32 * W2D_WARPS)
sweep_warp2dr_kernel(const T* __restrict__ in, T* __restrict__ out, Grid g, Domain dom,
                     long long ob, long long oe, int nstrips, int chunk_rows, long long nitems,
                     T avg, const SlabLinkDev link) {
    constexpr int HL = w2r_halo_left(K, R);
    constexpr int V = w2r_valid(K, R);
    constexpr int D = w2r_depth(R, SHAPE);
    const int lane = threadIdx.x & 31;
    const long long item = (long long)blockIdx.x * W2D_WARPS + (threadIdx.x >> 5);
    if (item >= nitems) return;
    const int strip = (int)(item % nstrips);
    long long chunk = item / nstrips;
    // LINKED: as in sweep_warp2d_kernel (faces are g = R * k rows, their owners the warps of the
    // first / last row chunk, dispatched first)
    if (LINKED && link.link_hi) chunk = chunk == 0 ? link.nchunks - 1 : chunk - 1;
    const bool is_lo = LINKED && link.link_lo && chunk == 0;
    const bool is_hi = LINKED && link.link_hi && chunk == link.nchunks - 1;
    const long long yb = LINKED ? ob + (oe - ob) * chunk / link.nchunks : ob + chunk * chunk_rows;
    const long long ye = LINKED ? ob + (oe - ob) * (chunk + 1) / link.nchunks
                                : ((yb + chunk_rows < oe) ? yb + chunk_rows : oe);
    if (LINKED) w2d_link_acquire(link, is_lo, is_hi, lane);

    const long long gx = (long long)strip * V - HL + lane * W2D_PX;
    long long gx_ld = gx < -(long long)g.ox ? -(long long)g.ox : gx;
    if (gx_ld + W2D_PX > g.pitch_x - g.ox) gx_ld = g.pitch_x - g.ox - W2D_PX;
    unsigned mxmask = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (gx + i >= 0 && gx + i < g.n[0]) mxmask |= 1u << i;
    const bool lane_valid = lane * W2D_PX >= HL && lane * W2D_PX < HL + V && gx < g.n[0];
    const long long rem = g.n[0] - gx;
    const int ncols_valid = rem >= W2D_PX ? W2D_PX : (rem > 0 ? (int)rem : 0);

    const long long y_first = yb - K * R;
    const int nsteps = (int)(ye - yb) + 2 * K * R;
    constexpr int SLACK = K * R + 8;
    auto rel = [&](long long y) -> int {
        const long long v = y - y_first;
        return v < -SLACK ? -SLACK : (v > nsteps + SLACK ? nsteps + SLACK : (int)v);
    };
    const int j_dom_lo = rel(dom.lo), j_dom_hi = rel(dom.hi);
    const int j_row_lo = rel(-(long long)g.oy), j_row_hi = rel(g.pitch_y - g.oy);
    const T* __restrict__ src = in + g.idx(gx_ld, y_first, 0);
    T* __restrict__ dst = out + g.idx(gx, y_first, 0);

    __shared__ __align__(16) T ring_all[W2D_WARPS][D][W2D_WIDTH];
    T* ring = &ring_all[threadIdx.x >> 5][0][lane * W2D_PX];
    // interior: every point this warp touches is updatable at every level (emitted rows span
    // j - K*R .. j - R over the run, the x trapezoid stays inside the domain)
    const bool interior =
        __all_sync(0xffffffffu, mxmask == 0xFu) && j_dom_lo <= -K * R && j_dom_hi >= nsteps;
    if (interior) {
        w2r_stream<T, R, K, false, SHAPE>(src, dst, g.pitch_x, nsteps, j_dom_lo, j_dom_hi, j_row_lo, j_row_hi,
                                   mxmask, ring, lane_valid, ncols_valid, avg);
    } else {
        w2r_stream<T, R, K, true, SHAPE>(src, dst, g.pitch_x, nsteps, j_dom_lo, j_dom_hi, j_row_lo, j_row_hi,
                                  mxmask, ring, lane_valid, ncols_valid, avg);
    }
    if (LINKED) w2d_link_push<T>(out, g, gx, lane_valid, ncols_valid, link, is_lo, is_hi, lane);
}

template <class T, int R, int K, int SHAPE = STAR>
static cudaError_t launch_warp2dr_k(const T* in, T* out, const Grid& g, Domain dom, long long ob,
                                    long long oe, const Tuning& tn, cudaStream_t st, const SlabLinkDev* link_in) {
    if (oe <= ob) return cudaSuccess;
    const W2dPlan p = w2d_plan(g.n[0], oe - ob, K, tn.sm_count, R, tn.prefer_chunk);
    const T avg = Avg<T>::value(SHAPE, 2, R);
    if (link_in && (link_in->link_lo || link_in->link_hi)) {
        SlabLinkDev link = *link_in;
        long long nc = p.nchunks;
        if (nc > (oe - ob) / link.face_planes) nc = (oe - ob) / link.face_planes;
        if (nc < 1) nc = 1;
        link.nchunks = (int)nc;
        const long long nitems = nc * p.nstrips;
        sweep_warp2dr_kernel<T, R, K, SHAPE, true><<<(unsigned)((nitems + W2D_WARPS - 1) / W2D_WARPS), 32 * W2D_WARPS, 0, st>>>(
            in, out, g, dom, ob, oe, p.nstrips, p.chunk_rows, nitems, avg, link);
        return cudaGetLastError();
    }
    const unsigned blocks = (unsigned)((p.nitems + W2D_WARPS - 1) / W2D_WARPS);
    sweep_warp2dr_kernel<T, R, K, SHAPE, false><<<blocks, 32 * W2D_WARPS, 0, st>>>(
        in, out, g, dom, ob, oe, p.nstrips, p.chunk_rows, p.nitems, avg, SlabLinkDev{});
    return cudaGetLastError();
}

// 2-D star, radius 2..4, K fused steps (1 .. w2r_max_k(R, sizeof(T))); 2-D box, radius 2..4
// (25-, 49-, 81-point), K fused steps (1 .. w2r_box_max_k).
template <class T>
static cudaError_t launch_warp2dr(const T* in, T* out, const Grid& g, int shape, Domain dom, int K, long long ob,
                                  long long oe, const Tuning& tn, cudaStream_t st, const SlabLinkDev* link = nullptr) {
    if (shape == BOX) {
        if (g.r < 2 || g.r > 4 || K < 1 || K > w2r_box_max_k(g.r, (int)sizeof(T))) return cudaErrorInvalidValue;
        if (g.r == 3) return launch_warp2dr_k<T, 3, 1, BOX>(in, out, g, dom, ob, oe, tn, st, link);
        if (g.r == 4) return launch_warp2dr_k<T, 4, 1, BOX>(in, out, g, dom, ob, oe, tn, st, link);
        if (K == 1) return launch_warp2dr_k<T, 2, 1, BOX>(in, out, g, dom, ob, oe, tn, st, link);
        if constexpr (w2r_box_max_k(2, (int)sizeof(T)) >= 2)
            return launch_warp2dr_k<T, 2, 2, BOX>(in, out, g, dom, ob, oe, tn, st, link);
        return cudaErrorInvalidValue;
    }
    if (K < 1 || g.r < 2 || g.r > 4 || K > w2r_max_k(g.r, (int)sizeof(T))) return cudaErrorInvalidValue;
#define SB200_W2R(RR, KK) \
    if constexpr (KK <= w2r_max_k(RR, (int)sizeof(T))) \
        if (g.r == RR && K == KK) return launch_warp2dr_k<T, RR, KK>(in, out, g, dom, ob, oe, tn, st, link)
    SB200_W2R(2, 1); SB200_W2R(2, 2); SB200_W2R(2, 3); SB200_W2R(2, 4);
    SB200_W2R(3, 1); SB200_W2R(3, 2);
    SB200_W2R(4, 1); SB200_W2R(4, 2);
#undef SB200_W2R
    return cudaErrorInvalidValue;
}

}  // namespace sb200
