// tma3d.cuh — 3-D radius-1 star (7-point) sweep, K time steps fused per HBM round trip.
//
// B200 design: a CTA owns an xy tile (32*PXV x 64 points incl. a K-deep halo) and
// streams along z, the slowest axis.  Level-0 planes are staged into a shared-memory
// ring by TMA (cp.async.bulk.tensor.3d + mbarrier expect-tx — the sm_100a counterpart
// of the reference's athread_dma_iget + reply counter, stencil_dma.cpp:236-247,448),
// out-of-range box parts are zero-filled by the TMA unit so loads need no bounds
// code.  Each thread owns PXV consecutive x (one 128-bit shared load) by 4 rows;
// x-neighbours come from adjacent lanes by warp shuffle (a warp spans the tile's full
// x extent), y-neighbours from shared memory, and the z-1 / z+1 planes are folded into
// per-thread register partial sums, so every plane of every time level is read from
// shared memory exactly once.  Time levels 1..K-1 ping-pong through double-buffered
// shared planes: ONE __syncthreads per streamed plane regardless of K.  The final level
// is stored straight from registers with 128-bit coalesced stores.
//
// Per-point order is the oracle's (generalising reference
// stencil.cpp:103-125):  ((((((0+xl)+xr)+yl)+yr)+zl)+zr) * avg, hence bit-identical for
// any K.  Points outside the update domain pass through unchanged at every level.
#pragma once
#include <cuda.h>

#include <algorithm>
#include <array>
#include <functional>
#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "slablink.cuh"

namespace sb200 {

template <class T, int K> struct T3 {
    static constexpr int PXV = 16 / (int)sizeof(T);  // x points per thread (128-bit)
    static constexpr int PY = 4;                     // rows per thread
#ifdef SB200_T3_TYN  // experiment: 64 x 32 tiles, two CTAs per SM
    static constexpr int TXN = 32, TYN = SB200_T3_TYN;
    static constexpr int CTAS_PER_SM = 16 / SB200_T3_TYN;
#else
    // K = 3 keeps nine plane-sized arrays of state per thread: 12 warps (64 x 48 tiles) leave
    // every thread 168 registers instead of 128
    static constexpr int TXN = 32, TYN = K >= 3 ? 12 : 16;
    static constexpr int CTAS_PER_SM = 1;
#endif
    static constexpr int THREADS = TXN * TYN;
    static constexpr int BX = TXN * PXV, BY = TYN * PY;
    static constexpr int HLX = (K + PXV - 1) / PXV * PXV;
    static constexpr int VX = (BX - HLX - K) / PXV * PXV;
    static constexpr int VY = BY - 2 * K;
    static constexpr int STAGES = 4;
    static constexpr int PLANE = BX * BY;  // elements
    static constexpr size_t PLANE_BYTES = (size_t)PLANE * sizeof(T);
    // levels 1..K-1 only publish each thread's first and last row (what its y-neighbour
    // threads need); the rows in between stay in the owner's registers
    static constexpr int EDGE = TYN * 2 * BX;  // elements per edge buffer
    // the last level's planes leave through a dense VX x VY staging tile in shared memory and
    // a TMA store (two slots: the store of step j reads its slot while step j+1 fills the other)
    static constexpr int STAGE = VX * VY;  // elements
    static constexpr size_t STAGE_BYTES = ((size_t)STAGE * sizeof(T) + 127) / 128 * 128;
    static constexpr size_t SMEM =
        PLANE_BYTES * STAGES + (size_t)EDGE * sizeof(T) * 2 * (K - 1) + 2 * STAGE_BYTES + 8 * STAGES + 128;
};
// tools/sm_bound_probe.sh builds a second library with -DSB200_PROBE_ZWRAP=8: every CTA then
// streams the same 8 planes over and over (loads and stores stay in L2, results are garbage),
// which measures the SM-side rate of the kernel without the HBM limit.  Never defined in
// the product build.
#ifdef SB200_PROBE_ZWRAP
#define T3_PROBE_Z(j) ((j) & (SB200_PROBE_ZWRAP - 1))
#else
#define T3_PROBE_Z(j) (j)
#endif
constexpr int T3_MAXK = 2;      // star: K = 3 (12-warp tiles, 168 registers) is correct but slower, 599 vs 705
                                // GStencil/s on config 2 (profiles/r02e_tma_store_probe.log); not instantiated
constexpr int T3_MAXK_BOX = 2;  // box (27-point)

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// Store of a dense smem box into the tensor (bulk async-group of the calling thread); parts of
// the box outside the tensor's extents are clipped by the TMA unit.
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, const void* src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map),
                 "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read() {  // every committed store has read its smem
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_all() {  // ... and has been written
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar,
                                            int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

template <class T, int N> struct VecN;
template <> struct VecN<float, 4> {
    static __device__ __forceinline__ void ld(const float* p, float* v) {
        const float4 f = *reinterpret_cast<const float4*>(p);
        v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    }
    static __device__ __forceinline__ void ldcg(const float* p, float* v) {
        const float4 f = __ldcg(reinterpret_cast<const float4*>(p));
        v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    }
    static __device__ __forceinline__ void st(float* p, const float* v) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <> struct VecN<double, 2> {
    static __device__ __forceinline__ void ld(const double* p, double* v) {
        const double2 f = *reinterpret_cast<const double2*>(p);
        v[0] = f.x; v[1] = f.y;
    }
    static __device__ __forceinline__ void ldcg(const double* p, double* v) {
        const double2 f = __ldcg(reinterpret_cast<const double2*>(p));
        v[0] = f.x; v[1] = f.y;
    }
    static __device__ __forceinline__ void st(double* p, const double* v) {
        *reinterpret_cast<double2*>(p) = make_double2(v[0], v[1]);
    }
};

// Per-thread partial-sum state of one time level (PY rows x PXV points).  The centre rows of
// the two most recent planes of every level live in two-slot register rings owned by the
// kernel (indexed by the plane-loop phase, a compile-time constant after unrolling by two),
// so advancing a plane never copies registers.
template <class T, int PXV, int PY> struct T3Star {
    T t[PY][PXV];  // (in-plane partial of plane a-1) + centre of plane a-2
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int i = 0; i < PY; ++i)
#pragma unroll
            for (int k = 0; k < PXV; ++k) t[i][k] = T(0);
    }
};

// Plane a of the level below arrives: centre rows c[0..PY), the row above (up) and below
// (dn) this thread's rows; cm1 = centre rows of plane a-1.  Emits plane a-1 of this level
// into o[][] for every point as if it were updatable (the caller restores pass-through
// points), and advances the partial sums to plane a.
template <class T, int PXV, int PY>
__device__ __forceinline__ void t3_level(const T (&up)[PXV], const T (&c)[PY][PXV], const T (&dn)[PXV],
                                         const T (&cm1)[PY][PXV], T3Star<T, PXV, PY>& L, T avg,
                                         T (&o)[PY][PXV]) {
#pragma unroll
    for (int i = 0; i < PY; ++i) {
#if defined(SB200_T3_EXP) && (SB200_T3_EXP & 8)  // ablation (wrong results): no shuffles
        const T lft = c[i][PXV - 1], rgt = c[i][0];
#else
        const T lft = __shfl_up_sync(0xffffffffu, c[i][PXV - 1], 1);
        const T rgt = __shfl_down_sync(0xffffffffu, c[i][0], 1);
#endif
#pragma unroll
        for (int k = 0; k < PXV; ++k) {
            const T xl = k == 0 ? lft : c[i][k - 1];
            const T xr = k == PXV - 1 ? rgt : c[i][k + 1];
            const T yl = i == 0 ? up[k] : c[i - 1][k];
            const T yr = i == PY - 1 ? dn[k] : c[i + 1][k];
#if defined(SB200_T3_EXP) && (SB200_T3_EXP & 2)
            const T sum = add_rn(add_rn(add_rn(xl, xr), yl), yr);  // experiment: without the oracle's 0 + xl
#else
            const T sum = add_rn(add_rn(add_rn(add_rn(T(0), xl), xr), yl), yr);
#endif
            o[i][k] = mul_rn(add_rn(L.t[i][k], c[i][k]), avg);
            L.t[i][k] = add_rn(sum, cm1[i][k]);
        }
    }
}

// ---- 27-point box ---------------------------------------------------------------------
// Oracle order: sum over dz (outer), dy, dx (inner), ascending, starting
// from 0, then one multiply.  An arriving plane is the dz=-1 plane of output a+1, the dz=0
// plane of output a and the dz=+1 plane of output a-1, so three accumulator chains advance
// by nine sequential adds each; no partial sums are shared (that would reorder the adds).
template <class T, int PXV, int PY> struct T3Box {
    T a[PY][PXV];  // chain of the output whose first plane (dz=-1) has been added
    T b[PY][PXV];  // chain of the output whose first two planes have been added
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int i = 0; i < PY; ++i)
#pragma unroll
            for (int k = 0; k < PXV; ++k) a[i][k] = b[i][k] = T(0);
    }
};

template <class T, int PXV, int PY>
__device__ __forceinline__ void t3_level(const T (&up)[PXV], const T (&c)[PY][PXV], const T (&dn)[PXV],
                                         const T (&)[PY][PXV], T3Box<T, PXV, PY>& L, T avg,
                                         T (&o)[PY][PXV]) {
    T lft[PY + 2], rgt[PY + 2];
    lft[0] = __shfl_up_sync(0xffffffffu, up[PXV - 1], 1);
    rgt[0] = __shfl_down_sync(0xffffffffu, up[0], 1);
#pragma unroll
    for (int i = 0; i < PY; ++i) {
        lft[i + 1] = __shfl_up_sync(0xffffffffu, c[i][PXV - 1], 1);
        rgt[i + 1] = __shfl_down_sync(0xffffffffu, c[i][0], 1);
    }
    lft[PY + 1] = __shfl_up_sync(0xffffffffu, dn[PXV - 1], 1);
    rgt[PY + 1] = __shfl_down_sync(0xffffffffu, dn[0], 1);
#pragma unroll
    for (int i = 0; i < PY; ++i) {
#pragma unroll
        for (int k = 0; k < PXV; ++k) {
            T f = L.b[i][k], nb = L.a[i][k], na = T(0);
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) {
                const int r = i + dy;  // 0 = up, 1..PY = c, PY+1 = dn
                const T* row = r == 0 ? up : (r == PY + 1 ? dn : c[r - 1]);
                const T xl = k == 0 ? lft[r] : row[k - 1];
                const T xc = row[k];
                const T xr = k == PXV - 1 ? rgt[r] : row[k + 1];
                f = add_rn(add_rn(add_rn(f, xl), xc), xr);
                nb = add_rn(add_rn(add_rn(nb, xl), xc), xr);
                na = add_rn(add_rn(add_rn(na, xl), xc), xr);
            }
            o[i][k] = mul_rn(f, avg);
            L.b[i][k] = nb;
            L.a[i][k] = na;
        }
    }
}

// ---- 27-point box, separable sums (SB200_FLAG_REASSOCIATE) -----------------------------------
// row sums rs = (xl + xc) + xr of the PY + 2 rows a thread sees, plane sums sxy = (rs(y-1) +
// rs(y)) + rs(y+1), then the z chain with the same two accumulators as the exact variant:
// out(a-1) = ((sxy(a-2) + sxy(a-1)) + sxy(a)) * avg.  8 operations per point instead of 28, in
// an order the reference does not use: equal within 1e-12 (fp64) / 1e-5 (fp32) relative.
template <class T, int PXV, int PY> struct T3BoxSep : T3Box<T, PXV, PY> {};

template <class T, int PXV, int PY>
__device__ __forceinline__ void t3_level(const T (&up)[PXV], const T (&c)[PY][PXV], const T (&dn)[PXV],
                                         const T (&)[PY][PXV], T3BoxSep<T, PXV, PY>& L, T avg,
                                         T (&o)[PY][PXV]) {
    T rs_prev[PXV], rs_cur[PXV], rs_next[PXV];
    auto row_sum = [&](const T (&row)[PXV], T (&rs)[PXV]) {
        const T lft = __shfl_up_sync(0xffffffffu, row[PXV - 1], 1);
        const T rgt = __shfl_down_sync(0xffffffffu, row[0], 1);
#pragma unroll
        for (int k = 0; k < PXV; ++k)
            rs[k] = add_rn(add_rn(k == 0 ? lft : row[k - 1], row[k]), k == PXV - 1 ? rgt : row[k + 1]);
    };
    row_sum(up, rs_prev);
    row_sum(c[0], rs_cur);
#pragma unroll
    for (int i = 0; i < PY; ++i) {
        if (i + 1 < PY) row_sum(c[i + 1 < PY ? i + 1 : i], rs_next);
        else row_sum(dn, rs_next);
#pragma unroll
        for (int k = 0; k < PXV; ++k) {
            const T sxy = add_rn(add_rn(rs_prev[k], rs_cur[k]), rs_next[k]);
            o[i][k] = mul_rn(add_rn(L.b[i][k], sxy), avg);
            L.b[i][k] = add_rn(L.a[i][k], sxy);
            L.a[i][k] = sxy;
            rs_prev[k] = rs_cur[k];
            rs_cur[k] = rs_next[k];
        }
    }
}

template <class T, int PXV, int PY, int SHAPE> struct T3LevelSel { using type = T3Star<T, PXV, PY>; };
template <class T, int PXV, int PY> struct T3LevelSel<T, PXV, PY, BOXSEP> { using type = T3BoxSep<T, PXV, PY>; };
template <class T, int PXV, int PY> struct T3LevelSel<T, PXV, PY, BOX> { using type = T3Box<T, PXV, PY>; };

template <int V> struct IC { static constexpr int value = V; };

template <class T, int K, int SHAPE, bool LINKED>
__global__ void __launch_bounds__(T3<T, K>::THREADS, T3<T, K>::CTAS_PER_SM)
sweep_tma3d_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap omap,
                   const T* __restrict__ in, T* __restrict__ out, Grid g, Domain dom, long long ob, long long oe, int ntx, int nty,
                   int chunk_planes, T avg, const SlabLinkDev link) {
    using C = T3<T, K>;
    constexpr int PXV = C::PXV, PY = C::PY, S = C::STAGES;
    static_assert((S & (S - 1)) == 0 && S >= 2, "stage count must be a power of two");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // 128-byte align the TMA destinations
    unsigned char* base = smem_raw + ((128 - (smem_u32(smem_raw) & 127)) & 127);
    T* ring = reinterpret_cast<T*>(base);
    T* ebuf = ring + (size_t)S * C::PLANE;
    unsigned char* stage = reinterpret_cast<unsigned char*>(ebuf + (size_t)2 * (K - 1) * C::EDGE);
    uint64_t* bars = reinterpret_cast<uint64_t*>(stage + 2 * C::STAGE_BYTES);

    const int tid = threadIdx.x;
    const int tx = tid & 31, ty = tid >> 5;
    const int tiles = ntx * nty;
    const int tile = (int)(blockIdx.x % (unsigned)tiles);
    const int tix = tile % ntx, tiy = tile / ntx;
    const int slot = (int)(blockIdx.x / (unsigned)tiles);
    // linked launches dispatch the LAST chunk first: its CTAs deliver the high face only at
    // the end of their z stream
    int chunk = slot;
    if (LINKED && link.link_hi) chunk = slot == 0 ? link.nchunks - 1 : slot - 1;
    const bool is_lo = LINKED && link.link_lo && chunk == 0;                // owns the low face
    const bool is_hi = LINKED && link.link_hi && chunk == link.nchunks - 1;  // owns the high face
    // linked launches balance the chunks (every chunk >= g planes, so the first / last chunk
    // contains the whole low / high face)
    const long long zb = LINKED ? ob + (oe - ob) * chunk / link.nchunks : ob + (long long)chunk * chunk_planes;
    const long long ze = LINKED ? ob + (oe - ob) * (chunk + 1) / link.nchunks
                                : (zb + chunk_planes < oe ? zb + chunk_planes : oe);
    const long long gx0 = (long long)tix * C::VX - C::HLX;  // interior coords of tile (0,0)
    const long long gy0 = (long long)tiy * C::VY - K;
    const long long z_first = zb - K;
    const int nin = (int)(ze - zb) + 2 * K;  // level-0 planes streamed
    const int nsteps = nin + K - 1;

    if (tid == 0) {
        for (int s = 0; s < S; ++s) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (LINKED && (is_lo || is_hi) && link.wait_value != 0) {
            // the neighbours' previous launch must have delivered its face planes into the
            // ghost planes this chunk is about to stream
            if (is_lo) link_wait(link.wait_lo, link.wait_value, link.err_flag);
            if (is_hi) link_wait(link.wait_hi, link.wait_value, link.err_flag);
            // TMA reads go through the async proxy
            asm volatile("fence.proxy.async.global;" ::: "memory");
        }
    }
    __syncthreads();
    const int cx = (int)(gx0 + g.ox), cy = (int)(gy0 + g.oy);
    const int cz0 = (int)(z_first + g.oz);
    if (tid == 0) {
        for (int s = 0; s < S - 1 && s < nin; ++s) {
            mbar_expect_tx(&bars[s], (uint32_t)C::PLANE_BYTES);
            tma_load_3d(ring + (size_t)s * C::PLANE, &tmap, &bars[s], cx, cy, cz0 + T3_PROBE_Z(s));
        }
    }

    // per-thread constants: update mask in x/y, store validity, shared-memory offsets
    const int lx = tx * PXV, ly = ty * PY;
    const long long gx = gx0 + lx, gy = gy0 + ly;
    unsigned mask = 0, rows_ok = 0;
#pragma unroll
    for (int i = 0; i < PY; ++i) {
        const bool yin = gy + i >= 0 && gy + i < g.n[1];
        if (yin && ly + i >= K && ly + i < K + C::VY) rows_ok |= 1u << i;
#pragma unroll
        for (int k = 0; k < PXV; ++k)
            if (yin && gx + k >= 0 && gx + k < g.n[0]) mask |= 1u << (i * PXV + k);
    }
    const bool warp_all = __all_sync(0xffffffffu, mask == (1u << (PY * PXV)) - 1u);
    const bool col_ok = lx >= C::HLX && lx < C::HLX + C::VX && gx < g.n[0];
    const bool col_full = gx + PXV <= g.n[0];
    if (!col_ok) rows_ok = 0;
    // staging tile of the TMA store: rows / columns of the tile's valid VX x VY region, whether
    // or not they lie inside the grid (the TMA unit clips)
    unsigned s_rows = 0;
#pragma unroll
    for (int i = 0; i < PY; ++i)
        if (ly + i >= K && ly + i < K + C::VY) s_rows |= 1u << i;
    if (!(lx >= C::HLX && lx < C::HLX + C::VX)) s_rows = 0;
    const int s_off = (ly - K) * C::VX + (lx - C::HLX);
    const int x_tail = gx == g.n[0] / PXV * PXV ? (int)(g.n[0] - gx) : 0;  // 0 .. PXV-1
    // store coordinates in the output map (x, y: interior origin; z: plane index of the allocation)
    const int ox0 = tix * C::VX, oy0 = tiy * C::VY, oz0 = (int)(z_first + g.oz);
    const int toff = ly * C::BX + lx;
    const int up_off = ly > 0 ? -C::BX : 0;
    const int dn_off = ly + PY < C::BY ? PY * C::BX : (PY - 1) * C::BX;
    // edge buffers: [thread row ty][0 = first row, 1 = last row][BX]
    const int e_own = ty * 2 * C::BX + lx;
    const int e_up = (ty > 0 ? (ty - 1) * 2 + 1 : 0) * C::BX + lx;
    const int e_dn = (ty + 1 < C::TYN ? (ty + 1) * 2 : ty * 2 + 1) * C::BX + lx;
    // plane bookkeeping in local indices p (global z = z_first + p)
    const long long dlo = dom.lo - z_first, dhi = dom.hi - z_first;
    const int p_dom_lo = dlo < -4 ? -4 : (int)dlo;
    const int p_dom_hi = dhi > nin + 4 ? nin + 4 : (int)dhi;
    const int p_out_lo = K, p_out_hi = nin - K;

    // Register state.  R: level-0 centre rows of the two newest planes; O[l]: the two newest
    // planes emitted by level l+1 (l < K-1); slot = plane-loop phase.  Every level runs on
    // every step: before its first real plane a level only sees the zero-initialised rings
    // (its state stays zero), and what the levels compute after the last streamed plane lies
    // beyond the last stored plane.
    // The star needs the previous plane of every level for its arithmetic (the z-low term is
    // added after the in-plane ones), so R and O are two-slot rings indexed by the phase.  The
    // box variants only ever read the previous plane in the pass-through fix-up of a level that
    // is not the last: they keep ONE slot (6 plane-sized arrays at K = 2 instead of 8: no
    // spills) and the fix-up of level 0 re-reads the plane from the TMA ring, which therefore
    // runs S-1 planes ahead, not S.
    constexpr bool HIST = SHAPE == STAR;
    static_assert(HIST || K <= 2, "box variants: the only level that is not the last one is level 0");
    constexpr int NS = HIST ? 2 : 1;
    T R[NS][PY][PXV];
    T O[K > 1 ? K - 1 : 1][NS][PY][PXV];
    typename T3LevelSel<T, PXV, PY, SHAPE>::type lv[K];
#pragma unroll
    for (int l = 0; l < K; ++l) lv[l].init();
#pragma unroll
    for (int s = 0; s < NS; ++s)
#pragma unroll
        for (int i = 0; i < PY; ++i)
#pragma unroll
            for (int k = 0; k < PXV; ++k) {
                R[s][i][k] = T(0);
#pragma unroll
                for (int l = 0; l < K - 1; ++l) O[l][s][i][k] = T(0);
            }

    const long long plane_stride = g.pitch_x * g.pitch_y;
    const long long row_stride = g.pitch_x;
    // One time level of one plane step.  L = level index (level L+1 of the fused block),
    // P = j & 1.  Level L consumes the plane its lower level emitted on the PREVIOUS step
    // (registers + the edge rows published in shared memory), so one barrier per step
    // orders everything.
    auto level = [&](auto Lc, auto Pc, int j) {
        constexpr int L = decltype(Lc)::value, P = decltype(Pc)::value;
        T up[PXV], dn[PXV];
        if constexpr (L == 0) {
            const int st = j & (S - 1);
            if (j < nin) mbar_wait(&bars[st], (uint32_t)(j / S) & 1u);
            const T* src = ring + (size_t)st * C::PLANE + toff;
            VecN<T, PXV>::ld(src + up_off, up);
#pragma unroll
            for (int i = 0; i < PY; ++i) VecN<T, PXV>::ld(src + i * C::BX, R[NS == 2 ? P : 0][i]);
            VecN<T, PXV>::ld(src + dn_off, dn);
        } else {
            const T* eb = ebuf + (size_t)(2 * (L - 1) + (P ^ 1)) * C::EDGE;
            VecN<T, PXV>::ld(eb + e_up, up);
            VecN<T, PXV>::ld(eb + e_dn, dn);
        }
        const T (&c)[PY][PXV] = *(L == 0 ? &R[NS == 2 ? P : 0] : &O[L > 0 ? L - 1 : 0][NS == 2 ? P ^ 1 : 0]);
        const T (&cm1)[PY][PXV] = *(L == 0 ? &R[NS == 2 ? P ^ 1 : 0] : &O[L > 0 ? L - 1 : 0][NS == 2 ? P : 0]);
        const int pe = j - 2 * L - 1;  // local index of the plane being emitted
        const bool upd = pe >= p_dom_lo && pe < p_dom_hi;
        if constexpr (L < K - 1) {
            T (&o)[PY][PXV] = O[L][NS == 2 ? P : 0];
            t3_level<T, PXV, PY>(up, c, dn, cm1, lv[L], avg, o);
            if (!(upd && warp_all)) {
                const unsigned m = upd ? mask : 0u;
                // plane j-1 of the TMA ring (box, level 0): still there, the ring runs S-1 ahead
                const T* prev = ring + (size_t)((j - 1) & (S - 1)) * C::PLANE + toff;
#pragma unroll
                for (int i = 0; i < PY; ++i)
#pragma unroll
                    for (int k = 0; k < PXV; ++k)
                        if (!((m >> (i * PXV + k)) & 1u)) o[i][k] = HIST ? cm1[i][k] : prev[i * C::BX + k];
            }
            T* wb = ebuf + (size_t)(2 * L + P) * C::EDGE + e_own;
            VecN<T, PXV>::st(wb, o[0]);
            VecN<T, PXV>::st(wb + C::BX, o[PY - 1]);
        } else {
            // The last level needs no pass-through fix-up: only updatable points of planes
            // inside the update domain are ever stored.  The box variant drops it (and with it
            // the register history of the previous plane, which only the fix-up reads: no
            // spills at K=2); the star variant carries that history anyway for its
            // arithmetic and keeps the fix-up — without it ptxas spills inside the plane loop.
            T o[PY][PXV];
            t3_level<T, PXV, PY>(up, c, dn, cm1, lv[L], avg, o);
            if (SHAPE == STAR && !(upd && warp_all)) {
                const unsigned m = upd ? mask : 0u;
#pragma unroll
                for (int i = 0; i < PY; ++i)
#pragma unroll
                    for (int k = 0; k < PXV; ++k)
                        if (!((m >> (i * PXV + k)) & 1u)) o[i][k] = cm1[i][k];
            }
            // out through the staging tile: dense VX x VY, this thread's rows at (ly - K, lx - HLX)
            if (pe >= p_out_lo && pe < p_out_hi && s_rows) {
                T* sp = reinterpret_cast<T*>(stage + (size_t)P * C::STAGE_BYTES) + s_off;
#pragma unroll
                for (int i = 0; i < PY; ++i)
                    if (s_rows & (1u << i)) VecN<T, PXV>::st(sp + i * C::VX, o[i]);
                if (x_tail) {
                    // the TMA unit clips in 16-byte units, so the output map ends at the last
                    // full 16 bytes of a row: the thread that owns the ragged end of the rows
                    // (n[0] mod PXV points) stores it itself
                    T* gp = out + g.idx(gx, gy, z_first + pe);
#pragma unroll
                    for (int i = 0; i < PY; ++i)
                        if ((s_rows & (1u << i)) && gy + i < g.n[1])
#pragma unroll
                            for (int k = 0; k < PXV - 1; ++k)
                                if (k < x_tail) gp[(long long)i * g.pitch_x + k] = o[i][k];
                }
            }
        }
    };
    auto step = [&](auto Pc, int j) {
        // higher levels first: they consume what the level below emitted one step earlier
#if defined(SB200_T3_EXP) && (SB200_T3_EXP & 1)
        if (K == 2 && ((ty >> 2) & 1)) {  // experiment: half of the warps of a sub-partition run the levels in the other order
            level(IC<0>{}, Pc, j);
            if constexpr (K > 1) level(IC<1>{}, Pc, j);
        } else
#endif
        {
        if constexpr (K > 2) level(IC<2>{}, Pc, j);
        if constexpr (K > 1) level(IC<1>{}, Pc, j);
        level(IC<0>{}, Pc, j);
        }
        // The plane the last level just staged (local index j - 2K + 1) leaves through a TMA
        // store issued by thread 0 after the barrier: 512-byte rows straight into L2 without
        // passing the LSU (an STG version of this kernel spent a third of its time on its
        // stores: profiles/r02d_kernel_experiments.log), clipped to the interior by the TMA
        // unit, so ragged tiles need no predication.  Slot P is refilled two steps later; the
        // store that read the OTHER slot is waited for before the barrier that lets step j+1
        // overwrite it.
        constexpr int P = decltype(Pc)::value;
        const int po = j - 2 * K + 1;
        const bool storing = po >= p_out_lo && po < p_out_hi;
        if (storing) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (tid == 0) tma_store_wait_read();
        __syncthreads();
        if (tid == 0) {
            if (storing && g.n[0] >= PXV)  // narrower grids are all "ragged end"
                tma_store_3d(&omap, stage + (size_t)P * C::STAGE_BYTES, ox0, oy0,
#ifdef SB200_PROBE_ZWRAP
                             oz0 + K + ((po - K) & (SB200_PROBE_ZWRAP - 1)));
#else
                             oz0 + po);
#endif
            if (j + S - 1 < nin) {  // into the slot of plane j-1, which step j was the last to read
                const int st = (j + S - 1) & (S - 1);
                mbar_expect_tx(&bars[st], (uint32_t)C::PLANE_BYTES);
                tma_load_3d(ring + (size_t)st * C::PLANE, &tmap, &bars[st], cx, cy, cz0 + T3_PROBE_Z(j + S - 1));
            }
        }
    };
    // steps past nsteps (loop rounded up to the phase count) wait for nothing and store nothing
#pragma unroll 1
    for (int j = 0; j < nsteps; j += 2) {
        step(IC<0>{}, j);
        step(IC<1>{}, j + 1);
    }
    if (tid == 0) tma_store_wait_all();  // the staging slots outlive the CTA otherwise
    if (LINKED && (is_lo || is_hi)) {
        // Face push, after the z stream so the plane loop carries no link code at all: every
        // thread copies the face-plane points of its part of the tile (written by this CTA's TMA
        // stores, all complete: thread 0 waited, then the barrier and the proxy fence) into the
        // neighbour's ghost planes with 128-bit stores over NVLink; after the next barrier
        // thread 0 publishes them system-wide and counts this CTA in on the neighbour's
        // arrival counter.
        __syncthreads();
        asm volatile("fence.proxy.async.global;" ::: "memory");
        T* home = out + g.idx(gx, gy, 0);
#pragma unroll 1
        for (int side = 0; side < 2; ++side) {
            if (side == 0 ? !is_lo : !is_hi) continue;
            const long long z0 = side == 0 ? 0 : g.n[2] - link.face_planes;
            const long long rdelta = (static_cast<T*>(side == 0 ? link.peer_out_lo : link.peer_out_hi) - out) +
                                     (side == 0 ? link.shift_lo : link.shift_hi);
#pragma unroll 1
            for (int p = 0; p < link.face_planes; ++p) {
                T* src = home + (z0 + p) * plane_stride;
#pragma unroll
                for (int i = 0; i < PY; ++i) {
                    if (rows_ok & (1u << i)) {
                        if (col_full) {
                            T v[PXV];
                            VecN<T, PXV>::ldcg(src + i * row_stride, v);
                            VecN<T, PXV>::st(src + rdelta + i * row_stride, v);
                        } else {
#pragma unroll
                            for (int k = 0; k < PXV; ++k)
                                if (gx + k < g.n[0]) src[rdelta + i * row_stride + k] = __ldcg(src + i * row_stride + k);
                        }
                    }
                }
            }
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence_system();
            if (is_lo) red_release_sys_add_u32(link.signal_lo, 1u);
            if (is_hi) red_release_sys_add_u32(link.signal_hi, 1u);
        }
    }
}

struct T3Plan {
    int ntx, nty, chunk_planes;
    long long nchunks, blocks;
};

// Chunk count along z.  Every CTA streams its chunk plus `fill` pipeline planes (3K-1 for the
// radius-1 kernels), CTAs are dispatched in blockIdx order to the first free SM: pick the chunk
// count with the smallest simulated makespan (in plane-steps).
static inline T3Plan t3_plan_core(long long nx, long long ny, long long planes, int sm_slots, int vx, int vy,
                                  int fill, long long variant, int prefer_planes) {
    T3Plan p;
    p.ntx = (int)((nx + vx - 1) / vx);
    p.nty = (int)((ny + vy - 1) / vy);
    const long long tiles = (long long)p.ntx * p.nty;
    if (prefer_planes > 0) {  // the caller's tile preference wins over the makespan search
        p.chunk_planes = (int)(prefer_planes < planes ? prefer_planes : planes);
        p.nchunks = (planes + p.chunk_planes - 1) / p.chunk_planes;
        p.blocks = tiles * p.nchunks;
        return p;
    }
    // one cache shared by every host thread (slab workers, planners)
    static std::map<std::array<long long, 6>, std::pair<int, long long>> cache;
    static std::mutex cache_mu;
    std::lock_guard<std::mutex> cache_lock(cache_mu);
    const std::array<long long, 6> key = {tiles, planes, sm_slots, fill, variant, 0};
    auto it = cache.find(key);
    if (it == cache.end()) {
        const double startup = 3.0;  // CTA launch + first TMA round trip, in plane-steps
        double best = 1e300;
        int best_cp = (int)planes;
        long long best_nc = 1;
        std::vector<double> heap;
        for (long long c = 1; c <= 256 && c <= planes; ++c) {
            const long long cp = (planes + c - 1) / c;
            const long long nc = (planes + cp - 1) / cp;
            if (c > 1 && nc == (planes + (planes + c - 2) / (c - 1) - 1) / ((planes + c - 2) / (c - 1)) &&
                cp == (planes + c - 2) / (c - 1))
                continue;  // same split as the previous candidate
            const double dfull = cp + fill + startup;
            const double dlast = (planes - (nc - 1) * cp) + fill + startup;
            // min-heap of SM free times
            heap.assign((size_t)sm_slots, 0.0);
            auto push_cta = [&](double d) {
                std::pop_heap(heap.begin(), heap.end(), std::greater<double>());
                heap.back() += d;
                std::push_heap(heap.begin(), heap.end(), std::greater<double>());
            };
            for (long long ch = 0; ch < nc; ++ch)
                for (long long i = 0; i < tiles; ++i) push_cta(ch + 1 < nc ? dfull : dlast);
            const double makespan = *std::max_element(heap.begin(), heap.end());
            if (makespan < best - 1e-9) {
                best = makespan;
                best_cp = (int)cp;
                best_nc = nc;
            }
        }
        if (cache.size() > 256) cache.clear();
        it = cache.emplace(key, std::make_pair(best_cp, best_nc)).first;
    }
    p.chunk_planes = it->second.first;
    p.nchunks = it->second.second;
    p.blocks = tiles * p.nchunks;
    return p;
}

template <class T, int K>
static T3Plan t3_plan(long long nx, long long ny, long long planes, int sm_count, int prefer_planes = 0) {
    using C = T3<T, K>;
    return t3_plan_core(nx, ny, planes, sm_count * C::CTAS_PER_SM, C::VX, C::VY, 3 * K - 1, 0, prefer_planes);
}

}  // namespace sb200
