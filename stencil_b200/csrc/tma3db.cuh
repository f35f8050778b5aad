// tma3db.cuh — 3-D box stencil of radius R >= 2 (125-point at R = 2), one time step per launch, in
// the tma3d framework (`-r 2` on a 3-D grid with the box footprint: program_options.cpp:22-23; the
// reference's general-radius tile kernel is neighbors_impl, stencil_dma.cpp:571-1099).
//
// (2R+1)^3 ordered adds per point make this kernel FP-issue bound, not memory bound (125 adds for
// 8 or 16 algorithmic bytes), so the design minimises everything that is not an add:
//   * a CTA owns a 128/64 x 64 xy tile with an R-wide halo and streams along z; planes arrive
//     through the 4-slot TMA ring of tma3d.cuh, the finished plane leaves through its staging
//     tile + TMA store;
//   * a thread owns PXV x PY points.  The oracle's sum of a point is ONE chain over dz (outer),
//     dy, dx (inner) — SURVEY.md 8(c) — so a plane contributes a contiguous run of (2R+1)^2 adds
//     to each of the 2R+1 output planes around it.  Every thread keeps 2R+1 running sums per
//     point (slot q = output plane j-R+q while plane j is being consumed); the plane's rows are
//     read ONCE, row by row (one 128-bit shared load + 2R shuffles for the x window), and each
//     value read feeds all 2R+1 chains: 1 shared load per ~150 adds.  After the plane, slot 0 is
//     complete (multiply by avg, store), the slots shift down and slot 2R restarts from zero.
// Each chain is exactly the oracle's order: bit-identical.
#pragma once
#include "tma3dr.cuh"

namespace sb200 {

// 8-byte vectors of floats: the narrow fp32 tile (VecN is declared in tma3d.cuh)
template <> struct VecN<float, 2> {
    static __device__ __forceinline__ void ld(const float* p, float* v) {
        const float2 f = *reinterpret_cast<const float2*>(p);
        v[0] = f.x; v[1] = f.y;
    }
    static __device__ __forceinline__ void ldcg(const float* p, float* v) {
        const float2 f = __ldcg(reinterpret_cast<const float2*>(p));
        v[0] = f.x; v[1] = f.y;
    }
    static __device__ __forceinline__ void st(float* p, const float* v) {
        *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
    }
};

// Tile variants (PXV_ points of x per lane, PY_ rows per thread; PXV_ * PY_ * (2R+1) running sums
// must fit the register file): fp64 2 x 4 (64 x 64 tile, 60 x 60 stored); fp32 4 x 4 (128 x 64,
// 120 x 60 stored) or 2 x 8 (64 x 128, 56 x 124 stored).
// x is quantised per LANE (a warp spans the tile's width, every lane computes), y per WARP (a warp
// that stores nothing skips the adds), so the host picks the width that wastes fewer columns.
template <class T, int R, int PXV_, int PY_> struct T3B {
    static constexpr int PXV = PXV_;
    static constexpr int PY = PY_;
    static constexpr int TXN = 32, TYN = 16;
    static constexpr int THREADS = TXN * TYN;
    static constexpr int BX = TXN * PXV, BY = TYN * PY;
    // x origin of a tile's TMA load and width of its TMA store in 16-byte units: a load whose
    // rows start 8 bytes into a 16-byte unit faults (measured: "illegal instruction")
    static constexpr int HLQ = 16 / (int)sizeof(T);
    static constexpr int HLX = (R + HLQ - 1) / HLQ * HLQ;
    static constexpr int VX = (BX - HLX - R) / HLQ * HLQ;
    static constexpr int VY = BY - 2 * R;
    static constexpr int STAGES = 4;
    static constexpr int NSLOT = 2 * R + 1;
    static constexpr int PLANE = BX * BY;
    static constexpr size_t PLANE_BYTES = (size_t)PLANE * sizeof(T);
    static constexpr int STAGE = VX * VY;
    static constexpr size_t STAGE_BYTES = ((size_t)STAGE * sizeof(T) + 127) / 128 * 128;
    static constexpr size_t SMEM = PLANE_BYTES * STAGES + 2 * STAGE_BYTES + 8 * STAGES + 128;
};

template <class T, int R, int PXV_, int PY_, bool LINKED = false>
__global__ void __launch_bounds__(T3B<T, R, PXV_, PY_>::THREADS, 1)
sweep_tma3db_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap omap,
                    T* __restrict__ out, Grid g, long long ob, long long oe, int ntx, int nty, int chunk_planes,
                    T avg, const SlabLinkDev link) {
    using C = T3B<T, R, PXV_, PY_>;
    constexpr int PXV = C::PXV, PY = C::PY, S = C::STAGES, NQ = C::NSLOT;
    constexpr int Q = 16 / (int)sizeof(T);  // the TMA unit clips stores in 16-byte units
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* base = smem_raw + ((128 - (smem_u32(smem_raw) & 127)) & 127);
    T* ring = reinterpret_cast<T*>(base);
    unsigned char* stage = reinterpret_cast<unsigned char*>(ring + (size_t)S * C::PLANE);
    uint64_t* bars = reinterpret_cast<uint64_t*>(stage + 2 * C::STAGE_BYTES);

    const int tid = threadIdx.x;
    const int tx = tid & 31, ty = tid >> 5;
    const int tiles = ntx * nty;
    const int tile = (int)(blockIdx.x % (unsigned)tiles);
    const int tix = tile % ntx, tiy = tile / ntx;
    int chunk = (int)(blockIdx.x / (unsigned)tiles);
    if (LINKED && link.link_hi) chunk = chunk == 0 ? link.nchunks - 1 : chunk - 1;  // the last chunk first
    const bool is_lo = LINKED && link.link_lo && chunk == 0;
    const bool is_hi = LINKED && link.link_hi && chunk == link.nchunks - 1;
    const long long zb = LINKED ? ob + (oe - ob) * chunk / link.nchunks : ob + (long long)chunk * chunk_planes;
    const long long ze = LINKED ? ob + (oe - ob) * (chunk + 1) / link.nchunks
                                : (zb + chunk_planes < oe ? zb + chunk_planes : oe);
    const long long gx0 = (long long)tix * C::VX - C::HLX;
    const long long gy0 = (long long)tiy * C::VY - R;
    const long long z_first = zb - R;
    const int nin = (int)(ze - zb) + 2 * R;  // planes streamed; plane j-R is emitted when plane j arrives

    if (tid == 0) {
        for (int s = 0; s < S; ++s) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (LINKED && (is_lo || is_hi) && link.wait_value != 0) {
            if (is_lo) link_wait(link.wait_lo, link.wait_value, link.err_flag);
            if (is_hi) link_wait(link.wait_hi, link.wait_value, link.err_flag);
            asm volatile("fence.proxy.async.global;" ::: "memory");  // TMA reads go through the async proxy
        }
    }
    __syncthreads();
    const int cx = (int)(gx0 + g.ox), cy = (int)(gy0 + g.oy), cz0 = (int)(z_first + g.oz);
    if (tid == 0) {
        for (int s = 0; s < S && s < nin; ++s) {
            mbar_expect_tx(&bars[s], (uint32_t)C::PLANE_BYTES);
            tma_load_3d(ring + (size_t)s * C::PLANE, &tmap, &bars[s], cx, cy, cz0 + s);
        }
    }
    const int lx = tx * PXV, ly = ty * PY;
    const long long gx = gx0 + lx, gy = gy0 + ly;
    unsigned s_rows = 0;
#pragma unroll
    for (int i = 0; i < PY; ++i)
        if (ly + i >= R && ly + i < R + C::VY) s_rows |= 1u << i;
    if (!(lx >= C::HLX && lx < C::HLX + C::VX)) s_rows = 0;
    const int s_off = (ly - R) * C::VX + (lx - C::HLX);
    // points of the rows' ragged end (past the last full 16 bytes of a row) are stored by their threads
    unsigned x_tail = 0;
#pragma unroll
    for (int k = 0; k < PXV; ++k)
        if (gx + k >= g.n[0] / Q * Q && gx + k < g.n[0]) x_tail |= 1u << k;
    // a warp owns PY rows of the tile's whole width: when none of them is stored (rows of the tile's
    // halo ring, rows past the end of the grid) the warp skips the plane's arithmetic
    bool warp_on = false;
#pragma unroll
    for (int i = 0; i < PY; ++i)
        if (ly + i >= R && ly + i < R + C::VY && gy + i < g.n[1]) warp_on = true;
    const int ox0 = tix * C::VX, oy0 = tiy * C::VY, oz0 = (int)(z_first + g.oz);
    // rows ly-R .. ly+PY+R-1 of the tile, clamped into it (a clamped row only ever feeds points
    // of the tile's own halo ring, which are not stored): only the first and the last thread
    // row clamp, so two offsets describe it
    const int row_lo = ly - R < 0 ? R - ly : 0;                          // rows below this index read row 0
    const int row_hi = ly + PY + R > C::BY ? C::BY - 1 - (ly - R) : PY + 2 * R;  // rows above read row BY-1

    // acc[q]: running sum of output plane (j - R + q) while plane j is consumed
    T acc[NQ][PY][PXV];
#pragma unroll
    for (int q = 0; q < NQ; ++q)
#pragma unroll
        for (int i = 0; i < PY; ++i)
#pragma unroll
            for (int m = 0; m < PXV; ++m) acc[q][i][m] = T(0);

#pragma unroll 1
    for (int j = 0; j < nin; ++j) {
        const int st = j & (S - 1);
        const int pe = j - R;  // local index of the plane this step finishes (slot 0)
        const bool storing = pe >= R && pe < nin - R;
        if (warp_on) {  // (warp-uniform)
            mbar_wait(&bars[st], (uint32_t)(j / S) & 1u);
            const T* src = ring + (size_t)st * C::PLANE + (ly - R) * C::BX + lx;
#pragma unroll
            for (int a = 0; a < PY + 2 * R; ++a) {
                const int ra = a < row_lo ? row_lo : (a > row_hi ? row_hi : a);
                T w[PXV + 2 * R];
                {
                    T c[PXV];
                    VecN<T, PXV>::ld(src + ra * C::BX, c);
#pragma unroll
                    for (int m = 0; m < PXV; ++m) w[R + m] = c[m];
                    auto fill = [&](auto self, auto Kc) {
                        constexpr int k = decltype(Kc)::value;
                        if constexpr (k <= R) {
                            w[R - k] = t3r_at<-k, T, PXV>(c);                    // x - k of element 0
                            w[R + PXV - 1 + k] = t3r_at<PXV - 1 + k, T, PXV>(c);  // x + k of element PXV-1
                            self(self, IC<k + 1>{});
                        }
                    };
                    fill(fill, IC<1>{});
                }
                // this row is row dy = a - R - i of owned row i
#pragma unroll
                for (int i = 0; i < PY; ++i) {
                    if (a - i < 0 || a - i > 2 * R) continue;
#pragma unroll
                    for (int m = 0; m < PXV; ++m)
#pragma unroll
                        for (int dx = 0; dx <= 2 * R; ++dx) {
                            const T v = w[m + dx];
#pragma unroll
                            for (int q = 0; q < NQ; ++q) acc[q][i][m] = add_rn(acc[q][i][m], v);
                        }
                }
            }
            T* sp = reinterpret_cast<T*>(stage + (size_t)(pe & 1) * C::STAGE_BYTES) + s_off;
#pragma unroll
            for (int i = 0; i < PY; ++i) {
                T o[PXV];
#pragma unroll
                for (int m = 0; m < PXV; ++m) {
                    o[m] = mul_rn(acc[0][i][m], avg);
#pragma unroll
                    for (int q = 0; q + 1 < NQ; ++q) acc[q][i][m] = acc[q + 1][i][m];
                    acc[NQ - 1][i][m] = T(0);
                }
                if (storing && (s_rows & (1u << i))) {
                    VecN<T, PXV>::st(sp + i * C::VX, o);
                    if (x_tail && gy + i < g.n[1]) {  // the ragged end of the rows (tma3d.cuh)
                        T* gp = out + g.idx(gx, gy + i, z_first + pe);
#pragma unroll
                        for (int k = 0; k < PXV; ++k)
                            if (x_tail & (1u << k)) gp[k] = o[k];
                    }
                }
            }
        }  // warp_on
        if (storing) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (tid == 0) tma_store_wait_read();
        __syncthreads();
        if (tid == 0) {
            if (storing && g.n[0] >= Q)
                tma_store_3d(&omap, stage + (size_t)(pe & 1) * C::STAGE_BYTES, ox0, oy0, oz0 + pe);
            if (j + S < nin) {
                mbar_expect_tx(&bars[st], (uint32_t)C::PLANE_BYTES);
                tma_load_3d(ring + (size_t)st * C::PLANE, &tmap, &bars[st], cx, cy, cz0 + j + S);
            }
        }
    }
    if (tid == 0) tma_store_wait_all();
    if (LINKED && (is_lo || is_hi))
        t3r_push_faces<T, PXV, PY>(out, g, link, is_lo, is_hi, gx, gy, s_rows, tid == 0);
}

template <class C>
static T3Plan t3b_plan(long long nx, long long ny, long long planes, int sm_count, int prefer_planes = 0) {
    return t3_plan_core(nx, ny, planes, sm_count, C::VX, C::VY, C::NSLOT - 1, 16 + C::PY, prefer_planes);
}

}  // namespace sb200
