// tma3dr.cuh — 3-D star stencils of radius R = 2..4 (13-, 19-, 25-point), one time step per launch,
// in the tma3d framework (what the reference's `-r` flag asks of a 3-D grid: program_options.cpp:22-23;
// its general-radius tile kernel is neighbors_impl, stencil_dma.cpp:571-1099).
//
// Same machinery as tma3d.cuh: a CTA owns a 64(32*PXV) x 64 xy tile with an R-wide halo and
// streams along z; planes arrive through a 4-slot TMA ring (OOB zero-filled), every thread owns
// PXV x 4 points, x neighbours come from adjacent lanes by shuffle (R per side), y neighbours
// from the staged plane (R rows above and below the thread's rows), and the z direction never
// touches shared memory: the last R planes' centre values and the R pending partial sums live
// in registers, in R-slot rings indexed by the plane's phase (plane mod R, a compile-time
// constant because the plane loop is unrolled by R).  The finished plane leaves through the
// staging tile + TMA store of tma3d.cuh.
//
// Per-point order is the oracle's: x low side far->near, x high side near->far,
// the same for y, then z low far->near (from the history ring when the plane arrives) and z high
// near->far (one add per later plane); one multiply by avg = 1/(6R).  Bit-identical.
//
// One step per launch (K = 1): 2R plane-sized arrays of state per level leave no room for a
// second level at 512 threads x 128 registers — at R >= 3 not even for four rows per thread, so
// those radii use 64 x 32 tiles (two rows per thread).  The win over the one-thread-per-vector
// generic kernel is the memory system (TMA tiles, z re-use in registers), not fusion.
#pragma once
#include "tma3d.cuh"

namespace sb200 {

template <class T, int R> struct T3R {
    static constexpr int PXV = 16 / (int)sizeof(T);
    static constexpr int PY = R == 2 ? 4 : 2;  // rows per thread: the register budget (2R state arrays)
    static constexpr int TXN = 32, TYN = 16;
    static constexpr int THREADS = TXN * TYN;
    static constexpr int BX = TXN * PXV, BY = TYN * PY;
    static constexpr int HLX = (R + PXV - 1) / PXV * PXV;
    static constexpr int VX = (BX - HLX - R) / PXV * PXV;
    static constexpr int VY = BY - 2 * R;
    static constexpr int STAGES = 4;
    static constexpr int PLANE = BX * BY;
    static constexpr size_t PLANE_BYTES = (size_t)PLANE * sizeof(T);
    static constexpr int STAGE = VX * VY;
    static constexpr size_t STAGE_BYTES = ((size_t)STAGE * sizeof(T) + 127) / 128 * 128;
    static constexpr size_t SMEM = PLANE_BYTES * STAGES + 2 * STAGE_BYTES + 8 * STAGES + 128;
};

// value at x offset E (in elements, relative to this thread's first point) of a row whose own
// PXV points are v[]: from a neighbouring lane when E falls outside [0, PXV)
template <int E, class T, int PXV>
__device__ __forceinline__ T t3r_at(const T (&v)[PXV]) {
    constexpr int LANE = E >= 0 ? E / PXV : -((-E + PXV - 1) / PXV);
    constexpr int EL = E - LANE * PXV;
    if constexpr (LANE == 0) return v[EL];
    else if constexpr (LANE < 0) return __shfl_up_sync(0xffffffffu, v[EL], -LANE);
    else return __shfl_down_sync(0xffffffffu, v[EL], LANE);
}

// Face push of a LINKED launch (as in tma3d.cuh), shared by the radius-R star and box kernels:
// this CTA's TMA stores are complete (thread 0 waited; then the barrier and the proxy fence);
// every thread copies its part (PY rows of PXV points at gx, gy; s_rows = the rows it stores) of
// the face planes into the neighbour's ghost planes, then thread 0 publishes them and counts
// the CTA in.
template <class T, int PXV, int PY>
__device__ __forceinline__ void t3r_push_faces(T* __restrict__ out, const Grid& g, const SlabLinkDev& link, bool is_lo,
                                               bool is_hi, long long gx, long long gy, unsigned s_rows, bool leader) {
    __syncthreads();
    asm volatile("fence.proxy.async.global;" ::: "memory");
    const long long plane_stride = g.pitch_x * g.pitch_y;
    T* home = out + g.idx(gx, gy, 0);
    const bool col_in = gx < g.n[0], col_full = gx + PXV <= g.n[0];
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
                if ((s_rows & (1u << i)) && col_in && gy + i < g.n[1]) {
                    if (col_full) {
                        T v[PXV];
                        VecN<T, PXV>::ldcg(src + i * g.pitch_x, v);
                        VecN<T, PXV>::st(src + rdelta + i * g.pitch_x, v);
                    } else {
#pragma unroll
                        for (int k = 0; k < PXV; ++k)
                            if (gx + k < g.n[0]) src[rdelta + i * g.pitch_x + k] = __ldcg(src + i * g.pitch_x + k);
                    }
                }
            }
        }
    }
    __syncthreads();
    if (leader) {
        __threadfence_system();
        if (is_lo) red_release_sys_add_u32(link.signal_lo, 1u);
        if (is_hi) red_release_sys_add_u32(link.signal_hi, 1u);
    }
}

// LINKED: the slab link of tma3d.cuh / slablink.cuh (faces are g = R planes; the CTAs of the first
// / last z chunk own them, are dispatched first, acquire the neighbours' arrival counter before
// their first TMA load and push their face planes after the z stream).
template <class T, int R, bool LINKED = false>
__global__ void __launch_bounds__(T3R<T, R>::THREADS, 1)
sweep_tma3dr_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap omap,
                    T* __restrict__ out, Grid g, long long ob, long long oe, int ntx, int nty, int chunk_planes,
                    T avg, const SlabLinkDev link) {
    using C = T3R<T, R>;
    constexpr int PXV = C::PXV, PY = C::PY, S = C::STAGES;
    static_assert(R >= 2 && R <= 4, "radius 2..4");
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
    const int nin = (int)(ze - zb) + 2 * R;  // planes streamed; plane a-R is emitted when plane a arrives

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
    const int toff = ly * C::BX + lx;
    // rows above / below the thread's rows, clamped into the tile (a clamped row only ever
    // feeds points of the tile's own halo ring, which are not stored)
    int up_off[R], dn_off[R];
#pragma unroll
    for (int k = 0; k < R; ++k) {
        up_off[k] = (ly - 1 - k >= 0 ? -(1 + k) : -ly) * C::BX;
        dn_off[k] = (ly + PY + k < C::BY ? PY + k : C::BY - 1 - ly) * C::BX;
    }
    unsigned s_rows = 0;
#pragma unroll
    for (int i = 0; i < PY; ++i)
        if (ly + i >= R && ly + i < R + C::VY) s_rows |= 1u << i;
    if (!(lx >= C::HLX && lx < C::HLX + C::VX)) s_rows = 0;
    const int s_off = (ly - R) * C::VX + (lx - C::HLX);
    const int x_tail = gx == g.n[0] / PXV * PXV ? (int)(g.n[0] - gx) : 0;
    const int ox0 = tix * C::VX, oy0 = tiy * C::VY, oz0 = (int)(z_first + g.oz);

    // hist[s]: centre values of the newest plane with phase s; pend[s]: its partial sum (in-plane
    // part + z low side + the z high neighbours that have arrived so far)
    T hist[R][PY][PXV], pend[R][PY][PXV];
#pragma unroll
    for (int s = 0; s < R; ++s)
#pragma unroll
        for (int i = 0; i < PY; ++i)
#pragma unroll
            for (int k = 0; k < PXV; ++k) hist[s][i][k] = pend[s][i][k] = T(0);

    auto step = [&](auto Pc, int j) {
        constexpr int P = decltype(Pc)::value;  // j mod R
        const int st = j & (S - 1);
        if (j < nin) mbar_wait(&bars[st], (uint32_t)(j / S) & 1u);
        const T* src = ring + (size_t)st * C::PLANE + toff;
        // rows ly-R .. ly+PY+R-1 of the thread's columns
        T rows[PY + 2 * R][PXV];
#pragma unroll
        for (int k = 0; k < R; ++k) VecN<T, PXV>::ld(src + up_off[k], rows[R - 1 - k]);
#pragma unroll
        for (int i = 0; i < PY; ++i) VecN<T, PXV>::ld(src + i * C::BX, rows[R + i]);
#pragma unroll
        for (int k = 0; k < R; ++k) VecN<T, PXV>::ld(src + dn_off[k], rows[R + PY + k]);
        const int pe = j - R;  // local index of the plane this step finishes
        const bool storing = pe >= R && pe < nin - R;
        T* sp = reinterpret_cast<T*>(stage + (size_t)(pe & 1) * C::STAGE_BYTES) + s_off;
#pragma unroll
        for (int i = 0; i < PY; ++i) {
            const T (&c)[PXV] = rows[R + i];
            T o[PXV];
            // x window of this row: 2R values from the neighbouring lanes
            T wl[R], wr[R];
            auto fill = [&](auto self, auto Kc) {
                constexpr int k = decltype(Kc)::value;
                if constexpr (k <= R) {
                    wl[k - 1] = t3r_at<-k, T, PXV>(c);            // x - k of element 0
                    wr[k - 1] = t3r_at<PXV - 1 + k, T, PXV>(c);   // x + k of element PXV-1
                    self(self, IC<k + 1>{});
                }
            };
            fill(fill, IC<1>{});
#pragma unroll
            for (int m = 0; m < PXV; ++m) {
                auto xat = [&](int e) -> T {  // value at element offset e in [-R, PXV + R)
                    return e < 0 ? wl[-e - 1] : (e >= PXV ? wr[e - PXV] : c[e]);
                };
                T sum = T(0);
#pragma unroll
                for (int k = R; k >= 1; --k) sum = add_rn(sum, xat(m - k));
#pragma unroll
                for (int k = 1; k <= R; ++k) sum = add_rn(sum, xat(m + k));
#pragma unroll
                for (int k = R; k >= 1; --k) sum = add_rn(sum, rows[R + i - k][m]);
#pragma unroll
                for (int k = 1; k <= R; ++k) sum = add_rn(sum, rows[R + i + k][m]);
                // z low side, far -> near: planes j-R .. j-1 (phases P, P+1, .. mod R)
#pragma unroll
                for (int k = R; k >= 1; --k) sum = add_rn(sum, hist[(P + R - k) % R][i][m]);
                // this plane is the +k neighbour of plane j-k
#pragma unroll
                for (int k = 1; k <= R; ++k)
                    pend[(P + R - k) % R][i][m] = add_rn(pend[(P + R - k) % R][i][m], c[m]);
                // plane j-R (slot P) is complete; its slot takes this plane
                o[m] = mul_rn(pend[P][i][m], avg);
                pend[P][i][m] = sum;
                hist[P][i][m] = c[m];
            }
            if (storing && (s_rows & (1u << i))) {
                VecN<T, PXV>::st(sp + i * C::VX, o);
                if (x_tail && gy + i < g.n[1]) {  // the ragged end of the rows (tma3d.cuh)
                    T* gp = out + g.idx(gx, gy + i, z_first + pe);
#pragma unroll
                    for (int k = 0; k < PXV - 1; ++k)
                        if (k < x_tail) gp[k] = o[k];
                }
            }
        }
        if (storing) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (tid == 0) tma_store_wait_read();
        __syncthreads();
        if (tid == 0) {
            if (storing && g.n[0] >= PXV)
                tma_store_3d(&omap, stage + (size_t)(pe & 1) * C::STAGE_BYTES, ox0, oy0, oz0 + pe);
            if (j + S < nin) {
                mbar_expect_tx(&bars[st], (uint32_t)C::PLANE_BYTES);
                tma_load_3d(ring + (size_t)st * C::PLANE, &tmap, &bars[st], cx, cy, cz0 + j + S);
            }
        }
    };
    // the plane loop is unrolled by R so that the phase (plane mod R) is a compile-time constant;
    // steps past nin wait for nothing and store nothing
    auto steps_R = [&](auto self, auto Pc, int j) {
        constexpr int P = decltype(Pc)::value;
        if constexpr (P < R) {
            step(IC<P>{}, j + P);
            self(self, IC<P + 1>{}, j);
        }
    };
#pragma unroll 1
    for (int j = 0; j < nin; j += R) steps_R(steps_R, IC<0>{}, j);
    if (tid == 0) tma_store_wait_all();
    if (LINKED && (is_lo || is_hi))
        t3r_push_faces<T, PXV, PY>(out, g, link, is_lo, is_hi, gx, gy, s_rows, tid == 0);
}

template <class T, int R>
static T3Plan t3r_plan(long long nx, long long ny, long long planes, int sm_count, int prefer_planes = 0) {
    using C = T3R<T, R>;
    return t3_plan_core(nx, ny, planes, sm_count, C::VX, C::VY, 2 * R, R, prefer_planes);
}

}  // namespace sb200
