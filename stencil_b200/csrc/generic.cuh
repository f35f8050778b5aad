// generic.cuh — one time step per launch, any ndim / shape / radius.
//
// The correctness workhorse: one thread per output point, neighbours fetched through
// L1/L2 (read-only path), summation in exactly the oracle's order
// (reference src/stencil/stencil.cpp:103-125 for the 2-D star; SURVEY.md §8(c) for the
// rest), so results are bit-identical to the CPU loop.  Radius is a template parameter
// (1..4, cf. the reference's compile-time-radius idea in
// stencil_dma_static_unroll.cpp:502) with a runtime-radius instantiation (RT = 0).
#pragma once
#include "common.cuh"

namespace sb200 {

template <class T, int NDIM, int SHAPE, int RT>
__global__ void __launch_bounds__(256)
sweep_generic_kernel(const T* __restrict__ in, T* __restrict__ out, Grid g, long long ob,
                     long long oe, T avg, int r_rt) {
    const int r = RT > 0 ? RT : r_rt;
    const long long x = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= g.n[0]) return;
    // rows = (outermost range) x (middle axes)
    const long long ny = NDIM == 3 ? g.n[1] : 1;
    const long long nrows = NDIM == 1 ? 1 : (oe - ob) * ny;
    const long long sy = g.pitch_x, sz = g.pitch_x * g.pitch_y;
    for (long long row = blockIdx.y; row < nrows; row += gridDim.y) {
        long long y = 0, z = 0;
        if (NDIM == 2) y = ob + row;
        if (NDIM == 3) {
            z = ob + row / ny;
            y = row % ny;
        }
        const long long c = g.idx(x, y, z);
        const T* p = in + c;
        T sum = T(0);
        if (SHAPE == STAR) {
#pragma unroll
            for (int a = 0; a < NDIM; ++a) {
                const long long st = a == 0 ? 1 : (a == 1 ? sy : sz);
#pragma unroll(RT > 0 ? RT : 1)
                for (int k = r; k >= 1; --k) sum = add_rn(sum, __ldg(p - k * st));
#pragma unroll(RT > 0 ? RT : 1)
                for (int k = 1; k <= r; ++k) sum = add_rn(sum, __ldg(p + k * st));
            }
        } else {
            const int rz = NDIM > 2 ? r : 0, ry = NDIM > 1 ? r : 0;
            for (int dz = -rz; dz <= rz; ++dz)
                for (int dy = -ry; dy <= ry; ++dy)
#pragma unroll(RT > 0 ? 2 * RT + 1 : 1)
                    for (int dx = -r; dx <= r; ++dx)
                        sum = add_rn(sum, __ldg(p + dz * sz + dy * sy + dx));
        }
        out[c] = mul_rn(sum, avg);
    }
}

// ---- vectorised variant (compile-time radius 1..4) -------------------------------------
// Each thread owns 4 consecutive x (one 128-bit vector of floats, two of doubles) of one row.
// All loads are aligned 128-bit loads through the read-only path: the x window comes from
// the 2Q+1 vectors around the thread's own, y/z neighbours from one vector per neighbour
// row; a block is 32 x 8 threads (128 or 64 x by 8 consecutive rows), so the +-r neighbour
// rows are shared through L1 and the +-r planes through L2, and HBM sees each point about
// once.  The summation order per point is unchanged (oracle order), so the result is
// bit-identical to the scalar kernel above.
template <class T> struct GVec;
template <> struct GVec<float> {
    static constexpr int V = 4;
    static __device__ __forceinline__ void ld(const float* p, float* v) {
        const float4 f = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    }
    static __device__ __forceinline__ void st(float* p, const float* v) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <> struct GVec<double> {
    static constexpr int V = 4;  // two 128-bit vectors per thread: as many points as the fp32 path
    static __device__ __forceinline__ void ld(const double* p, double* v) {
        const double2 a = __ldg(reinterpret_cast<const double2*>(p));
        const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
    }
    static __device__ __forceinline__ void st(double* p, const double* v) {
        reinterpret_cast<double2*>(p)[0] = make_double2(v[0], v[1]);
        reinterpret_cast<double2*>(p)[1] = make_double2(v[2], v[3]);
    }
};

template <class T, int NDIM, int SHAPE, int R>
__global__ void __launch_bounds__(256)
sweep_vec_kernel(const T* __restrict__ in, T* __restrict__ out, Grid g, long long ob, long long oe,
                 T avg) {
    constexpr int V = GVec<T>::V;
    constexpr int Q = (R + V - 1) / V;  // extra vectors per side of the x window
    constexpr int W = V * (2 * Q + 1);
    constexpr int C = Q * V;            // window index of the thread's first point
    const long long xv = ((long long)blockIdx.x * 32 + threadIdx.x) * V;
    if (xv >= g.n[0]) return;
    const long long ny = NDIM == 3 ? g.n[1] : 1;
    const long long nrows = NDIM == 1 ? 1 : (oe - ob) * ny;
    const long long sy = g.pitch_x, sz = g.pitch_x * g.pitch_y;
    // vectors right of the row's allocation are not loaded (they only feed points >= n)
    const long long x_room = g.pitch_x - g.ox - xv;
    auto load_window = [&](const T* q, T (&w)[W]) {
#pragma unroll
        for (int v = -Q; v <= Q; ++v) {
            if (v <= 0 || (long long)(v + 1) * V <= x_room) {
                GVec<T>::ld(q + v * V, &w[(v + Q) * V]);
            } else {
#pragma unroll
                for (int e = 0; e < V; ++e) w[(v + Q) * V + e] = T(0);
            }
        }
    };
    const bool full = xv + V <= g.n[0];
    for (long long row = (long long)blockIdx.y * blockDim.y + threadIdx.y; row < nrows;
         row += (long long)gridDim.y * blockDim.y) {
        long long y = 0, z = 0;
        if (NDIM == 2) y = ob + row;
        if (NDIM == 3) {
            z = ob + row / ny;
            y = row % ny;
        }
        const long long c = g.idx(xv, y, z);
        const T* p = in + c;
        T sum[V];
#pragma unroll
        for (int e = 0; e < V; ++e) sum[e] = T(0);
        if (SHAPE == STAR) {
            T w[W];
            load_window(p, w);
#pragma unroll
            for (int e = 0; e < V; ++e) {
#pragma unroll
                for (int k = R; k >= 1; --k) sum[e] = add_rn(sum[e], w[C + e - k]);
#pragma unroll
                for (int k = 1; k <= R; ++k) sum[e] = add_rn(sum[e], w[C + e + k]);
            }
#pragma unroll
            for (int a = 1; a < NDIM; ++a) {
                const long long st = a == 1 ? sy : sz;
                T nb[2 * R][V];
#pragma unroll
                for (int k = R; k >= 1; --k) GVec<T>::ld(p - k * st, nb[R - k]);
#pragma unroll
                for (int k = 1; k <= R; ++k) GVec<T>::ld(p + k * st, nb[R + k - 1]);
#pragma unroll
                for (int j = 0; j < 2 * R; ++j)
#pragma unroll
                    for (int e = 0; e < V; ++e) sum[e] = add_rn(sum[e], nb[j][e]);
            }
        } else {
            constexpr int RZ = NDIM > 2 ? R : 0, RY = NDIM > 1 ? R : 0;
#pragma unroll
            for (int dz = -RZ; dz <= RZ; ++dz)
#pragma unroll
                for (int dy = -RY; dy <= RY; ++dy) {
                    T w[W];
                    load_window(p + dz * sz + dy * sy, w);
#pragma unroll
                    for (int e = 0; e < V; ++e)
#pragma unroll
                        for (int dx = -R; dx <= R; ++dx) sum[e] = add_rn(sum[e], w[C + e + dx]);
                }
        }
#pragma unroll
        for (int e = 0; e < V; ++e) sum[e] = mul_rn(sum[e], avg);
        if (full) {
            GVec<T>::st(out + c, sum);
        } else {
#pragma unroll
            for (int e = 0; e < V; ++e)
                if (xv + e < g.n[0]) out[c + e] = sum[e];
        }
    }
}

template <class T, int NDIM, int SHAPE, int R>
static void launch_vec(const T* in, T* out, const Grid& g, long long ob, long long oe, T avg,
                       cudaStream_t st) {
    constexpr int V = GVec<T>::V;
    const long long ny = NDIM == 3 ? g.n[1] : 1;
    const long long nrows = NDIM == 1 ? 1 : (oe - ob) * ny;
    dim3 block(32, NDIM == 1 ? 1 : 8);
    const long long gy = (nrows + block.y - 1) / block.y;
    dim3 grid((unsigned)((g.n[0] + 32 * V - 1) / (32 * V)), (unsigned)(gy < 65535 ? gy : 65535));
    sweep_vec_kernel<T, NDIM, SHAPE, R><<<grid, block, 0, st>>>(in, out, g, ob, oe, avg);
}

template <class T, int NDIM, int SHAPE>
static cudaError_t launch_generic_r(const T* in, T* out, const Grid& g, long long ob,
                                    long long oe, cudaStream_t st) {
    if (oe <= ob || g.n[0] <= 0) return cudaSuccess;
    const T avg = Avg<T>::value(SHAPE, NDIM, g.r);
    const long long ny = NDIM == 3 ? g.n[1] : 1;
    const long long nrows = NDIM == 1 ? 1 : (oe - ob) * ny;
    dim3 block(256);
    dim3 grid((unsigned)((g.n[0] + 255) / 256), (unsigned)(nrows < 65535 ? nrows : 65535));
    switch (g.r) {
        case 1: launch_vec<T, NDIM, SHAPE, 1>(in, out, g, ob, oe, avg, st); break;
        case 2: launch_vec<T, NDIM, SHAPE, 2>(in, out, g, ob, oe, avg, st); break;
        case 3: launch_vec<T, NDIM, SHAPE, 3>(in, out, g, ob, oe, avg, st); break;
        case 4: launch_vec<T, NDIM, SHAPE, 4>(in, out, g, ob, oe, avg, st); break;
        default: sweep_generic_kernel<T, NDIM, SHAPE, 0><<<grid, block, 0, st>>>(in, out, g, ob, oe, avg, g.r); break;
    }
    return cudaGetLastError();
}

// One step in -> out over outermost-axis range [ob, oe) (slab-interior coordinates).
template <class T>
static cudaError_t launch_generic(const T* in, T* out, const Grid& g, int shape, long long ob,
                                  long long oe, cudaStream_t st) {
    if (g.ndim == 1) {
        ob = 0;
        oe = 1;
    }
#define SB200_GEN(ND, SH) return launch_generic_r<T, ND, SH>(in, out, g, ob, oe, st)
    if (shape == STAR) {
        if (g.ndim == 1) SB200_GEN(1, STAR);
        if (g.ndim == 2) SB200_GEN(2, STAR);
        SB200_GEN(3, STAR);
    } else {
        if (g.ndim == 1) SB200_GEN(1, BOX);
        if (g.ndim == 2) SB200_GEN(2, BOX);
        SB200_GEN(3, BOX);
    }
#undef SB200_GEN
}

}  // namespace sb200
