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
        case 1: sweep_generic_kernel<T, NDIM, SHAPE, 1><<<grid, block, 0, st>>>(in, out, g, ob, oe, avg, g.r); break;
        case 2: sweep_generic_kernel<T, NDIM, SHAPE, 2><<<grid, block, 0, st>>>(in, out, g, ob, oe, avg, g.r); break;
        case 3: sweep_generic_kernel<T, NDIM, SHAPE, 3><<<grid, block, 0, st>>>(in, out, g, ob, oe, avg, g.r); break;
        case 4: sweep_generic_kernel<T, NDIM, SHAPE, 4><<<grid, block, 0, st>>>(in, out, g, ob, oe, avg, g.r); break;
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
