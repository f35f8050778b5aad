// util_kernels.cuh — grid initialisation and checksum kernels (not on the timed path).
#pragma once
#include "common.cuh"

namespace sb200 {

// Initial condition, generated on the device so multi-GB slabs never cross PCIe.
// "ref": reference Stencil::generate_initialized_matrix (src/stencil/stencil.cpp:190-207)
//   — interior 0, the two x faces 1 over the full padded extent of the other axes,
//   every other halo cell 0.  "random": interior = (mix64(seed, global padded linear
//   index) >> 11) * 2^-53 cast to T; halo as "ref".
// gn = global interior extents; off = this slab's offset along the outermost axis.
template <class T>
__global__ void fill_kernel(T* __restrict__ grid, Grid g, int random, uint64_t seed,
                            long long gn0, long long gn1, long long gn2, long long off) {
    const long long total = g.elems();
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int r = g.r;
    const long long h1 = g.ndim > 1 ? r : 0, h2 = g.ndim > 2 ? r : 0;
    const long long gp0 = gn0 + 2 * r, gp1 = gn1 + 2 * h1;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long px = i % g.pitch_x;
        const long long py = (i / g.pitch_x) % g.pitch_y;
        const long long pz = i / (g.pitch_x * g.pitch_y);
        long long x = px - g.ox, y = py - g.oy, z = pz - g.oz;
        if (g.ndim == 3) z += off;
        if (g.ndim == 2) y += off;
        T v = T(0);
        const bool in_padded = x >= -r && x < gn0 + r && y >= -h1 && y < gn1 + h1 && z >= -h2 &&
                               z < gn2 + h2;
        if (in_padded) {
            const bool interior = x >= 0 && x < gn0 && y >= 0 && y < gn1 && z >= 0 && z < gn2;
            if (!interior) {
                if (x < 0 || x >= gn0) v = T(1);
            } else if (random) {
                const uint64_t lin =
                    ((uint64_t)(z + h2) * (uint64_t)gp1 + (uint64_t)(y + h1)) * (uint64_t)gp0 +
                    (uint64_t)(x + r);
                v = (T)((double)(mix64_seeded(seed, lin) >> 11) * 0x1.0p-53);
            }
        }
        grid[i] = v;
    }
}

// Dense host-layout image (row pitch = hx, as uploaded with one contiguous copy) -> the
// pitched device layout; padding columns are zeroed so no uninitialised bits are ever read.
template <class T>
__global__ void unpack_dense_kernel(const T* __restrict__ dense, T* __restrict__ a, Grid g,
                                    long long hx) {
    const long long rows = g.pitch_y * g.planes;
    const long long x0 = g.ox - g.r;  // pitched column of dense column 0
    for (long long row = blockIdx.y; row < rows; row += gridDim.y) {
        const T* src = dense + row * hx;
        T* da = a + row * g.pitch_x;
        for (long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x; px < g.pitch_x;
             px += (long long)gridDim.x * blockDim.x) {
            const long long dx = px - x0;
            const T v = (dx >= 0 && dx < hx) ? src[dx] : T(0);
            da[px] = v;
        }
    }
}

// Same for the rows [row_begin, row_end) only, written into BOTH device grids (the two
// reference matrices carry the same halo, stencil.cpp:190-207): the unpack step of the
// pipelined host path, one call per uploaded chunk.
template <class T>
__global__ void unpack_dense2_kernel(const T* __restrict__ dense, T* __restrict__ a, T* __restrict__ b,
                                     Grid g, long long hx, long long row_begin, long long row_end) {
    const long long x0 = g.ox - g.r;
    for (long long row = row_begin + blockIdx.y; row < row_end; row += gridDim.y) {
        const T* src = dense + row * hx;
        T* da = a + row * g.pitch_x;
        T* db = b + row * g.pitch_x;
        for (long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x; px < g.pitch_x;
             px += (long long)gridDim.x * blockDim.x) {
            const long long dx = px - x0;
            const T v = (dx >= 0 && dx < hx) ? src[dx] : T(0);
            da[px] = v;
            db[px] = v;
        }
    }
}

template <class T> struct Bits;
template <> struct Bits<float> {
    static __device__ uint64_t of(float v) { return (uint64_t)__float_as_uint(v); }
};
template <> struct Bits<double> {
    static __device__ uint64_t of(double v) { return (uint64_t)__double_as_longlong(v); }
};

// acc += sum over slab-interior points of mix64(bits(v) ^ mix64(global interior index)).
template <class T>
__global__ void checksum_kernel(const T* __restrict__ grid, Grid g, long long gn0, long long gn1,
                                long long off, unsigned long long* acc) {
    const long long total = g.n[0] * g.n[1] * g.n[2];
    const long long stride = (long long)gridDim.x * blockDim.x;
    uint64_t local = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long x = i % g.n[0];
        const long long y = (i / g.n[0]) % g.n[1];
        const long long z = i / (g.n[0] * g.n[1]);
        long long gy = y, gz = z;
        if (g.ndim == 3) gz += off;
        if (g.ndim == 2) gy += off;
        const uint64_t lin = ((uint64_t)gz * (uint64_t)gn1 + (uint64_t)gy) * (uint64_t)gn0 + (uint64_t)x;
        local += mix64(Bits<T>::of(grid[g.idx(x, y, z)]) ^ mix64(lin));
    }
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(acc, (unsigned long long)local);
}

}  // namespace sb200
