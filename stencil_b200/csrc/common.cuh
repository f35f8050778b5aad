// common.cuh — shared device/host definitions for the sm_100a stencil kernels.
//
// Device grid layout (chosen here, not the reference's): like the reference's padded
// row-major array (include/stencil/boundary_matrix.hpp:54-59,115-123) but with the
// first interior x at a 128-byte boundary and a 128-byte-multiple row pitch, so warps
// issue aligned 128-bit loads/stores and TMA boxes start on 16-byte addresses.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace sb200 {

// BOXSEP: the box stencil with separable partial sums (x, then y, then z: 8 operations per
// point instead of 28), an internal kernel variant chosen under SB200_FLAG_REASSOCIATE; agrees
// with the reference order within the north-star tolerance, not bit for bit.
enum : int { STAR = 0, BOX = 1, BOXSEP = 2 };
enum : int { F32 = 0, F64 = 1 };

// Geometry of one device grid (one slab).  Coordinates passed to idx() are interior
// coordinates: 0..n-1 inside, negative / >= n in the halo or ghost planes.
struct Grid {
    long long pitch_x;  // elements per row
    long long pitch_y;  // rows per plane
    long long planes;   // planes allocated
    int ox, oy, oz;     // position of interior (0,0,0)
    long long n[3];     // interior extents of THIS slab (x, y, z)
    int ndim;
    int r;      // stencil radius = Dirichlet halo width
    int ghost;  // extra planes/rows on each side of the outermost axis (slabs)
    __host__ __device__ long long idx(long long x, long long y, long long z) const {
        return ((z + oz) * pitch_y + (y + oy)) * pitch_x + (x + ox);
    }
    __host__ __device__ long long elems() const { return pitch_x * pitch_y * planes; }
};

// Update domain along the outermost axis, in slab-interior coordinates.  Points of
// the outermost axis outside [lo, hi) are never updated at ANY time level: they are
// the global Dirichlet halo (value passes through unchanged).  For a slab whose low
// face is internal, lo = -ghost (ghost planes are ordinary updatable points holding
// neighbour data); same for hi.
struct Domain {
    long long lo, hi;
};

// What the planners need to know about the device and the caller's preferences.
// prefer_chunk > 0: rows / planes of the outermost axis one CTA (tma3d) or warp (warp2d) should
// stream — the reference's -b/--block-size (the side of the tile one CPE owns,
// program_options.cpp:16-18) reinterpreted for this backend (SB200_FLAG_BLOCK_TILE).
struct Tuning {
    int sm_count;
    int prefer_chunk;
};

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t mix64_seeded(uint64_t seed, uint64_t idx) {
    return mix64(seed + (idx + 1) * 0x9E3779B97F4A7C15ull);
}

template <class T>
struct Avg {
    // star: 1 / T(2*d*r)  (reference stencil.cpp:85-86);  box: 1 / T((2r+1)^d).
    static __host__ __device__ T value(int shape, int ndim, int r) {
        if (shape == STAR) return T(1) / T(2 * ndim * r);
        int c = 1;
        for (int a = 0; a < ndim; ++a) c *= (2 * r + 1);
        return T(1) / T(c);
    }
};

// IEEE add/mul that the compiler may not contract into an FMA or reassociate.  The
// oracle's order is a chain of adds followed by ONE multiply (stencil.cpp:103-125), so
// there is no legal FMA; these intrinsics make that explicit (no -use_fast_math, no ftz).
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }

}  // namespace sb200
