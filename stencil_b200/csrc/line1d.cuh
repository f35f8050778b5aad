// line1d.cuh — 1-D star / box stencils of radius 1..4 with K time steps fused per HBM round
// trip (the north star names 1-D grids; the reference's loop nest
// with one axis, stencil.cpp:94-131).
//
// A WARP owns a segment of 512 consecutive points held entirely in registers: lane l keeps
// points [128 j + 4 l, +4) for j = 0..3, so every load and store is a coalesced 128-bit
// access.  A time step is a register-to-register pass over the segment: the R neighbours
// beyond a lane's 4 points come from the adjacent lane by a rotating shuffle (lane 31 hands
// lane 0 the end of the previous 128-point chunk).  K steps run back to back in a runtime
// loop — the step body does not depend on the step index, the wrap-around garbage entering
// at the two segment ends just moves inwards R points per step — and only the middle
// 512 - 2*K*R points are stored.  No shared memory, no barrier.
//
// Per-point order is the oracle's: star = low side far->near then high side near->far,
// box = ascending offsets including the centre; one multiply by avg.  Points outside [0, n)
// (the Dirichlet halo and whatever lies beyond it) pass through unchanged at every level.
#pragma once
#include "warp2d.cuh"

namespace sb200 {

constexpr int L1D_J = 4;              // 128-point chunks per segment
constexpr int L1D_SEG = 128 * L1D_J;  // points per warp
constexpr int L1D_WARPS = 4;
constexpr int L1D_MAXK = 16;

static inline int l1d_halo_left(int K, int R) { return (K * R + 3) & ~3; }
static inline int l1d_valid(int K, int R) { return (L1D_SEG - l1d_halo_left(K, R) - K * R) & ~3; }

template <class T> __device__ __forceinline__ void l1d_load4(const T* p, T (&v)[4]);
template <> __device__ __forceinline__ void l1d_load4<float>(const float* p, float (&v)[4]) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
}
template <> __device__ __forceinline__ void l1d_load4<double>(const double* p, double (&v)[4]) {
    const double2 a = __ldg(reinterpret_cast<const double2*>(p));
    const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void l1d_store4(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void l1d_store4(double* p, const double (&v)[4]) {
    reinterpret_cast<double2*>(p)[0] = make_double2(v[0], v[1]);
    reinterpret_cast<double2*>(p)[1] = make_double2(v[2], v[3]);
}

// K time steps on a warp's register-resident segment.  EDGE: some points lie outside [0, n)
// and pass through (bit 4j+i of upd clear).
template <class T, int SHAPE, int R, bool EDGE>
__device__ __forceinline__ void l1d_steps(T (&v)[L1D_J][4], unsigned upd, int K, int lane, T avg) {
    constexpr unsigned FULL = 0xffffffffu;
    const int from_left = (lane + 31) & 31, from_right = (lane + 1) & 31;
#pragma unroll 1
    for (int s = 0; s < K; ++s) {
        T o[L1D_J][4];
#pragma unroll
        for (int j = 0; j < L1D_J; ++j) {
            T w[4 + 2 * R];
#pragma unroll
            for (int k = 1; k <= R; ++k) {
                // lane l reads its low-side neighbours from lane l-1 and its high-side ones from
                // lane l+1 (mod 32).  What lane 0 reads from lane 31 must be the tail of chunk
                // j-1, what lane 31 reads from lane 0 the head of chunk j+1 (at the two segment
                // ends: wrap-around values, never used by a point that is stored)
                const T sl = (lane == 31 && j > 0) ? v[j > 0 ? j - 1 : 0][4 - k] : v[j][4 - k];
                const T sr = (lane == 0 && j < L1D_J - 1) ? v[j < L1D_J - 1 ? j + 1 : j][k - 1] : v[j][k - 1];
                w[R - k] = __shfl_sync(FULL, sl, from_left);
                w[R + 3 + k] = __shfl_sync(FULL, sr, from_right);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) w[R + i] = v[j][i];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                T sum = T(0);
                if (SHAPE == STAR) {
#pragma unroll
                    for (int k = R; k >= 1; --k) sum = add_rn(sum, w[R + i - k]);
#pragma unroll
                    for (int k = 1; k <= R; ++k) sum = add_rn(sum, w[R + i + k]);
                } else {
#pragma unroll
                    for (int k = -R; k <= R; ++k) sum = add_rn(sum, w[R + i + k]);
                }
                o[j][i] = (!EDGE || ((upd >> (4 * j + i)) & 1u)) ? mul_rn(sum, avg) : v[j][i];
            }
        }
#pragma unroll
        for (int j = 0; j < L1D_J; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) v[j][i] = o[j][i];
    }
}

template <class T, int SHAPE, int R>
__global__ void __launch_bounds__(32 * L1D_WARPS)
sweep_line1d_kernel(const T* __restrict__ in, T* __restrict__ out, Grid g, int K, int HL, int V,
                    long long nseg, T avg) {
    constexpr unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const long long seg = (long long)blockIdx.x * L1D_WARPS + (threadIdx.x >> 5);
    if (seg >= nseg) return;
    const long long x0 = seg * V - HL;  // first point of the segment (a multiple of 4)
    T v[L1D_J][4];
    unsigned upd = 0;  // bit 4j+i: point i of chunk j lies in [0, n)
#pragma unroll
    for (int j = 0; j < L1D_J; ++j) {
        const long long gx = x0 + j * 128 + lane * 4;
        // stay inside the allocation; clamped lanes hold points nothing valid depends on
        long long gl = gx < -(long long)g.ox ? -(long long)g.ox : gx;
        if (gl + 4 > g.pitch_x - g.ox) gl = g.pitch_x - g.ox - 4;
        l1d_load4(in + g.ox + gl, v[j]);
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (gx + i >= 0 && gx + i < g.n[0]) upd |= 1u << (4 * j + i);
    }
    // interior segments (every point of the warp updatable) skip the pass-through selects,
    // which otherwise cost as many issue slots as the arithmetic
    if (__all_sync(FULL, upd == 0xffffu))
        l1d_steps<T, SHAPE, R, false>(v, upd, K, lane, avg);
    else
        l1d_steps<T, SHAPE, R, true>(v, upd, K, lane, avg);
#pragma unroll
    for (int j = 0; j < L1D_J; ++j) {
        const int c = j * 128 + lane * 4;  // segment-relative
        const long long gx = x0 + c;
        if (c < HL || c >= HL + V || gx >= g.n[0]) continue;
        T* p = out + g.ox + gx;
        if (gx + 4 <= g.n[0]) {
            l1d_store4(p, v[j]);
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (gx + i < g.n[0]) p[i] = v[j][i];
        }
    }
}

// Fused depth when the caller does not choose: the deepest whose halo K*R stays within 32
// points (measured: r=1 peaks at 16, r=4 at 8; profiles/r01e_line1d_probe.log).
static inline int l1d_default_k(int R) { return 32 / R < L1D_MAXK ? 32 / R : L1D_MAXK; }

struct L1dPlan {
    int halo_left, valid;
    long long nseg;
};
static inline L1dPlan l1d_plan(long long nx, int K, int R) {
    L1dPlan p;
    p.halo_left = l1d_halo_left(K, R);
    p.valid = l1d_valid(K, R);
    p.nseg = (nx + p.valid - 1) / p.valid;
    return p;
}

template <class T, int SHAPE, int R>
static cudaError_t launch_line1d_r(const T* in, T* out, const Grid& g, int K, cudaStream_t st) {
    const L1dPlan p = l1d_plan(g.n[0], K, R);
    const T avg = Avg<T>::value(SHAPE, 1, R);
    const unsigned blocks = (unsigned)((p.nseg + L1D_WARPS - 1) / L1D_WARPS);
    sweep_line1d_kernel<T, SHAPE, R><<<blocks, 32 * L1D_WARPS, 0, st>>>(in, out, g, K, p.halo_left, p.valid,
                                                                       p.nseg, avg);
    return cudaGetLastError();
}

// 1-D star or box, radius 1..4, K = 1..L1D_MAXK fused steps.
template <class T>
static cudaError_t launch_line1d(const T* in, T* out, const Grid& g, int shape, int K, cudaStream_t st) {
    if (K < 1 || K > L1D_MAXK || g.ndim != 1) return cudaErrorInvalidValue;
#define SB200_L1D(RR)                                                                      \
    case RR:                                                                               \
        return shape == STAR ? launch_line1d_r<T, STAR, RR>(in, out, g, K, st)             \
                             : launch_line1d_r<T, BOX, RR>(in, out, g, K, st)
    switch (g.r) {
        SB200_L1D(1); SB200_L1D(2); SB200_L1D(3); SB200_L1D(4);
        default: return cudaErrorInvalidValue;
    }
#undef SB200_L1D
}

}  // namespace sb200
