// resident2d.cu — host launcher of the on-chip resident 2-D sweep (own translation unit so
// its template instantiations compile in parallel with capi.cu).
#include "resident2d.cuh"

#include "resident2d_api.h"

namespace sb200 {

namespace {

template <class T, int SHAPE, int R, bool CLUSTER>
cudaError_t launch_one(const T* in, T* out, const Grid& g, int steps, const RsPlan& p, cudaStream_t st) {
    auto kern = sweep_resident2d_kernel<T, SHAPE, R, CLUSTER>;
    const size_t smem = 2 * RS_WARPS * sizeof(RsInbox<T>);
    // function attributes belong to the current device's context, and this is one launch per
    // call: set them every time rather than caching per process (sb200_init_multi, re-init)
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (CLUSTER) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return e;
    }
    const T avg = Avg<T>::value(SHAPE, 2, 1);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)p.ncta);
    cfg.blockDim = dim3(RS_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)p.ncta;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = CLUSTER ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, in, out, g, steps, p.cx, p.wr, avg);
}

template <class T, int SHAPE, int R>
cudaError_t launch_r(const T* in, T* out, const Grid& g, int steps, const RsPlan& p, cudaStream_t st) {
    if constexpr (sizeof(T) == 8 && (R > 8)) {
        return cudaErrorInvalidValue;
    } else {
        return p.ncta > 1 ? launch_one<T, SHAPE, R, true>(in, out, g, steps, p, st)
                          : launch_one<T, SHAPE, R, false>(in, out, g, steps, p, st);
    }
}

template <class T, int SHAPE>
cudaError_t launch_s(const T* in, T* out, const Grid& g, int steps, const RsPlan& p, cudaStream_t st) {
    switch (p.R) {
        case 1: return launch_r<T, SHAPE, 1>(in, out, g, steps, p, st);
        case 2: return launch_r<T, SHAPE, 2>(in, out, g, steps, p, st);
        case 4: return launch_r<T, SHAPE, 4>(in, out, g, steps, p, st);
        case 8: return launch_r<T, SHAPE, 8>(in, out, g, steps, p, st);
        case 12: return launch_r<T, SHAPE, 12>(in, out, g, steps, p, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace

int resident2d_fits(long long nx, long long ny, int elem_size) {
    return rs_plan(nx, ny, elem_size).ok ? 1 : 0;
}

void resident2d_layout(long long nx, long long ny, int elem_size, int* ncta, int* cx, int* wr, int* R) {
    const RsPlan p = rs_plan(nx, ny, elem_size);
    *ncta = p.ok ? p.ncta : 0;
    *cx = p.ok ? p.cx : 0;
    *wr = p.ok ? p.wr : 0;
    *R = p.ok ? p.R : 0;
}

cudaError_t launch_resident2d(const void* in, void* out, const Grid& g, int shape, int elem_size, int steps,
                              cudaStream_t st) {
    const RsPlan p = rs_plan(g.n[0], g.n[1], elem_size);
    if (!p.ok || g.ndim != 2 || g.r != 1 || steps < 1) return cudaErrorInvalidValue;
    if (elem_size == 8)
        return shape == STAR ? launch_s<double, STAR>((const double*)in, (double*)out, g, steps, p, st)
                             : launch_s<double, BOX>((const double*)in, (double*)out, g, steps, p, st);
    return shape == STAR ? launch_s<float, STAR>((const float*)in, (float*)out, g, steps, p, st)
                         : launch_s<float, BOX>((const float*)in, (float*)out, g, steps, p, st);
}

}  // namespace sb200
