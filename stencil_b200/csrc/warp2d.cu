// warp2d.cu — instantiations of the 2-D warp kernels (warp2d.cuh, warp2dr.cuh) behind the
// non-template launchers of warp2d_api.h (own translation unit: compiles in parallel with capi.cu).
#include "warp2d_api.h"
#include "warp2dr.cuh"  // includes warp2d.cuh

namespace sb200 {

cudaError_t launch_warp2d_any(const void* in, void* out, int elem_size, const Grid& g, int shape, Domain dom, int K,
                              long long ob, long long oe, const Tuning& tn, cudaStream_t st, const SlabLinkDev* link) {
    if (elem_size == 8)
        return launch_warp2d<double>((const double*)in, (double*)out, g, shape, dom, K, ob, oe, tn, st, link);
    return launch_warp2d<float>((const float*)in, (float*)out, g, shape, dom, K, ob, oe, tn, st, link);
}

cudaError_t launch_warp2dr_any(const void* in, void* out, int elem_size, const Grid& g, int shape, Domain dom, int K,
                               long long ob, long long oe, const Tuning& tn, cudaStream_t st, const SlabLinkDev* link) {
    if (elem_size == 8)
        return launch_warp2dr<double>((const double*)in, (double*)out, g, shape, dom, K, ob, oe, tn, st, link);
    return launch_warp2dr<float>((const float*)in, (float*)out, g, shape, dom, K, ob, oe, tn, st, link);
}

}  // namespace sb200
