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

namespace {
template <class T, int SHAPE>
int resident_k(int K) {
    switch (K) {
        case 1: return w2d_resident_warps<T, SHAPE, 1>();
        case 2: return w2d_resident_warps<T, SHAPE, 2>();
        case 3: return w2d_resident_warps<T, SHAPE, 3>();
        case 4: return w2d_resident_warps<T, SHAPE, 4>();
        case 5: return w2d_resident_warps<T, SHAPE, 5>();
        case 6: return w2d_resident_warps<T, SHAPE, 6>();
    }
    return 0;
}
template <class T>
int resident_s(int shape, int K) {
    return shape == STAR ? resident_k<T, STAR>(K) : shape == BOX ? resident_k<T, BOX>(K) : shape == BOXSEP ? resident_k<T, BOXSEP>(K) : 0;
}
}  // namespace

int warp2d_resident_warps(int elem_size, int shape, int K) {
    return elem_size == 8 ? resident_s<double>(shape, K) : resident_s<float>(shape, K);
}

cudaError_t launch_warp2dr_any(const void* in, void* out, int elem_size, const Grid& g, int shape, Domain dom, int K,
                               long long ob, long long oe, const Tuning& tn, cudaStream_t st, const SlabLinkDev* link) {
    if (elem_size == 8)
        return launch_warp2dr<double>((const double*)in, (double*)out, g, shape, dom, K, ob, oe, tn, st, link);
    return launch_warp2dr<float>((const float*)in, (float*)out, g, shape, dom, K, ob, oe, tn, st, link);
}

}  // namespace sb200
