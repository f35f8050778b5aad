// warp2d_api.h — interface between capi.cu and warp2d.cu (internal): the 2-D warp kernels are
// instantiated in their own translation unit so that the library compiles in parallel.
#pragma once
#include <cuda_runtime.h>

#include "common.cuh"
#include "slablink.cuh"

namespace sb200 {
// Radius-1 star / box / separable box (shape = STAR, BOX, BOXSEP), K fused steps (1 .. W2D_MAXK);
// link == nullptr or without internal faces: plain launch, else the LINKED variant.
cudaError_t launch_warp2d_any(const void* in, void* out, int elem_size, const Grid& g, int shape, Domain dom, int K,
                              long long ob, long long oe, const Tuning& tn, cudaStream_t st, const SlabLinkDev* link);
// Warps of the radius-1 kernel <elem_size, shape, K> one SM holds at once (what the launcher hands to
// w2d_plan); 0 when no device is available or the combination does not exist.
int warp2d_resident_warps(int elem_size, int shape, int K);
// Star radius 2..4 and box radius 2..4 (warp2dr.cuh).
cudaError_t launch_warp2dr_any(const void* in, void* out, int elem_size, const Grid& g, int shape, Domain dom, int K,
                               long long ob, long long oe, const Tuning& tn, cudaStream_t st, const SlabLinkDev* link);
}  // namespace sb200
