// resident2d_api.h — interface between capi.cu and resident2d.cu (internal).
#pragma once
#include <cuda_runtime.h>

#include "common.cuh"

namespace sb200 {
// 1 if a (nx, ny) grid of this element size fits the register-resident cluster kernel.
int resident2d_fits(long long nx, long long ny, int elem_size);
// Layout of the launch: cx x wr warp tiles of 128 x R points on ncta CTAs (0s if it does not fit).
void resident2d_layout(long long nx, long long ny, int elem_size, int* ncta, int* cx, int* wr, int* R);
// `steps` time steps in -> out in ONE launch (2-D, radius 1); the result is in `out`.
cudaError_t launch_resident2d(const void* in, void* out, const Grid& g, int shape, int elem_size, int steps,
                              cudaStream_t st);
}  // namespace sb200
