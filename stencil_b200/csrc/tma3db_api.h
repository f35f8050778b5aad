// tma3db_api.h — interface between capi.cu and tma3db.cu (internal): the 3-D radius-2 box kernel
// (tma3db.cuh) is instantiated in its own translation unit.
#pragma once
#include "tma3d.cuh"

namespace sb200 {
// Tile geometry the kernel uses for this element size and row length nx (fp32 has a 128- and a
// 64-wide variant; the narrower one when it wastes fewer columns): TMA load box (bx, by), stored
// part (vx, vy).
void tma3db_tile(int elem_size, long long nx, int* bx, int* by, int* vx, int* vy);
T3Plan tma3db_plan(int elem_size, long long nx, long long ny, long long planes, int sm_count, int prefer_planes);
// One step of the 125-point box over planes [ob, oe).  imap / omap: tensor maps of `in` (box bx x by)
// and of `out`'s interior (box vx x vy); link == nullptr or without internal faces: plain launch.
cudaError_t launch_tma3db_any(const CUtensorMap& imap, const CUtensorMap& omap, void* out, int elem_size, const Grid& g,
                              long long ob, long long oe, const T3Plan& p, cudaStream_t st, const SlabLinkDev* link);
}  // namespace sb200
