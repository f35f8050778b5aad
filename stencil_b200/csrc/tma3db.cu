// tma3db.cu — instantiations of the 3-D radius-2 box kernel (tma3db.cuh) behind the non-template
// launcher of tma3db_api.h (own translation unit: compiles in parallel with capi.cu).
#include "tma3db_api.h"
#include "tma3db.cuh"

namespace sb200 {
namespace {
constexpr int R = 2;
using TileF64 = T3B<double, R, 2, 4>;   // 64 x 64
using TileF32W = T3B<float, R, 4, 4>;   // 128 x 64
using TileF32N = T3B<float, R, 2, 8>;   // 64 x 128

// columns the tiles of a row compute (a lane computes whether or not its points are stored)
template <class C> long long computed_columns(long long nx) { return (nx + C::VX - 1) / C::VX * C::BX; }
bool narrow_f32(long long nx) { return computed_columns<TileF32N>(nx) < computed_columns<TileF32W>(nx); }

template <class T, class C, bool LINKED>
cudaError_t launch_one(const CUtensorMap& imap, const CUtensorMap& omap, T* out, const Grid& g, long long ob,
                       long long oe, const T3Plan& p, unsigned blocks, cudaStream_t st, const SlabLinkDev& link) {
    auto kernel = sweep_tma3db_kernel<T, R, C::PXV, C::PY, LINKED>;
    // function attributes are per device: set on every launch (cheap next to a 125-add-per-point step)
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
    if (e != cudaSuccess) return e;
    kernel<<<blocks, C::THREADS, C::SMEM, st>>>(imap, omap, out, g, ob, oe, p.ntx, p.nty, p.chunk_planes,
                                                 Avg<T>::value(BOX, 3, R), link);
    return cudaGetLastError();
}

template <class T, class C>
cudaError_t launch_t(const CUtensorMap& imap, const CUtensorMap& omap, T* out, const Grid& g, long long ob, long long oe,
                     const T3Plan& p, cudaStream_t st, const SlabLinkDev* link_in) {
    if (link_in && (link_in->link_lo || link_in->link_hi)) {
        SlabLinkDev link = *link_in;
        long long nc = p.nchunks;
        if (nc > (oe - ob) / link.face_planes) nc = (oe - ob) / link.face_planes;
        if (nc < 1) nc = 1;
        link.nchunks = (int)nc;
        return launch_one<T, C, true>(imap, omap, out, g, ob, oe, p, (unsigned)(nc * p.ntx * p.nty), st, link);
    }
    return launch_one<T, C, false>(imap, omap, out, g, ob, oe, p, (unsigned)p.blocks, st, SlabLinkDev{});
}

template <class C> void tile_of(int* bx, int* by, int* vx, int* vy) { *bx = C::BX, *by = C::BY, *vx = C::VX, *vy = C::VY; }
}  // namespace

void tma3db_tile(int elem_size, long long nx, int* bx, int* by, int* vx, int* vy) {
    if (elem_size == 8) tile_of<TileF64>(bx, by, vx, vy);
    else if (narrow_f32(nx)) tile_of<TileF32N>(bx, by, vx, vy);
    else tile_of<TileF32W>(bx, by, vx, vy);
}

T3Plan tma3db_plan(int elem_size, long long nx, long long ny, long long planes, int sm_count, int prefer_planes) {
    if (elem_size == 8) return t3b_plan<TileF64>(nx, ny, planes, sm_count, prefer_planes);
    return narrow_f32(nx) ? t3b_plan<TileF32N>(nx, ny, planes, sm_count, prefer_planes)
                          : t3b_plan<TileF32W>(nx, ny, planes, sm_count, prefer_planes);
}

cudaError_t launch_tma3db_any(const CUtensorMap& imap, const CUtensorMap& omap, void* out, int elem_size, const Grid& g,
                              long long ob, long long oe, const T3Plan& p, cudaStream_t st, const SlabLinkDev* link) {
    if (oe <= ob) return cudaSuccess;
    if (elem_size == 8) return launch_t<double, TileF64>(imap, omap, (double*)out, g, ob, oe, p, st, link);
    if (narrow_f32(g.n[0])) return launch_t<float, TileF32N>(imap, omap, (float*)out, g, ob, oe, p, st, link);
    return launch_t<float, TileF32W>(imap, omap, (float*)out, g, ob, oe, p, st, link);
}

}  // namespace sb200
