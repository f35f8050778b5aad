// capi.cu — the C ABI of include/stencil_b200.h over the sm_100a kernels.
//
// Replaces the reference's accelerator dispatch (src/stencil/stencil.cpp:33-54:
// athread_spawn + athread_join of one of four slave kernels).  No CPU fallback: every
// entry point fails with SB200_ERR_NO_DEVICE when no CUDA device is usable.
#include <cuda.h>
#include <cuda_runtime.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/stencil_b200.h"
#include "common.cuh"
#include "generic.cuh"
#include "resident2d_api.h"
#include "tma3d.cuh"
#include "tma3dr.cuh"
#include "tma3db_api.h"
#include "util_kernels.cuh"
#include "warp2d_api.h"
#include "warp2d.cuh"   // planner and limits only: the kernels are instantiated in warp2d.cu
#include "warp2dr.cuh"
#include "line1d.cuh"

using namespace sb200;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CU_TRY(expr)                                                                       \
    do {                                                                                   \
        cudaError_t e_ = (expr);                                                           \
        if (e_ != cudaSuccess)                                                             \
            return fail(SB200_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), \
                        __FILE__, __LINE__);                                               \
    } while (0)

// Owns the CUDA events of one call: destroyed on every return path, errors included.
struct Events {
    std::vector<cudaEvent_t> all;
    cudaError_t err = cudaSuccess;
    cudaEvent_t make(unsigned flags = cudaEventDisableTiming) {
        cudaEvent_t e = nullptr;
        const cudaError_t rc = cudaEventCreateWithFlags(&e, flags);
        if (rc != cudaSuccess) {
            if (err == cudaSuccess) err = rc;
            return nullptr;
        }
        all.push_back(e);
        return e;
    }
    ~Events() {
        for (cudaEvent_t e : all) cudaEventDestroy(e);
    }
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct MapKey {
    const void* ptr;
    int dtype;
    long long px, py, pl;
    int box_x, box_y;  // TMA box; box_y < 0 marks a store map (interior extents, origin at interior (0,0))
};
struct MapEntry {
    MapKey key;
    CUtensorMap map;
};

// Everything the library owns on ONE device.  A process binds one device (sb200_init) or a
// group (sb200_init_multi); every host thread works on the State its thread-local pointer
// names: the caller's thread on the first device of the group, the worker threads of a
// multi-GPU sb200_run_host each on their own.
struct State {
    bool ready = false;
    int device = -1;
    int sm_count = 0;
    std::vector<MapEntry> maps;
    std::vector<const void*> attr_done;  // kernels whose launch attributes are set on this device
    unsigned long long* noalloc_scratch = nullptr;  // device address of g_noalloc_scratch on this device
    // cached device buffers of sb200_run_host
    void* buf[2] = {nullptr, nullptr};
    size_t buf_bytes = 0;
    unsigned long long* d_acc = nullptr;
    // pipelined host path: dense staging image + copy-in / copy-out streams
    void* staging = nullptr;
    size_t staging_bytes = 0;
    cudaStream_t st_in = nullptr, st_out = nullptr, st_sweep[4] = {nullptr, nullptr, nullptr, nullptr};
};
std::vector<State*> g_group;          // devices bound by this process, in slab order
thread_local State* tls_state = nullptr;
State g_unbound;                      // placeholder before sb200_init (device -1, not ready)
EncodeTiledFn g_encode = nullptr;
std::mutex g_mu;

State& DS() {
    if (tls_state) return *tls_state;
    return g_group.empty() ? g_unbound : *g_group[0];
}

size_t esize(int dtype) { return dtype == SB200_F64 ? 8 : 4; }

int validate(const sb200_desc* d) {
    if (!d) return fail(SB200_ERR_INVALID, "null descriptor");
    if (d->ndim < 1 || d->ndim > 3) return fail(SB200_ERR_INVALID, "ndim must be 1, 2 or 3 (got %d)", d->ndim);
    if (d->shape != SB200_STAR && d->shape != SB200_BOX) return fail(SB200_ERR_INVALID, "bad shape %d", d->shape);
    if (d->dtype != SB200_F32 && d->dtype != SB200_F64) return fail(SB200_ERR_INVALID, "bad dtype %d", d->dtype);
    if (d->radius < 1 || d->radius > 16) return fail(SB200_ERR_INVALID, "radius must be in 1..16 (got %u)", d->radius);
    for (int a = 0; a < 3; ++a) {
        if (a < d->ndim && d->n[a] == 0) return fail(SB200_ERR_INVALID, "n[%d] is zero", a);
        if (a >= d->ndim && d->n[a] > 1) return fail(SB200_ERR_INVALID, "n[%d] must be 1 for ndim=%d", a, d->ndim);
        if (d->n[a] > (1ull << 40)) return fail(SB200_ERR_INVALID, "n[%d] too large", a);
    }
    if (d->kernel < SB200_KERNEL_AUTO || d->kernel > SB200_KERNEL_LINE1D)
        return fail(SB200_ERR_INVALID, "bad kernel selector %d", d->kernel);
    if (d->ngpus > SB200_MAX_GPUS) return fail(SB200_ERR_INVALID, "ngpus must be 0..%d (got %u)", SB200_MAX_GPUS, d->ngpus);
    if (d->flags & ~(uint32_t)(SB200_FLAG_REASSOCIATE | SB200_FLAG_BLOCK_TILE)) return fail(SB200_ERR_INVALID, "unknown flag bits 0x%x", d->flags);
    return SB200_OK;
}

// Device layout, see sb200_layout / common.cuh.
Grid make_grid(const sb200_desc* d, uint32_t ghost) {
    Grid g;
    const int r = (int)d->radius;
    const long long al = 128 / (long long)esize(d->dtype);  // elements per 128 B
    g.ndim = d->ndim;
    g.r = r;
    g.ghost = (int)ghost;
    for (int a = 0; a < 3; ++a) g.n[a] = a < d->ndim ? (long long)d->n[a] : 1;
    g.ox = (int)al;  // >= radius (<=16) and >= any temporal-halo depth
    g.pitch_x = (g.ox + g.n[0] + r + al - 1) / al * al;
    const int h1 = d->ndim >= 2 ? r + (d->ndim == 2 ? (int)ghost : 0) : 0;
    const int h2 = d->ndim >= 3 ? r + (int)ghost : 0;
    g.oy = h1;
    g.oz = h2;
    g.pitch_y = g.n[1] + 2 * h1;
    g.planes = g.n[2] + 2 * h2;
    return g;
}

// Bring one device up: context, SM count, the TMA descriptor encoder, the checksum
// accumulator and the streams of the pipelined host path (one-time costs belong here, not
// in the first timed call).  Leaves `dev` the calling thread's current device.
int bring_up(State& st, int dev) {
    CU_TRY(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, dev));
    if (prop.major != 10)
        return fail(SB200_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only",
                    dev, prop.major, prop.minor);
    st.device = dev;
    st.sm_count = prop.multiProcessorCount;
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        CU_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        if (!fn || qres != cudaDriverEntryPointSuccess)
            return fail(SB200_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
        g_encode = (EncodeTiledFn)fn;
    }
    CU_TRY(cudaMalloc(&st.d_acc, sizeof(unsigned long long)));
    CU_TRY(cudaStreamCreateWithFlags(&st.st_in, cudaStreamNonBlocking));
    CU_TRY(cudaStreamCreateWithFlags(&st.st_out, cudaStreamNonBlocking));
    for (auto& q : st.st_sweep) CU_TRY(cudaStreamCreateWithFlags(&q, cudaStreamNonBlocking));
    st.ready = true;
    return SB200_OK;
}

void tear_down(State& st) {
    if (!st.ready) return;
    cudaSetDevice(st.device);
    cudaDeviceSynchronize();
    for (int i = 0; i < 2; ++i) {
        if (st.buf[i]) cudaFree(st.buf[i]);
        st.buf[i] = nullptr;
    }
    st.buf_bytes = 0;
    if (st.d_acc) cudaFree(st.d_acc);
    st.d_acc = nullptr;
    if (st.staging) cudaFree(st.staging);
    st.staging = nullptr;
    st.staging_bytes = 0;
    if (st.st_in) cudaStreamDestroy(st.st_in);
    if (st.st_out) cudaStreamDestroy(st.st_out);
    for (auto& q : st.st_sweep) {
        if (q) cudaStreamDestroy(q);
        q = nullptr;
    }
    st.st_in = st.st_out = nullptr;
    st.maps.clear();
    st.attr_done.clear();  // function attributes are per device context
    st.ready = false;
}

int device_count_checked(int* n) {
    // load every kernel when the context is created (not inside the first timed launch);
    // has no effect if the process initialised CUDA before
    setenv("CUDA_MODULE_LOADING", "EAGER", 0);
    *n = 0;
    cudaError_t e = cudaGetDeviceCount(n);
    if (e != cudaSuccess || *n == 0) {
        cudaGetLastError();
        return fail(SB200_ERR_NO_DEVICE,
                    "no CUDA device available (%s); this backend has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    return SB200_OK;
}

// Bind `n` devices (devices == nullptr: 0..n-1).  Caller holds g_mu.
int bind_group(int n, const int* devices) {
    int have = 0;
    if (int rc = device_count_checked(&have)) return rc;
    if (n < 1 || n > SB200_MAX_GPUS) return fail(SB200_ERR_INVALID, "device group of %d (1..%d)", n, SB200_MAX_GPUS);
    for (int i = 0; i < n; ++i) {
        const int d = devices ? devices[i] : i;
        if (d < 0 || d >= have) return fail(SB200_ERR_INVALID, "device %d of %d visible", d, have);
    }
    std::vector<State*> group;
    int rc = SB200_OK;
    for (int i = 0; i < n && rc == SB200_OK; ++i) {
        group.push_back(new State);
        rc = bring_up(*group.back(), devices ? devices[i] : i);
    }
    if (rc != SB200_OK) {
        for (State* st : group) {
            tear_down(*st);
            delete st;
        }
        return rc;
    }
    g_group.swap(group);
    tls_state = nullptr;
    CU_TRY(cudaSetDevice(g_group[0]->device));
    return SB200_OK;
}

int ensure_ready() {
    if (DS().ready) return SB200_OK;
    return bind_group(1, nullptr);  // a caller that never called sb200_init gets device 0
}

// Launch attributes are per (function, device context): remember them per State.
template <class F>
int set_smem_attr_once(F* kernel, size_t bytes) {
    State& st = DS();
    const void* key = reinterpret_cast<const void*>(kernel);
    for (const void* k : st.attr_done)
        if (k == key) return SB200_OK;
    CU_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    st.attr_done.push_back(key);
    return SB200_OK;
}

// Tensor maps of one device grid.  Load maps (store == false) span the whole allocation, halo
// included, with box_x x box_y boxes (out-of-range parts are zero-filled).  Store maps span the
// INTERIOR in x and y — origin at interior (0, 0), every plane of the allocation in z — so that
// the TMA unit clips what a ragged tile holds beyond the grid and the Dirichlet halo is never
// written.
int get_tensor_map(const void* ptr, int dtype, const Grid& g, int box_x, int box_y, bool store,
                   const CUtensorMap** out) {
    State& D = DS();
    const int key_by = store ? -box_y : box_y;
    for (auto& m : D.maps)
        if (m.key.ptr == ptr && m.key.dtype == dtype && m.key.px == g.pitch_x && m.key.py == g.pitch_y &&
            m.key.pl == g.planes && m.key.box_x == box_x && m.key.box_y == key_by) {
            *out = &m.map;
            return SB200_OK;
        }
    MapEntry me;
    me.key = {ptr, dtype, g.pitch_x, g.pitch_y, g.planes, box_x, key_by};
    const size_t es = esize(dtype);
    // store maps end at the last full 16 bytes of an interior row (the unit the TMA clips in);
    // the kernel stores the ragged end itself
    const long long per16 = 16 / (long long)esize(dtype);
    cuuint64_t dims[3] = {(cuuint64_t)(store ? g.n[0] / per16 * per16 : g.pitch_x),
                          (cuuint64_t)(store ? g.n[1] : g.pitch_y), (cuuint64_t)g.planes};
    if (store && dims[0] == 0) dims[0] = (cuuint64_t)per16;  // a valid descriptor; nothing of it is ever in range
    cuuint64_t strides[2] = {(cuuint64_t)g.pitch_x * es, (cuuint64_t)g.pitch_x * g.pitch_y * es};
    cuuint32_t box[3] = {(cuuint32_t)box_x, (cuuint32_t)box_y, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    char* base = (char*)const_cast<void*>(ptr) + (store ? ((size_t)g.oy * g.pitch_x + g.ox) * es : 0);
    CUresult r = g_encode(&me.map, dtype == SB200_F64 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                          3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(SB200_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    if (D.maps.size() >= 64) D.maps.erase(D.maps.begin());
    D.maps.push_back(me);
    *out = &D.maps.back().map;
    return SB200_OK;
}

template <class T, int K, int SHAPE>
int launch_tma3d_k(const T* in, T* out, int dtype, const Grid& g, Domain dom, long long ob,
                   long long oe, cudaStream_t st, const SlabLinkDev& link_in, int prefer_chunk) {
    using C = T3<T, K>;
    if (oe <= ob) return SB200_OK;
    const bool linked = link_in.link_lo || link_in.link_hi;
    if (int rc = linked ? set_smem_attr_once(sweep_tma3d_kernel<T, K, SHAPE, true>, C::SMEM)
                        : set_smem_attr_once(sweep_tma3d_kernel<T, K, SHAPE, false>, C::SMEM))
        return rc;
    // copies: a later insertion may evict / move cache entries
    const CUtensorMap* mp = nullptr;
    if (int rc = get_tensor_map(in, dtype, g, C::BX, C::BY, false, &mp)) return rc;
    const CUtensorMap imap = *mp;
    if (int rc = get_tensor_map(out, dtype, g, C::VX, C::VY, true, &mp)) return rc;
    const CUtensorMap omap = *mp;
    const T3Plan p = t3_plan<T, K>(g.n[0], g.n[1], oe - ob, DS().sm_count, prefer_chunk);
    const T avg = Avg<T>::value(SHAPE, 3, 1);
    if (linked) {
        SlabLinkDev link = link_in;
        long long nc = p.nchunks;
        if (nc > (oe - ob) / link.face_planes) nc = (oe - ob) / link.face_planes;
        if (nc < 1) nc = 1;
        link.nchunks = (int)nc;
        sweep_tma3d_kernel<T, K, SHAPE, true><<<(unsigned)(nc * p.ntx * p.nty), C::THREADS, C::SMEM, st>>>(
            imap, omap, in, out, g, dom, ob, oe, p.ntx, p.nty, p.chunk_planes, avg, link);
    } else {
        sweep_tma3d_kernel<T, K, SHAPE, false><<<(unsigned)p.blocks, C::THREADS, C::SMEM, st>>>(
            imap, omap, in, out, g, dom, ob, oe, p.ntx, p.nty, p.chunk_planes, avg, link_in);
    }
    CU_TRY(cudaGetLastError());
    return SB200_OK;
}

const SlabLinkDev kNoLink = {};

template <class T>
int launch_tma3d(const T* in, T* out, int dtype, const Grid& g, int shape, Domain dom, int K,
                 long long ob, long long oe, cudaStream_t st, const SlabLinkDev& link = kNoLink,
                 int prefer_chunk = 0) {
    if (shape == BOXSEP) {  // SB200_FLAG_REASSOCIATE
        switch (K) {
            case 1: return launch_tma3d_k<T, 1, BOXSEP>(in, out, dtype, g, dom, ob, oe, st, link, prefer_chunk);
            case 2: return launch_tma3d_k<T, 2, BOXSEP>(in, out, dtype, g, dom, ob, oe, st, link, prefer_chunk);
            default: return fail(SB200_ERR_INVALID, "TMA3D box time block %d out of range", K);
        }
    }
    if (shape == SB200_BOX) {
        switch (K) {
            case 1: return launch_tma3d_k<T, 1, BOX>(in, out, dtype, g, dom, ob, oe, st, link, prefer_chunk);
            case 2: return launch_tma3d_k<T, 2, BOX>(in, out, dtype, g, dom, ob, oe, st, link, prefer_chunk);
            default: return fail(SB200_ERR_INVALID, "TMA3D box time block %d out of range", K);
        }
    }
    switch (K) {
        case 1: return launch_tma3d_k<T, 1, STAR>(in, out, dtype, g, dom, ob, oe, st, link, prefer_chunk);
        case 2: return launch_tma3d_k<T, 2, STAR>(in, out, dtype, g, dom, ob, oe, st, link, prefer_chunk);
        default: return fail(SB200_ERR_INVALID, "TMA3D time block %d out of range", K);
    }
}

// 3-D star of radius 2..4 (13-, 19-, 25-point), one step per launch (tma3dr.cuh).
template <class T, int R>
int launch_tma3dr_r(const T* in, T* out, int dtype, const Grid& g, long long ob, long long oe, cudaStream_t st,
                    int prefer_chunk, const SlabLinkDev* link_in) {
    using C = T3R<T, R>;
    if (oe <= ob) return SB200_OK;
    const bool linked = link_in && (link_in->link_lo || link_in->link_hi);
    if (int rc = linked ? set_smem_attr_once(sweep_tma3dr_kernel<T, R, true>, C::SMEM)
                        : set_smem_attr_once(sweep_tma3dr_kernel<T, R, false>, C::SMEM))
        return rc;
    const CUtensorMap* mp = nullptr;
    if (int rc = get_tensor_map(in, dtype, g, C::BX, C::BY, false, &mp)) return rc;
    const CUtensorMap imap = *mp;
    if (int rc = get_tensor_map(out, dtype, g, C::VX, C::VY, true, &mp)) return rc;
    const CUtensorMap omap = *mp;
    const T3Plan p = t3r_plan<T, R>(g.n[0], g.n[1], oe - ob, DS().sm_count, prefer_chunk);
    const T avg = Avg<T>::value(STAR, 3, R);
    if (linked) {
        SlabLinkDev link = *link_in;
        long long nc = p.nchunks;
        if (nc > (oe - ob) / link.face_planes) nc = (oe - ob) / link.face_planes;
        if (nc < 1) nc = 1;
        link.nchunks = (int)nc;
        sweep_tma3dr_kernel<T, R, true><<<(unsigned)(nc * p.ntx * p.nty), C::THREADS, C::SMEM, st>>>(
            imap, omap, out, g, ob, oe, p.ntx, p.nty, p.chunk_planes, avg, link);
    } else {
        sweep_tma3dr_kernel<T, R, false><<<(unsigned)p.blocks, C::THREADS, C::SMEM, st>>>(
            imap, omap, out, g, ob, oe, p.ntx, p.nty, p.chunk_planes, avg, SlabLinkDev{});
    }
    CU_TRY(cudaGetLastError());
    return SB200_OK;
}

template <class T>
int launch_tma3dr(const T* in, T* out, int dtype, const Grid& g, long long ob, long long oe, cudaStream_t st,
                  int prefer_chunk, const SlabLinkDev* link = nullptr) {
    switch (g.r) {
        case 2: return launch_tma3dr_r<T, 2>(in, out, dtype, g, ob, oe, st, prefer_chunk, link);
        case 3: return launch_tma3dr_r<T, 3>(in, out, dtype, g, ob, oe, st, prefer_chunk, link);
        case 4: return launch_tma3dr_r<T, 4>(in, out, dtype, g, ob, oe, st, prefer_chunk, link);
    }
    return fail(SB200_ERR_INVALID, "TMA3D star kernel: radius %d out of range", g.r);
}

// 3-D box of radius 2 (125-point), one step per launch (tma3db.cuh, instantiated in tma3db.cu).
int launch_tma3db(const void* in, void* out, int dtype, const Grid& g, long long ob, long long oe, cudaStream_t st,
                  int prefer_chunk, const SlabLinkDev* link = nullptr) {
    if (oe <= ob) return SB200_OK;
    if (g.r != 2) return fail(SB200_ERR_INVALID, "TMA3D box kernel: radius %d out of range", g.r);
    const int es = (int)esize(dtype);
    int bx, by, vx, vy;
    tma3db_tile(es, g.n[0], &bx, &by, &vx, &vy);
    const CUtensorMap* mp = nullptr;
    if (int rc = get_tensor_map(in, dtype, g, bx, by, false, &mp)) return rc;
    const CUtensorMap imap = *mp;
    if (int rc = get_tensor_map(out, dtype, g, vx, vy, true, &mp)) return rc;
    const CUtensorMap omap = *mp;
    const T3Plan p = tma3db_plan(es, g.n[0], g.n[1], oe - ob, DS().sm_count, prefer_chunk);
    CU_TRY(launch_tma3db_any(imap, omap, out, es, g, ob, oe, p, st, link));
    return SB200_OK;
}

// tiles / chunks of the radius-R star / radius-2 box kernels (host logic)
T3Plan t3rb_plan(const sb200_desc* d, const Grid& g, long long planes, int sm_count, int prefer);

template <class T>
T3Plan t3r_plan_any(int r, long long nx, long long ny, long long planes, int sm_count, int prefer) {
    return r == 2 ? t3r_plan<T, 2>(nx, ny, planes, sm_count, prefer)
           : r == 3 ? t3r_plan<T, 3>(nx, ny, planes, sm_count, prefer)
                    : t3r_plan<T, 4>(nx, ny, planes, sm_count, prefer);
}

// Kernel family + fused depth for a descriptor.
struct Choice {
    int kernel;
    int max_k;
    int k;
};

// WARP2D family: radius 1 star/box (warp2d.cuh), radius 2..4 star and box (warp2dr.cuh).
bool warp2d_supports(const sb200_desc* d) { return d->ndim == 2 && d->radius >= 1 && d->radius <= 4; }
// What AUTO takes from it: the 81-point box in fp32 loses to the generic kernel (198 vs 221 GStencil/s,
// profiles/r02u_box2d_probe.log); every other case wins.
bool warp2d_preferred(const sb200_desc* d) {
    return warp2d_supports(d) && !(d->shape == SB200_BOX && d->radius == 4 && d->dtype == SB200_F32);
}

// TMA3D family: radius 1 star / box (tma3d.cuh), radius 2..4 star (tma3dr.cuh), radius 2 box (tma3db.cuh).
bool tma3d_supports(const sb200_desc* d) {
    return d->ndim == 3 && (d->radius == 1 || (d->radius >= 2 && d->radius <= 4 && d->shape == SB200_STAR) ||
                            (d->radius == 2 && d->shape == SB200_BOX));
}
T3Plan t3rb_plan(const sb200_desc* d, const Grid& g, long long planes, int sm_count, int prefer) {
    if (d->shape == SB200_BOX) return tma3db_plan((int)esize(d->dtype), g.n[0], g.n[1], planes, sm_count, prefer);
    return d->dtype == SB200_F64 ? t3r_plan_any<double>(g.r, g.n[0], g.n[1], planes, sm_count, prefer)
                                 : t3r_plan_any<float>(g.r, g.n[0], g.n[1], planes, sm_count, prefer);
}
// What AUTO takes from it: at radius 3 and 4 the two-rows-per-thread tiles pay off in fp64 only
// (1.37-1.40x the generic kernel; fp32 1.03x / 0.96x: profiles/r02o_star3d_probe.log).
bool tma3d_preferred(const sb200_desc* d) {
    return tma3d_supports(d) && (d->radius <= 2 || d->dtype == SB200_F64);
}

// Kernel shape of the tma3d family and of the radius-1 warp2d kernel: the separable box under
// SB200_FLAG_REASSOCIATE.
int t3_shape(const sb200_desc* d) {
    return d->shape == SB200_BOX && (d->flags & SB200_FLAG_REASSOCIATE) ? (int)BOXSEP : d->shape;
}

Choice choose(const sb200_desc* d) {
    Choice c;
    const bool w2d_ok = warp2d_preferred(d);
    const bool t3_ok = tma3d_preferred(d);
    const bool l1_ok = d->ndim == 1 && d->radius <= 4;
    int kern = d->kernel;
    if (kern == SB200_KERNEL_AUTO)
        kern = w2d_ok ? SB200_KERNEL_WARP2D : t3_ok ? SB200_KERNEL_TMA3D : l1_ok ? SB200_KERNEL_LINE1D : SB200_KERNEL_GENERIC;
    c.kernel = kern;
    c.max_k = kern == SB200_KERNEL_WARP2D ? (d->radius > 1 ? (d->shape == SB200_BOX ? w2r_box_max_k((int)d->radius, (int)esize(d->dtype))
                                                                                     : w2r_max_k((int)d->radius, (int)esize(d->dtype)))
                                                           : W2D_MAXK)
              : kern == SB200_KERNEL_TMA3D ? (d->radius > 1 ? 1 : d->shape == SB200_BOX ? T3_MAXK_BOX : T3_MAXK)
              : kern == SB200_KERNEL_RESIDENT2D ? (1 << 30)
              : kern == SB200_KERNEL_LINE1D ? L1D_MAXK : 1;
    int k = (int)d->time_block;
    if (kern == SB200_KERNEL_RESIDENT2D) k = c.max_k;  // the whole call is one launch
    // radius-1 warp kernel: 4 fused steps (the separable box of SB200_FLAG_REASSOCIATE is cheap enough for 6)
    if (k == 0) k = kern == SB200_KERNEL_WARP2D ? (d->radius > 1 ? c.max_k : t3_shape(d) == BOXSEP ? 6 : 4) : kern == SB200_KERNEL_LINE1D ? l1d_default_k((int)d->radius) : kern == SB200_KERNEL_TMA3D ? 2 : 1;
    if (k > c.max_k) k = c.max_k;
    c.k = k;
    return c;
}

int check_choice(const sb200_desc* d, const Choice& c) {
    if (c.kernel == SB200_KERNEL_WARP2D && !warp2d_supports(d))
        return fail(SB200_ERR_UNSUPPORTED, "WARP2D kernel needs ndim=2 and radius=1 (star or box), radius 2..4 (star) or radius 2 (box)");
    if (c.kernel == SB200_KERNEL_LINE1D && !(d->ndim == 1 && d->radius <= 4))
        return fail(SB200_ERR_UNSUPPORTED, "LINE1D kernel needs ndim=1, radius 1..4");
    if (c.kernel == SB200_KERNEL_TMA3D && !tma3d_supports(d))
        return fail(SB200_ERR_UNSUPPORTED, "TMA3D kernel needs ndim=3 and radius=1 (star or box) or radius 2..4 (star)");
    if (c.kernel == SB200_KERNEL_RESIDENT2D &&
        !(d->ndim == 2 && d->radius == 1 && resident2d_fits((long long)d->n[0], (long long)d->n[1], (int)esize(d->dtype))))
        return fail(SB200_ERR_UNSUPPORTED, "RESIDENT2D kernel needs ndim=2, radius=1 and a grid that fits one "
                                           "thread-block cluster's registers (about 672^2 fp32 / 512^2 fp64)");
    return SB200_OK;
}

// -b/--block-size as a tile preference (opt-in): rows / planes one CTA or warp streams.
int tile_pref(const sb200_desc* d) {
    return (d->flags & SB200_FLAG_BLOCK_TILE) && d->block_hint > 0 ? (int)(d->block_hint > (1u << 20) ? (1u << 20) : d->block_hint) : 0;
}

// One fused pass of k steps: a -> b over outermost range [ob, oe).
template <class T>
int pass_t(const sb200_desc* d, const Choice& c, const Grid& g, Domain dom, const T* a, T* b, int k,
           long long ob, long long oe, cudaStream_t st) {
    switch (c.kernel) {
        case SB200_KERNEL_GENERIC:
            if (k != 1) return fail(SB200_ERR_INVALID, "generic kernel fuses 1 step");
            CU_TRY(launch_generic<T>(a, b, g, d->shape, ob, oe, st));
            return SB200_OK;
        case SB200_KERNEL_WARP2D:
            if (g.r > 1)
                CU_TRY(launch_warp2dr_any(a, b, (int)sizeof(T), g, d->shape, dom, k, ob, oe, Tuning{DS().sm_count, tile_pref(d)}, st, nullptr));
            else
                CU_TRY(launch_warp2d_any(a, b, (int)sizeof(T), g, t3_shape(d), dom, k, ob, oe, Tuning{DS().sm_count, tile_pref(d)}, st, nullptr));
            return SB200_OK;
        case SB200_KERNEL_TMA3D:
            if (g.r > 1) {
                if (k != 1) return fail(SB200_ERR_INVALID, "the radius-2..4 TMA3D kernels fuse 1 step");
                if (d->shape == SB200_BOX) return launch_tma3db(a, b, d->dtype, g, ob, oe, st, tile_pref(d));
                return launch_tma3dr<T>(a, b, d->dtype, g, ob, oe, st, tile_pref(d));
            }
            return launch_tma3d<T>(a, b, d->dtype, g, t3_shape(d),
                                   dom, k, ob, oe, st, kNoLink, tile_pref(d));
        case SB200_KERNEL_LINE1D:
            CU_TRY(launch_line1d<T>(a, b, g, d->shape, k, st));
            return SB200_OK;
    }
    return fail(SB200_ERR_INVALID, "kernel family %d cannot run a partial pass", c.kernel);
}

int pass(const sb200_desc* d, const Choice& c, const Grid& g, Domain dom, const void* a, void* b,
         int k, long long ob, long long oe, cudaStream_t st) {
    if (d->dtype == SB200_F64)
        return pass_t<double>(d, c, g, dom, (const double*)a, (double*)b, k, ob, oe, st);
    return pass_t<float>(d, c, g, dom, (const float*)a, (float*)b, k, ob, oe, st);
}

long long outer_n(const Grid& g) { return g.ndim == 3 ? g.n[2] : g.ndim == 2 ? g.n[1] : 1; }

Domain make_domain(const Grid& g, int lo_b, int hi_b) {
    Domain dom;
    const long long n = outer_n(g);
    dom.lo = lo_b ? 0 : -(long long)g.ghost;
    dom.hi = hi_b ? n : n + g.ghost;
    return dom;
}

int sweep_impl(const sb200_desc* d, uint32_t ghost, int lo_b, int hi_b, void* a, void* b,
               uint32_t steps, cudaStream_t st, int* result_in_b, uint32_t* launches,
               uint32_t* used_k, int* used_kernel) {
    Choice c = choose(d);
    if (int rc = check_choice(d, c)) return rc;
    const Grid g = make_grid(d, ghost);
    if (d->ndim == 1 && ghost) return fail(SB200_ERR_INVALID, "1-D grids cannot carry ghost cells");
    const int r = (int)d->radius;
    // Small 2-D grids, single GPU, many steps: keep the grid on chip for the whole call.
    const bool whole = ghost == 0 && lo_b && hi_b;
    bool resident = c.kernel == SB200_KERNEL_RESIDENT2D;
    if (resident && !whole) return fail(SB200_ERR_UNSUPPORTED, "RESIDENT2D kernel cannot sweep a slab with ghost planes");
    if (d->kernel == SB200_KERNEL_AUTO && d->time_block == 0 && c.kernel == SB200_KERNEL_WARP2D && whole &&
        d->radius == 1 && steps >= 16 && resident2d_fits(g.n[0], g.n[1], (int)esize(d->dtype)))
        resident = true;
    if (resident) {
        if (steps > 0) CU_TRY(launch_resident2d(a, b, g, d->shape, (int)esize(d->dtype), (int)steps, st));
        if (result_in_b) *result_in_b = steps > 0 ? 1 : 0;
        if (launches) *launches = steps > 0 ? 1 : 0;
        if (used_k) *used_k = steps;
        if (used_kernel) *used_kernel = SB200_KERNEL_RESIDENT2D;
        return SB200_OK;
    }
    if ((!lo_b || !hi_b) && (long long)steps * r > (long long)ghost)
        return fail(SB200_ERR_INVALID, "%u steps of radius %d need %lld ghost planes, have %u", steps, r,
                    (long long)steps * r, ghost);
    const Domain dom = make_domain(g, lo_b, hi_b);
    const long long n = outer_n(g);
    void* src = a;
    void* dst = b;
    uint32_t done = 0, nl = 0;
    while (done < steps) {
        const int k = (int)((steps - done) < (uint32_t)c.k ? (steps - done) : (uint32_t)c.k);
        done += k;
        // after `done` steps the still-valid region has shrunk by done*r on internal faces
        const long long ob = lo_b ? 0 : -((long long)ghost - (long long)done * r);
        const long long oe = hi_b ? n : n + ((long long)ghost - (long long)done * r);
        if (int rc = pass(d, c, g, dom, src, dst, k, ob, oe, st)) return rc;
        ++nl;
        void* t = src;
        src = dst;
        dst = t;
    }
    if (result_in_b) *result_in_b = (src == b) ? 1 : 0;
    if (launches) *launches = nl;
    if (used_k) *used_k = (uint32_t)c.k;
    if (used_kernel) *used_kernel = c.kernel;
    return SB200_OK;
}

}  // namespace

// ---- no-allocation checksum (sb200_checksum_device_noalloc) ---------------------------------
// Scratch of the variant below: a file-scope array in device memory, so the entry point never
// allocates.  Every GPU that loads the library holds these 65 792 bytes until the process
// exits.  Layout in 64-bit words: [0] ticket of the last-block election, [1] the result,
// [32 ..) one partial sum per block.  In HOST code this name is only a stand-in for the symbol,
// not a device address: the launcher resolves it with cudaGetSymbolAddress once per device.
constexpr int NOALLOC_HEADER_WORDS = 32;  // keeps the partials 256-byte aligned
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NOALLOC_HEADER_WORDS + SB200_NOALLOC_MAX_BLOCKS];

// Same digest as checksum_kernel (util_kernels.cuh), computed without atomics on the data:
// block b sums the points [b * span, (b+1) * span), stores ONE partial, and the block that
// draws the last ticket adds the partials up in index order (deterministic), publishes the
// result and puts the ticket back to zero for the next call.
template <class T>
__global__ void __launch_bounds__(256)
checksum_noalloc_kernel(const T* __restrict__ grid, Grid g, long long gn0, long long gn1, long long off,
                        long long span, unsigned long long* __restrict__ scratch) {
    __shared__ unsigned long long warp_sum[8];
    __shared__ bool is_last;
    const long long total = g.n[0] * g.n[1] * g.n[2];
    const long long lo = (long long)blockIdx.x * span;
    const long long hi = lo + span < total ? lo + span : total;
    uint64_t local = 0;
    for (long long i = lo + threadIdx.x; i < hi; i += 256) {
        const long long x = i % g.n[0];
        const long long y = (i / g.n[0]) % g.n[1];
        const long long z = i / (g.n[0] * g.n[1]);
        long long gy = y, gz = z;
        if (g.ndim == 3) gz += off;
        if (g.ndim == 2) gy += off;
        const uint64_t lin = ((uint64_t)gz * (uint64_t)gn1 + (uint64_t)gy) * (uint64_t)gn0 + (uint64_t)x;
        local += mix64(Bits<T>::of(grid[g.idx(x, y, z)]) ^ mix64(lin));
    }
    auto block_sum = [&](uint64_t v) -> uint64_t {  // valid in thread 0
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        __syncthreads();  // warp_sum may still be read from the previous use
        if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = v;
        __syncthreads();
        uint64_t t = 0;
        if (threadIdx.x == 0)
            for (int w = 0; w < 8; ++w) t += warp_sum[w];
        return t;
    };
    const uint64_t mine = block_sum(local);
    if (threadIdx.x == 0) {
        scratch[NOALLOC_HEADER_WORDS + blockIdx.x] = mine;
        __threadfence();
        is_last = atomicAdd(&scratch[0], 1ull) == (unsigned long long)gridDim.x - 1;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    uint64_t acc = 0;
    for (unsigned j = threadIdx.x; j < gridDim.x; j += 256) acc += __ldcg(&scratch[NOALLOC_HEADER_WORDS + j]);
    const uint64_t all = block_sum(acc);
    if (threadIdx.x == 0) {
        scratch[1] = all;
        scratch[0] = 0;
    }
}

// ---------------------------------------------------------------------------------
extern "C" {

int sb200_abi_version(void) { return SB200_ABI_VERSION; }
const char* sb200_last_error(void) { return g_err; }

int sb200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int sb200_init(int device) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_group.size() == 1 && g_group[0]->device == device) return SB200_OK;
    if (!g_group.empty())
        return fail(SB200_ERR_INVALID, "already bound to %zu device(s) starting at %d; sb200_finalize first",
                    g_group.size(), g_group[0]->device);
    return bind_group(1, &device);
}

int sb200_init_multi(int n, const int* devices) {
    std::lock_guard<std::mutex> lk(g_mu);
    if ((int)g_group.size() == n) {
        bool same = true;
        for (int i = 0; i < n; ++i) same = same && g_group[(size_t)i]->device == (devices ? devices[i] : i);
        if (same) return SB200_OK;
    }
    if (!g_group.empty())
        return fail(SB200_ERR_INVALID, "already bound to %zu device(s); sb200_finalize first", g_group.size());
    return bind_group(n, devices);
}

int sb200_group_size(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    return (int)g_group.size();
}

void sb200_finalize(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (State* st : g_group) {
        tear_down(*st);
        delete st;
    }
    g_group.clear();
    tls_state = nullptr;
}

int sb200_layout_for(const sb200_desc* desc, uint32_t ghost, sb200_layout* layout) {
    if (int rc = validate(desc)) return rc;
    if (!layout) return fail(SB200_ERR_INVALID, "null layout");
    if (desc->ndim == 1 && ghost) return fail(SB200_ERR_INVALID, "1-D grids cannot carry ghost cells");
    const Grid g = make_grid(desc, ghost);
    layout->pitch_x = (uint64_t)g.pitch_x;
    layout->pitch_y = (uint64_t)g.pitch_y;
    layout->planes = (uint64_t)g.planes;
    layout->ox = (uint64_t)g.ox;
    layout->oy = (uint64_t)g.oy;
    layout->oz = (uint64_t)g.oz;
    layout->bytes = (uint64_t)g.elems() * esize(desc->dtype);
    return SB200_OK;
}

static int pipeline_chunks(const sb200_desc* d, const Choice& c, const Grid& g, size_t dense_bytes);

int sb200_plan(const sb200_desc* desc, uint32_t steps, int sm_count, sb200_plan_info* info) {
    if (int rc = validate(desc)) return rc;
    if (!info || sm_count < 1) return fail(SB200_ERR_INVALID, "bad argument");
    memset(info, 0, sizeof *info);
    Choice c = choose(desc);
    if (int rc = check_choice(desc, c)) return rc;
    const Grid g = make_grid(desc, 0);
    const int es = (int)esize(desc->dtype);
    bool resident = c.kernel == SB200_KERNEL_RESIDENT2D;
    if (desc->kernel == SB200_KERNEL_AUTO && desc->time_block == 0 && c.kernel == SB200_KERNEL_WARP2D &&
        desc->radius == 1 && steps >= 16 && resident2d_fits(g.n[0], g.n[1], es))
        resident = true;
    info->pipeline_chunks = 1;
    if (resident) {
        int ncta, cx, wr, R;
        resident2d_layout(g.n[0], g.n[1], es, &ncta, &cx, &wr, &R);
        info->kernel = SB200_KERNEL_RESIDENT2D;
        info->time_block = steps;
        info->launches = steps > 0 ? 1 : 0;
        info->tiles_x = (uint32_t)cx;
        info->tiles_y = (uint32_t)wr;
        info->chunks = (uint32_t)ncta;
        info->chunk_extent = (uint32_t)R;
        return SB200_OK;
    }
    info->kernel = c.kernel;
    info->time_block = (uint32_t)c.k;
    info->launches = (steps + c.k - 1) / c.k;
    const int k = steps && steps < (uint32_t)c.k ? (int)steps : c.k;
    if (c.kernel == SB200_KERNEL_WARP2D) {
        // the launcher's own number when a device is bound (it lengthens chunks to fit one round of
        // resident warps); without one the report is the device-independent base geometry
        // (SB200_PLAN_RESIDENT_WARPS stands in for the device's answer when planning without one: CPU tests)
        const char* rw = getenv("SB200_PLAN_RESIDENT_WARPS");
        const int resident = g.r != 1 ? 0 : !g_group.empty() ? warp2d_resident_warps(es, t3_shape(desc), k) : rw ? atoi(rw) : 0;
        const W2dPlan p = w2d_plan(g.n[0], g.n[1], k, sm_count, g.r, tile_pref(desc), resident);
        info->tiles_x = (uint32_t)p.nstrips;
        info->tiles_y = 1;
        info->chunks = (uint32_t)p.nchunks;
        info->chunk_extent = (uint32_t)p.chunk_rows;
    } else if (c.kernel == SB200_KERNEL_TMA3D && desc->radius > 1) {
        const T3Plan p = t3rb_plan(desc, g, g.n[2], sm_count, tile_pref(desc));
        info->tiles_x = (uint32_t)p.ntx;
        info->tiles_y = (uint32_t)p.nty;
        info->chunks = (uint32_t)p.nchunks;
        info->chunk_extent = (uint32_t)p.chunk_planes;
    } else if (c.kernel == SB200_KERNEL_TMA3D) {
        T3Plan p;
        if (desc->dtype == SB200_F64)
            p = k == 1 ? t3_plan<double, 1>(g.n[0], g.n[1], g.n[2], sm_count, tile_pref(desc))
                : k == 2 ? t3_plan<double, 2>(g.n[0], g.n[1], g.n[2], sm_count, tile_pref(desc))
                         : t3_plan<double, 2>(g.n[0], g.n[1], g.n[2], sm_count, tile_pref(desc));
        else
            p = k == 1 ? t3_plan<float, 1>(g.n[0], g.n[1], g.n[2], sm_count, tile_pref(desc))
                : k == 2 ? t3_plan<float, 2>(g.n[0], g.n[1], g.n[2], sm_count, tile_pref(desc))
                         : t3_plan<float, 2>(g.n[0], g.n[1], g.n[2], sm_count, tile_pref(desc));
        info->tiles_x = (uint32_t)p.ntx;
        info->tiles_y = (uint32_t)p.nty;
        info->chunks = (uint32_t)p.nchunks;
        info->chunk_extent = (uint32_t)p.chunk_planes;
    } else if (c.kernel == SB200_KERNEL_LINE1D) {
        const L1dPlan p = l1d_plan(g.n[0], k, g.r);
        info->tiles_x = (uint32_t)p.nseg;
        info->tiles_y = 1;
        info->chunks = 1;
        info->chunk_extent = (uint32_t)p.valid;
    }
    sb200_desc d2 = *desc;
    d2.iterations = steps;
    const size_t dense = (size_t)(g.n[0] + 2 * g.r) * g.pitch_y * g.planes * (size_t)es;
    info->pipeline_chunks = (uint32_t)pipeline_chunks(&d2, c, g, dense);
    return SB200_OK;
}

int sb200_max_time_block(const sb200_desc* desc) {
    if (int rc = validate(desc)) return rc;
    const Choice c = choose(desc);
    if (int rc = check_choice(desc, c)) return rc;
    return c.max_k;
}

int sb200_sweep_device(const sb200_desc* desc, uint32_t ghost, int lo_is_boundary, int hi_is_boundary,
                       void* a, void* b, uint32_t steps, void* stream, int* result_in_b,
                       uint32_t* launches) {
    if (int rc = validate(desc)) return rc;
    if (!a || !b) return fail(SB200_ERR_INVALID, "null device buffer");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    return sweep_impl(desc, ghost, lo_is_boundary, hi_is_boundary, a, b, steps, (cudaStream_t)stream,
                      result_in_b, launches, nullptr, nullptr);
}

int sb200_sweep_device_range(const sb200_desc* desc, uint32_t ghost, int lo_is_boundary,
                             int hi_is_boundary, const void* a, void* b, uint32_t steps,
                             uint64_t z_begin, uint64_t z_end, void* stream, uint32_t* launches) {
    if (int rc = validate(desc)) return rc;
    if (!a || !b) return fail(SB200_ERR_INVALID, "null device buffer");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    const Choice c = choose(desc);
    if (int rc = check_choice(desc, c)) return rc;
    if (c.kernel == SB200_KERNEL_RESIDENT2D)
        return fail(SB200_ERR_UNSUPPORTED, "RESIDENT2D kernel sweeps whole grids only");
    if (steps < 1 || (int)steps > c.max_k)
        return fail(SB200_ERR_INVALID, "range sweep fuses 1..%d steps (got %u)", c.max_k, steps);
    if (desc->ndim == 1) return fail(SB200_ERR_INVALID, "range sweep needs ndim >= 2");
    const Grid g = make_grid(desc, ghost);
    const long long n = outer_n(g);
    if (z_end > (uint64_t)n || z_begin > z_end) return fail(SB200_ERR_INVALID, "bad plane range");
    if ((!lo_is_boundary || !hi_is_boundary) && (long long)steps * desc->radius > (long long)ghost)
        return fail(SB200_ERR_INVALID, "%u steps need %u ghost planes, have %u", steps, steps * desc->radius, ghost);
    const Domain dom = make_domain(g, lo_is_boundary, hi_is_boundary);
    if (launches) *launches = z_end > z_begin ? 1 : 0;
    return pass(desc, c, g, dom, a, b, (int)steps, (long long)z_begin, (long long)z_end, (cudaStream_t)stream);
}

int sb200_device_alloc(void** ptr, size_t bytes) {
    if (!ptr || bytes == 0) return fail(SB200_ERR_INVALID, "bad allocation request");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    CU_TRY(cudaMalloc(ptr, bytes));
    CU_TRY(cudaMemset(*ptr, 0, bytes));
    CU_TRY(cudaDeviceSynchronize());
    return SB200_OK;
}

int sb200_device_free(void* ptr) {
    if (!ptr) return SB200_OK;
    std::lock_guard<std::mutex> lk(g_mu);
    for (size_t i = 0; i < DS().maps.size();)  // tensor maps of a freed buffer must not be reused
        if (DS().maps[i].key.ptr == ptr) DS().maps.erase(DS().maps.begin() + i); else ++i;
    CU_TRY(cudaFree(ptr));
    return SB200_OK;
}

int sb200_ipc_export(void* dev_ptr, unsigned char handle[SB200_IPC_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == SB200_IPC_HANDLE_BYTES, "IPC handle size");
    if (!dev_ptr || !handle) return fail(SB200_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    cudaIpcMemHandle_t h;
    CU_TRY(cudaIpcGetMemHandle(&h, dev_ptr));
    memcpy(handle, &h, sizeof h);
    return SB200_OK;
}

int sb200_ipc_open(const unsigned char handle[SB200_IPC_HANDLE_BYTES], void** dev_ptr) {
    if (!dev_ptr || !handle) return fail(SB200_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    CU_TRY(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return SB200_OK;
}

int sb200_ipc_close(void* dev_ptr) {
    if (!dev_ptr) return SB200_OK;
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    CU_TRY(cudaIpcCloseMemHandle(dev_ptr));
    return SB200_OK;
}

int sb200_link_status(const uint32_t* flags_local, uint32_t words[SB200_LINK_FLAG_WORDS]) {
    if (!flags_local || !words) return fail(SB200_ERR_INVALID, "null argument");
    CU_TRY(cudaDeviceSynchronize());
    CU_TRY(cudaMemcpy(words, flags_local, sizeof(uint32_t) * SB200_LINK_FLAG_WORDS, cudaMemcpyDeviceToHost));
    return SB200_OK;
}

int sb200_link_reset(uint32_t* flags_local) {
    if (!flags_local) return fail(SB200_ERR_INVALID, "null argument");
    CU_TRY(cudaDeviceSynchronize());
    CU_TRY(cudaMemset(flags_local, 0, sizeof(uint32_t) * SB200_LINK_FLAG_WORDS));
    CU_TRY(cudaDeviceSynchronize());
    return SB200_OK;
}

static int slab_linked_impl(const sb200_desc* desc, uint32_t ghost, int lo_is_boundary, int hi_is_boundary,
                            const void* a, void* b, uint32_t steps, const sb200_slab_link* link, void* stream,
                            uint32_t* arrivals);

int sb200_sweep_slab_linked(const sb200_desc* desc, uint32_t ghost, int lo_is_boundary,
                            int hi_is_boundary, const void* a, void* b, uint32_t steps,
                            const sb200_slab_link* link, void* stream, uint32_t* arrivals) {
    if (int rc = validate(desc)) return rc;
    if (!a || !b || !link) return fail(SB200_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    return slab_linked_impl(desc, ghost, lo_is_boundary, hi_is_boundary, a, b, steps, link, stream, arrivals);
}

// On the calling thread's current State / device; the caller holds g_mu.
static int slab_linked_impl(const sb200_desc* desc, uint32_t ghost, int lo_is_boundary, int hi_is_boundary,
                            const void* a, void* b, uint32_t steps, const sb200_slab_link* link, void* stream,
                            uint32_t* arrivals) {
    const Choice c = choose(desc);
    if (int rc = check_choice(desc, c)) return rc;
    const bool w2d = c.kernel == SB200_KERNEL_WARP2D;
    if (!w2d && c.kernel != SB200_KERNEL_TMA3D)
        return fail(SB200_ERR_UNSUPPORTED, "linked slab sweeps need the WARP2D or the TMA3D family");
    if (steps < 1 || (int)steps > c.max_k)
        return fail(SB200_ERR_INVALID, "linked sweep fuses 1..%d steps (got %u)", c.max_k, steps);
    const Grid g = make_grid(desc, ghost);
    const long long n = outer_n(g);
    if (ghost < steps * desc->radius)
        return fail(SB200_ERR_INVALID, "%u steps need %u ghost planes, have %u", steps, steps * desc->radius, ghost);
    if (n < 2LL * ghost) return fail(SB200_ERR_INVALID, "slab of %lld planes is thinner than two ghost zones", n);
    if (!link->flags_local) return fail(SB200_ERR_INVALID, "null flag block");
    if ((!lo_is_boundary && (!link->peer_dst_lo || !link->flags_lo)) ||
        (!hi_is_boundary && (!link->peer_dst_hi || !link->flags_hi)))
        return fail(SB200_ERR_INVALID, "missing neighbour mapping for an internal face");
    SlabLinkDev L = {};
    L.link_lo = lo_is_boundary ? 0 : 1;
    L.link_hi = hi_is_boundary ? 0 : 1;
    L.face_planes = (int)ghost;
    const long long plane = w2d ? g.pitch_x : g.pitch_x * g.pitch_y;  // one index of the outermost axis
    L.peer_out_lo = link->peer_dst_lo;
    L.shift_lo = n * plane;  // my plane z is the lower neighbour's plane z + n (its upper ghosts)
    L.peer_out_hi = link->peer_dst_hi;
    L.shift_hi = -n * plane;
    L.wait_lo = link->flags_local + 0;
    L.wait_hi = link->flags_local + 1;
    L.signal_lo = link->flags_lo ? link->flags_lo + 1 : nullptr;  // I am its upper neighbour
    L.signal_hi = link->flags_hi ? link->flags_hi + 0 : nullptr;
    L.wait_value = link->wait_arrivals;
    L.err_flag = link->flags_local + 2;
    const Domain dom = make_domain(g, lo_is_boundary, hi_is_boundary);
    const long long ob = 0, oe = n;
    cudaStream_t st = (cudaStream_t)stream;
    if (w2d) {
        // face owners are warps: one per x strip and face
        if (arrivals) *arrivals = (uint32_t)w2d_plan(g.n[0], n, (int)steps, DS().sm_count, g.r, 0).nstrips;
        const Tuning tn{DS().sm_count, tile_pref(desc)};
        if (g.r > 1)
            CU_TRY(launch_warp2dr_any(a, b, (int)esize(desc->dtype), g, desc->shape, dom, (int)steps, ob, oe, tn, st, &L));
        else
            CU_TRY(launch_warp2d_any(a, b, (int)esize(desc->dtype), g, t3_shape(desc), dom, (int)steps, ob, oe, tn, st, &L));
        return SB200_OK;
    }
    if (g.r > 1) {  // radius 2..4 star (tma3dr.cuh), radius 2 box (tma3db.cuh)
        const T3Plan p = t3rb_plan(desc, g, 1, DS().sm_count, 0);
        if (arrivals) *arrivals = (uint32_t)(p.ntx * p.nty);
        if (desc->shape == SB200_BOX) return launch_tma3db(a, b, desc->dtype, g, ob, oe, st, tile_pref(desc), &L);
        if (desc->dtype == SB200_F64)
            return launch_tma3dr<double>((const double*)a, (double*)b, desc->dtype, g, ob, oe, st, tile_pref(desc), &L);
        return launch_tma3dr<float>((const float*)a, (float*)b, desc->dtype, g, ob, oe, st, tile_pref(desc), &L);
    }
    int ntx, nty;
    if (desc->dtype == SB200_F64) {
        switch (steps) {
            case 1: { auto p = t3_plan<double, 1>(g.n[0], g.n[1], 1, DS().sm_count); ntx = p.ntx; nty = p.nty; } break;
            case 2: { auto p = t3_plan<double, 2>(g.n[0], g.n[1], 1, DS().sm_count); ntx = p.ntx; nty = p.nty; } break;
            default: { auto p = t3_plan<double, 2>(g.n[0], g.n[1], 1, DS().sm_count); ntx = p.ntx; nty = p.nty; } break;
        }
    } else {
        switch (steps) {
            case 1: { auto p = t3_plan<float, 1>(g.n[0], g.n[1], 1, DS().sm_count); ntx = p.ntx; nty = p.nty; } break;
            case 2: { auto p = t3_plan<float, 2>(g.n[0], g.n[1], 1, DS().sm_count); ntx = p.ntx; nty = p.nty; } break;
            default: { auto p = t3_plan<float, 2>(g.n[0], g.n[1], 1, DS().sm_count); ntx = p.ntx; nty = p.nty; } break;
        }
    }
    if (arrivals) *arrivals = (uint32_t)(ntx * nty);
    if (desc->dtype == SB200_F64)
        return launch_tma3d<double>((const double*)a, (double*)b, desc->dtype, g, t3_shape(desc), dom, (int)steps, ob, oe, st, L);
    return launch_tma3d<float>((const float*)a, (float*)b, desc->dtype, g, t3_shape(desc), dom, (int)steps, ob, oe, st, L);
}

static int fill_impl(const sb200_desc* desc, uint32_t ghost, void* grid, int init, uint64_t seed, uint64_t z_offset,
                     uint64_t z_total, cudaStream_t st);

int sb200_fill_device(const sb200_desc* desc, uint32_t ghost, void* grid, int init, uint64_t seed,
                      uint64_t z_offset, uint64_t z_total, void* stream) {
    if (int rc = validate(desc)) return rc;
    if (!grid) return fail(SB200_ERR_INVALID, "null device buffer");
    if (init != SB200_INIT_REF && init != SB200_INIT_RANDOM) return fail(SB200_ERR_INVALID, "bad init %d", init);
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    return fill_impl(desc, ghost, grid, init, seed, z_offset, z_total, (cudaStream_t)stream);
}

static int fill_impl(const sb200_desc* desc, uint32_t ghost, void* grid, int init, uint64_t seed, uint64_t z_offset,
                     uint64_t z_total, cudaStream_t st) {
    const Grid g = make_grid(desc, ghost);
    long long gn[3] = {g.n[0], g.n[1], g.n[2]};
    if (z_total) gn[desc->ndim - 1] = (long long)z_total;
    const long long total = g.elems();
    long long blocks = (total + 255) / 256;
    if (blocks > DS().sm_count * 32) blocks = DS().sm_count * 32;
    if (desc->dtype == SB200_F64)
        fill_kernel<double><<<(unsigned)blocks, 256, 0, st>>>((double*)grid, g, init == SB200_INIT_RANDOM, seed, gn[0], gn[1], gn[2], (long long)z_offset);
    else
        fill_kernel<float><<<(unsigned)blocks, 256, 0, st>>>((float*)grid, g, init == SB200_INIT_RANDOM, seed, gn[0], gn[1], gn[2], (long long)z_offset);
    CU_TRY(cudaGetLastError());
    return SB200_OK;
}

// dense host padded array <-> device layout; `interior_only` copies n[] without halo.
static int copy_grid(const sb200_desc* desc, uint32_t ghost, const void* src, void* dst, bool to_device,
                     bool interior_only, cudaStream_t st) {
    const Grid g = make_grid(desc, ghost);
    const size_t es = esize(desc->dtype);
    const long long r = g.r;
    // host padded extents (ghost planes are part of the host slab image too)
    const long long hx = g.n[0] + 2 * r, hy = g.pitch_y, hz = g.planes;
    cudaMemcpy3DParms p;
    memset(&p, 0, sizeof p);
    void* host = to_device ? const_cast<void*>(src) : dst;
    void* dev = to_device ? dst : const_cast<void*>(src);
    cudaPitchedPtr hp = make_cudaPitchedPtr(host, (size_t)hx * es, (size_t)hx, (size_t)hy);
    cudaPitchedPtr dp = make_cudaPitchedPtr(dev, (size_t)g.pitch_x * es, (size_t)g.pitch_x, (size_t)g.pitch_y);
    cudaPos hpos, dpos;
    if (interior_only) {
        hpos = make_cudaPos((size_t)r * es, (size_t)g.oy, (size_t)g.oz);
        dpos = make_cudaPos((size_t)g.ox * es, (size_t)g.oy, (size_t)g.oz);
        p.extent = make_cudaExtent((size_t)g.n[0] * es, (size_t)g.n[1], (size_t)g.n[2]);
    } else {
        hpos = make_cudaPos(0, 0, 0);
        dpos = make_cudaPos((size_t)(g.ox - r) * es, 0, 0);
        p.extent = make_cudaExtent((size_t)hx * es, (size_t)hy, (size_t)hz);
    }
    p.srcPtr = to_device ? hp : dp;
    p.dstPtr = to_device ? dp : hp;
    p.srcPos = to_device ? hpos : dpos;
    p.dstPos = to_device ? dpos : hpos;
    p.kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
    CU_TRY(cudaMemcpy3DAsync(&p, st));
    return SB200_OK;
}

// Device interior -> host interior for the outermost-axis range [ob, oe) (interior coords).
static int download_interior_range(const sb200_desc* desc, const void* dev, void* host, long long ob,
                                   long long oe, cudaStream_t st) {
    if (oe <= ob) return SB200_OK;
    const Grid g = make_grid(desc, 0);
    const size_t es = esize(desc->dtype);
    const long long r = g.r;
    const long long hx = g.n[0] + 2 * r, hy = g.pitch_y;
    cudaMemcpy3DParms p;
    memset(&p, 0, sizeof p);
    p.srcPtr = make_cudaPitchedPtr(const_cast<void*>(dev), (size_t)g.pitch_x * es, (size_t)g.pitch_x, (size_t)g.pitch_y);
    p.dstPtr = make_cudaPitchedPtr(host, (size_t)hx * es, (size_t)hx, (size_t)hy);
    if (g.ndim == 3) {
        p.srcPos = make_cudaPos((size_t)g.ox * es, (size_t)g.oy, (size_t)(g.oz + ob));
        p.dstPos = make_cudaPos((size_t)r * es, (size_t)g.oy, (size_t)(g.oz + ob));
        p.extent = make_cudaExtent((size_t)g.n[0] * es, (size_t)g.n[1], (size_t)(oe - ob));
    } else {
        p.srcPos = make_cudaPos((size_t)g.ox * es, (size_t)(g.oy + ob), 0);
        p.dstPos = make_cudaPos((size_t)r * es, (size_t)(g.oy + ob), 0);
        p.extent = make_cudaExtent((size_t)g.n[0] * es, (size_t)(oe - ob), 1);
    }
    p.kind = cudaMemcpyDeviceToHost;
    CU_TRY(cudaMemcpy3DAsync(&p, st));
    return SB200_OK;
}

// Number of z (y in 2-D) parallelograms of the pipelined host path; 1 = not pipelined.
static int pipeline_chunks(const sb200_desc* d, const Choice& c, const Grid& g, size_t dense_bytes) {
    if (d->ndim < 2 || c.kernel == SB200_KERNEL_RESIDENT2D || d->iterations == 0) return 1;
    const long long n = outer_n(g);
    const long long reach = (long long)d->iterations * d->radius;  // planes a parallelogram drifts
    const long long wmin = 4LL * c.k * d->radius;                  // narrowest final region
    // starting edges are n/(C-1) apart and must be wider than their halos; nothing can be
    // downloaded early unless the grid is longer than the drift
    const long long last = reach > wmin ? reach : wmin;
    long long cmax = n > last ? n / wmin + 1 : 1;
    long long want = 11;
    if (const char* e = getenv("SB200_PIPELINE_CHUNKS")) want = atoll(e);
    else if (dense_bytes < (64u << 20)) want = 1;  // small grids: the copies are too short to hide
    if (want > cmax) want = cmax;
    if (want > 32) want = 32;
    return want < 2 ? 1 : (int)want;
}

// The drop-in call for grids that are worth pipelining: the outermost axis is cut into C
// parallelograms in (z, t) that lean towards low z by `radius` planes per time step.
// Parallelogram i only needs the planes below its right edge, so it starts as soon as upload
// chunk i has landed, and its final region [F_i, F_i+1) is complete — and starts its way
// back to the host — while the parallelograms to its right are still being swept.  H2D, the
// sweep and D2H run on separate streams; PCIe is used in both directions at once.  The slanted
// right edge of parallelogram i at time level t stays in the ping-pong buffer of that parity
// (later blocks of i write strictly below it, because block sizes never decrease), which is
// exactly the halo parallelogram i+1 needs: no extra storage, same kernels, same bits.
//
// The schedule is built first, as plain data (sb200_pipe_op: what, on which stream, over which
// planes, after which other ops), and executed afterwards: the same list is handed out by
// sb200_pipeline_schedule so that its geometry and its happens-before edges can be checked on a
// CPU (tests/test_pipeline_schedule.py proves every pair of unordered ops conflict-free).
static int pipeline_streams() {
    int NS = 2;
    if (const char* e = getenv("SB200_PIPELINE_STREAMS")) NS = atoi(e) < 1 ? 1 : (atoi(e) > 4 ? 4 : atoi(e));
    return NS;
}

// One slab of a multi-GPU call, in outermost interior indices LOCAL to its sub-grid: the slab
// downloads [own_lo, own_hi); planes outside are deep halo (see sb200_slab_plan).  The whole
// grid on one GPU is the slab {0, n}.
struct SlabRange {
    long long own_lo, own_hi;
};

static void build_pipeline(const sb200_desc* desc, const Choice& c, const Grid& g, int C, int NS,
                           std::vector<sb200_pipe_op>& ops, SlabRange own = {0, -1},
                           uint32_t* zone_uploads = nullptr) {
    const long long n = outer_n(g), r = g.r, T = desc->iterations;
    const long long n_outer_padded = g.ndim == 3 ? g.planes : g.pitch_y;
    if (own.own_hi < 0) own.own_hi = n;
    const bool low_zone = own.own_lo > 0, high_zone = own.own_hi < n;
    // Starting edges Z_i: C-1 equal upload chunks, then a last parallelogram that starts with
    // zero width at the high end (a triangle growing to T*r planes).  Final regions
    // F_i = Z_i - T*r clipped at 0: parallelograms that drift out of the grid before step T
    // (the first ~T*r planes) download nothing, so the first upload chunk does not have to
    // cover the whole drift before sweeping starts.
    std::vector<long long> F(C + 1), Z(C + 1);
    for (int i = 0; i < C; ++i) {
        Z[i] = n * i / (C - 1);
        F[i] = Z[i] - T * r > 0 ? Z[i] - T * r : 0;
    }
    Z[C] = F[C] = n;
    // block sizes: the short remainder block FIRST, so block sizes never decrease
    std::vector<int> blocks;
    if (T % c.k) blocks.push_back((int)(T % c.k));
    for (long long b = 0; b < T / c.k; ++b) blocks.push_back(c.k);
    const size_t nb = blocks.size();
    auto add = [&](int kind, int stream, int par, int block, uint32_t steps, int src, long long lo, long long hi,
                   int w0, int w1 = -1) {
        sb200_pipe_op o;
        o.kind = kind; o.stream = stream; o.parallelogram = par; o.block = block; o.steps = steps; o.src = src;
        o.lo = lo; o.hi = hi; o.wait[0] = w0; o.wait[1] = w1;
        ops.push_back(o);
        return (int)ops.size() - 1;
    };
    ops.clear();
    // upload chunk i = padded outer indices [U_i, U_i+1), all on the copy-in stream.  A slab
    // uploads its high deep-halo zone FIRST (and its low one is the start of chunk 0 anyway):
    // those are the host planes a neighbouring slab downloads into when T is even, and the
    // neighbours' downloads are held back until the zones have been read (zone_uploads).
    const long long hz = high_zone ? own.own_hi + r : n_outer_padded;  // first padded index of the high zone
    const int upH = high_zone ? add(SB200_PIPE_UPLOAD, 0, C - 1, -1, 0, -1, hz, n_outer_padded, -1) : -1;
    std::vector<int> up(C, -1);
    std::vector<char> needs_hz(C, 0);
    uint32_t zones = high_zone ? 1 : 0;
    bool low_covered = !low_zone;
    for (int i = 0; i < C; ++i) {
        const long long ub = i == 0 ? 0 : Z[i] + r, ue = i == C - 1 ? n_outer_padded : Z[i + 1] + r;
        const long long uc = ue < hz ? ue : hz;
        if (uc > ub) {
            up[i] = add(SB200_PIPE_UPLOAD, 0, i, -1, 0, -1, ub, uc, -1);
            if (!low_covered) {
                ++zones;
                low_covered = uc >= own.own_lo + r;
            }
        }
        needs_hz[i] = ue > hz;
    }
    if (zone_uploads) *zone_uploads = zones;
    // Consecutive parallelograms alternate between the sweep streams and run as a wavefront:
    // block b of parallelogram i+1 only waits for block b-1 of parallelogram i (its slanted
    // left halo; the same edge orders the write-after-read on the ping-pong buffer), so the
    // thin range launches of neighbouring parallelograms fill each other's partial waves.
    // last_before[b] = the newest op of the previous parallelogram issued up to its block b-1
    // (an empty block launches nothing; the op before it on that stream stands in for it).
    std::vector<int> prev_upto(nb + 1, -1), cur_upto(nb + 1, -1);
    for (int i = 0; i < C; ++i) {
        const int ss = 2 + i % NS;
        const long long ub = i == 0 ? 0 : Z[i] + r, ue = i == C - 1 ? n_outer_padded : Z[i + 1] + r;
        int last = add(SB200_PIPE_UNPACK, ss, i, -1, 0, -1, ub, ue, up[i] >= 0 ? up[i] : upH,
                       up[i] >= 0 && needs_hz[i] ? upH : -1);
        cur_upto[0] = last;  // what "block -1" of this parallelogram means to its right neighbour
        long long done_steps = 0;
        for (size_t b = 0; b < nb; ++b) {
            done_steps += blocks[b];
            long long L = i == 0 ? 0 : Z[i] - done_steps * r;
            long long R = i == C - 1 ? n : Z[i + 1] - done_steps * r;
            if (L < 0) L = 0;
            if (R > n) R = n;
            // a slab only needs the dependency cone of the planes it owns: after `done` steps
            // that is [own_lo - (T-done)*r, own_hi + (T-done)*r)
            const long long need_lo = own.own_lo - (T - done_steps) * r, need_hi = own.own_hi + (T - done_steps) * r;
            if (L < need_lo) L = need_lo;
            if (R > need_hi) R = need_hi;
            if (R > L) last = add(SB200_PIPE_SWEEP, ss, i, (int)b, (uint32_t)blocks[b], (int)(b & 1), L, R,
                                  i > 0 ? prev_upto[b] : -1);
            cur_upto[b + 1] = last;
        }
        const long long dl = F[i] > own.own_lo ? F[i] : own.own_lo, dh = F[i + 1] < own.own_hi ? F[i + 1] : own.own_hi;
        if (dh > dl) add(SB200_PIPE_DOWNLOAD, 1, i, -1, 0, (int)(nb & 1), dl, dh, last);
        prev_upto.swap(cur_upto);
    }
}

// Cross-device ordering of one slab's host reads against its neighbours' host writes.
struct SlabSync {
    cudaEvent_t zones_read = nullptr;           // recorded by this slab after its zone uploads
    cudaEvent_t wait[2] = {nullptr, nullptr};   // the neighbours' zones_read
    void (*host_barrier)(void*) = nullptr;      // returns once EVERY slab has recorded zones_read
    void* barrier_arg = nullptr;
    uint32_t zone_uploads = 0;
};

static int run_host_pipelined(const sb200_desc* desc, const Choice& c, void* in, void* out,
                              sb200_timing* timing, int C, SlabRange own = {0, -1},
                              const SlabSync* sync = nullptr) {
    const auto w0 = std::chrono::steady_clock::now();
    State& D = DS();
    const Grid g = make_grid(desc, 0);
    const size_t es = esize(desc->dtype);
    const long long r = g.r;
    const long long hx = g.n[0] + 2 * r;
    const long long rows_per_outer = g.ndim == 3 ? g.pitch_y : 1;
    const long long n_outer_padded = g.ndim == 3 ? g.planes : g.pitch_y;
    const size_t outer_bytes = (size_t)hx * rows_per_outer * es;  // dense bytes per outer index
    const size_t dense_bytes = outer_bytes * n_outer_padded;
    if (D.staging_bytes < dense_bytes) {
        if (D.staging) cudaFree(D.staging);
        D.staging = nullptr;
        D.staging_bytes = 0;
        CU_TRY(cudaMalloc(&D.staging, dense_bytes));
        D.staging_bytes = dense_bytes;
    }
    const int NS = pipeline_streams();
    std::vector<sb200_pipe_op> ops;
    build_pipeline(desc, c, g, C, NS, ops, own);
    cudaStream_t st = 0;  // the caller's stream: start / end of the whole pipeline
    auto stream_of = [&](const sb200_pipe_op& o) { return o.stream == 0 ? D.st_in : o.stream == 1 ? D.st_out : D.st_sweep[o.stream - 2]; };
    // one event per op that another op waits for
    Events events;
    cudaEvent_t ev[4];
    for (auto& e : ev) e = events.make(cudaEventDefault);
    std::vector<cudaEvent_t> op_done(ops.size(), nullptr);
    for (const auto& o : ops)
        for (int w : o.wait)
            if (w >= 0 && !op_done[(size_t)w]) op_done[(size_t)w] = events.make();
    // SB200_PIPELINE_TRACE=<file>: a start and an end event around every op, written out as CSV
    // after the call (diagnostics of the schedule; the extra events cost a few microseconds)
    const char* trace_path = getenv("SB200_PIPELINE_TRACE");
    std::vector<cudaEvent_t> t_begin, t_end;
    if (trace_path) {
        t_begin.resize(ops.size());
        t_end.resize(ops.size());
        for (size_t j = 0; j < ops.size(); ++j) {
            t_begin[j] = events.make(cudaEventDefault);
            t_end[j] = events.make(cudaEventDefault);
        }
    }
    CU_TRY(events.err);
    CU_TRY(cudaEventRecord(ev[0], st));
    CU_TRY(cudaStreamWaitEvent(D.st_in, ev[0], 0));
    CU_TRY(cudaStreamWaitEvent(D.st_out, ev[0], 0));
    for (int q = 0; q < NS; ++q) CU_TRY(cudaStreamWaitEvent(D.st_sweep[q], ev[0], 0));
    const Domain dom = make_domain(g, 1, 1);
    void* buf[2] = {D.buf[0], D.buf[1]};
    void* host_dst = (desc->iterations & 1u) ? out : in;
    uint32_t nl = 0;
    int uploads_left = 0;
    for (const auto& o : ops) uploads_left += o.kind == SB200_PIPE_UPLOAD;
    bool first_download = false;
    // Uploads have no predecessors, so they are all issued first, in list order (they share the
    // copy-in stream).  A slab then tells its neighbours when its deep-halo zones have been
    // read, and holds its own downloads back until theirs have (T even: slabs download into
    // the host array the others still upload from).
    uint32_t uploads_issued = 0;
    for (size_t j = 0; j < ops.size(); ++j) {
        const sb200_pipe_op& o = ops[j];
        if (o.kind != SB200_PIPE_UPLOAD) continue;
        if (trace_path) CU_TRY(cudaEventRecord(t_begin[j], D.st_in));
        CU_TRY(cudaMemcpyAsync((char*)D.staging + o.lo * outer_bytes, (const char*)in + o.lo * outer_bytes,
                               (size_t)(o.hi - o.lo) * outer_bytes, cudaMemcpyHostToDevice, D.st_in));
        if (trace_path) CU_TRY(cudaEventRecord(t_end[j], D.st_in));
        if (op_done[j]) CU_TRY(cudaEventRecord(op_done[j], D.st_in));
        if (--uploads_left == 0) CU_TRY(cudaEventRecord(ev[1], D.st_in));
        if (sync && ++uploads_issued == sync->zone_uploads) CU_TRY(cudaEventRecord(sync->zones_read, D.st_in));
    }
    if (sync) {
        if (sync->zone_uploads == 0) CU_TRY(cudaEventRecord(sync->zones_read, D.st_in));
        sync->host_barrier(sync->barrier_arg);  // every slab's zones_read is recorded from here on
        for (cudaEvent_t e : sync->wait)
            if (e) CU_TRY(cudaStreamWaitEvent(D.st_out, e, 0));
    }
    for (size_t j = 0; j < ops.size(); ++j) {
        const sb200_pipe_op& o = ops[j];
        if (o.kind == SB200_PIPE_UPLOAD) continue;
        cudaStream_t ss = stream_of(o);
        for (int w : o.wait)
            if (w >= 0) CU_TRY(cudaStreamWaitEvent(ss, op_done[(size_t)w], 0));
        if (trace_path) CU_TRY(cudaEventRecord(t_begin[j], ss));
        switch (o.kind) {
            case SB200_PIPE_UNPACK: {
                const long long rb = o.lo * rows_per_outer, re = o.hi * rows_per_outer;
                dim3 grid((unsigned)((g.pitch_x + 255) / 256), (unsigned)((re - rb) < 65535 ? (re - rb) : 65535));
                if (desc->dtype == SB200_F64)
                    unpack_dense2_kernel<double><<<grid, 256, 0, ss>>>((const double*)D.staging, (double*)buf[0], (double*)buf[1], g, hx, rb, re);
                else
                    unpack_dense2_kernel<float><<<grid, 256, 0, ss>>>((const float*)D.staging, (float*)buf[0], (float*)buf[1], g, hx, rb, re);
                CU_TRY(cudaGetLastError());
                break;
            }
            case SB200_PIPE_SWEEP:
                if (int rc = pass(desc, c, g, dom, buf[o.src], buf[o.src ^ 1], (int)o.steps, o.lo, o.hi, ss)) return rc;
                ++nl;
                break;
            case SB200_PIPE_DOWNLOAD:
                if (!first_download) {
                    CU_TRY(cudaEventRecord(ev[2], ss));
                    first_download = true;
                }
                if (int rc = download_interior_range(desc, buf[o.src], host_dst, o.lo, o.hi, ss)) return rc;
                break;
        }
        if (trace_path) CU_TRY(cudaEventRecord(t_end[j], ss));
        if (op_done[j]) CU_TRY(cudaEventRecord(op_done[j], ss));
    }
    if (!first_download) CU_TRY(cudaEventRecord(ev[2], D.st_out));
    CU_TRY(cudaEventRecord(ev[3], D.st_out));
    CU_TRY(cudaStreamWaitEvent(st, ev[3], 0));
    CU_TRY(cudaStreamSynchronize(D.st_out));
    for (int q = 0; q < NS; ++q) CU_TRY(cudaStreamSynchronize(D.st_sweep[q]));
    CU_TRY(cudaStreamSynchronize(st));
    CU_TRY(cudaStreamSynchronize(D.st_in));
    const auto w1 = std::chrono::steady_clock::now();
    if (trace_path) {
        if (FILE* f = fopen(trace_path, "w")) {
            fprintf(f, "op,kind,stream,parallelogram,block,steps,lo,hi,begin_ms,end_ms,device\n");
            for (size_t j = 0; j < ops.size(); ++j) {
                float b = 0, e = 0;
                cudaEventElapsedTime(&b, ev[0], t_begin[j]);
                cudaEventElapsedTime(&e, ev[0], t_end[j]);
                fprintf(f, "%zu,%d,%d,%d,%d,%u,%lld,%lld,%.4f,%.4f,%d\n", j, ops[j].kind, ops[j].stream, ops[j].parallelogram,
                        ops[j].block, ops[j].steps, (long long)ops[j].lo, (long long)ops[j].hi, b, e, D.device);
            }
            fclose(f);
        }
    }
    if (timing) {
        float ms = 0;
        cudaEventElapsedTime(&ms, ev[0], ev[1]); timing->h2d_ms = ms;    // all upload chunks
        cudaEventElapsedTime(&ms, ev[0], ev[3]); timing->sweep_ms = ms;  // whole pipeline on the device
        cudaEventElapsedTime(&ms, ev[2], ev[3]); timing->d2h_ms = ms;    // first to last download chunk
        timing->wall_ms = std::chrono::duration<double, std::milli>(w1 - w0).count();
        timing->launches = nl;
        timing->time_block = (uint32_t)c.k;
        timing->kernel = c.kernel;
        timing->result_in_out = (int)(desc->iterations & 1u);
        timing->ngpus = 1;
        timing->pipelined = 1;
    }
    return SB200_OK;
}

int sb200_pipeline_schedule(const sb200_desc* desc, sb200_pipe_op* ops_out, uint32_t capacity, uint32_t* count) {
    if (int rc = validate(desc)) return rc;
    if (!count || (capacity && !ops_out)) return fail(SB200_ERR_INVALID, "bad argument");
    const Choice c = choose(desc);
    if (int rc = check_choice(desc, c)) return rc;
    const Grid g = make_grid(desc, 0);
    const size_t dense = (size_t)(g.n[0] + 2 * g.r) * g.pitch_y * g.planes * esize(desc->dtype);
    const int C = pipeline_chunks(desc, c, g, dense);
    *count = 0;
    if (C <= 1) return SB200_OK;  // not pipelined: upload, sweep, download back to back
    std::vector<sb200_pipe_op> ops;
    build_pipeline(desc, c, g, C, pipeline_streams(), ops);
    *count = (uint32_t)ops.size();
    if (ops.size() > capacity) return capacity ? fail(SB200_ERR_INVALID, "schedule has %zu ops, capacity %u", ops.size(), capacity) : SB200_OK;
    memcpy(ops_out, ops.data(), ops.size() * sizeof(sb200_pipe_op));
    return SB200_OK;
}

// ---- one call, N GPUs ---------------------------------------------------------------------
// Slabs along the outermost axis with T*r planes of deep halo per internal side, read from
// the caller's host array (the host holds the whole grid, so the halo costs PCIe bytes and
// redundant sweeps of the shrinking cone, but no GPU ever waits for another one and every slab
// keeps the pipelined upload / sweep / download schedule on its own PCIe link).
static sb200_desc slab_desc(const sb200_desc* d, long long planes) {
    sb200_desc s = *d;
    s.n[d->ndim - 1] = (uint64_t)planes;
    s.ngpus = 1;
    return s;
}

static int slab_plan_impl(const sb200_desc* d, int ngpus, sb200_slab_plan* plan) {
    memset(plan, 0, sizeof *plan);
    const long long n = d->ndim >= 2 ? (long long)d->n[d->ndim - 1] : 1;
    const long long h = (long long)d->iterations * d->radius;
    plan->deep_halo = (uint32_t)h;
    const Choice c = choose(d);
    if (int rc = check_choice(d, c)) return rc;
    int G = ngpus < 1 ? 1 : (ngpus > SB200_MAX_GPUS ? SB200_MAX_GPUS : ngpus);
    if (d->ndim < 2 || d->iterations == 0 || c.kernel == SB200_KERNEL_RESIDENT2D) G = 1;
    for (; G > 1; --G) {
        // thick enough for the halo to stay a minority, and worth pipelining on its own
        if (n / G < 2 * h || n / G < 8LL * c.k * d->radius) continue;
        const long long sub = n / G + 2 * h;
        const sb200_desc sd = slab_desc(d, sub);
        const Grid g = make_grid(&sd, 0);
        const size_t dense = (size_t)(g.n[0] + 2 * g.r) * g.pitch_y * g.planes * esize(d->dtype);
        if (pipeline_chunks(&sd, c, g, dense) > 1) break;
    }
    plan->ngpus_used = (uint32_t)G;
    for (int i = 0; i < G; ++i) {
        const long long lo = n * i / G, hi = n * (i + 1) / G;
        plan->slab[i].own_lo = (uint64_t)lo;
        plan->slab[i].own_hi = (uint64_t)hi;
        plan->slab[i].sub_lo = (uint64_t)(G > 1 && lo - h > 0 ? lo - h : 0);
        plan->slab[i].sub_hi = (uint64_t)(G > 1 && hi + h < n ? hi + h : n);
    }
    return SB200_OK;
}

int sb200_slab_plan_for(const sb200_desc* desc, int ngpus, sb200_slab_plan* plan) {
    if (int rc = validate(desc)) return rc;
    if (!plan) return fail(SB200_ERR_INVALID, "null plan");
    return slab_plan_impl(desc, ngpus, plan);
}

int sb200_pipeline_schedule_slab(const sb200_desc* desc, int ngpus, int slab, sb200_pipe_op* ops_out,
                                 uint32_t capacity, uint32_t* count, uint32_t* zone_uploads) {
    if (int rc = validate(desc)) return rc;
    if (!count || (capacity && !ops_out)) return fail(SB200_ERR_INVALID, "bad argument");
    sb200_slab_plan plan;
    if (int rc = slab_plan_impl(desc, ngpus, &plan)) return rc;
    if (slab < 0 || slab >= (int)plan.ngpus_used) return fail(SB200_ERR_INVALID, "slab %d of %u", slab, plan.ngpus_used);
    *count = 0;
    if (zone_uploads) *zone_uploads = 0;
    if (plan.ngpus_used < 2) return sb200_pipeline_schedule(desc, ops_out, capacity, count);
    const auto& sl = plan.slab[slab];
    const sb200_desc sd = slab_desc(desc, (long long)(sl.sub_hi - sl.sub_lo));
    const Choice c = choose(&sd);
    const Grid g = make_grid(&sd, 0);
    const size_t dense = (size_t)(g.n[0] + 2 * g.r) * g.pitch_y * g.planes * esize(sd.dtype);
    const int C = pipeline_chunks(&sd, c, g, dense);
    std::vector<sb200_pipe_op> ops;
    uint32_t zones = 0;
    build_pipeline(&sd, c, g, C, pipeline_streams(), ops,
                   {(long long)(sl.own_lo - sl.sub_lo), (long long)(sl.own_hi - sl.sub_lo)}, &zones);
    if (zone_uploads) *zone_uploads = zones;
    *count = (uint32_t)ops.size();
    if (ops.size() > capacity) return capacity ? fail(SB200_ERR_INVALID, "schedule has %zu ops, capacity %u", ops.size(), capacity) : SB200_OK;
    memcpy(ops_out, ops.data(), ops.size() * sizeof(sb200_pipe_op));
    return SB200_OK;
}

int sb200_upload(const sb200_desc* desc, uint32_t ghost, const void* host, void* dev, void* stream) {
    if (int rc = validate(desc)) return rc;
    if (!host || !dev) return fail(SB200_ERR_INVALID, "null buffer");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    return copy_grid(desc, ghost, host, dev, true, false, (cudaStream_t)stream);
}

int sb200_download(const sb200_desc* desc, uint32_t ghost, const void* dev, void* host, void* stream) {
    if (int rc = validate(desc)) return rc;
    if (!host || !dev) return fail(SB200_ERR_INVALID, "null buffer");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    return copy_grid(desc, ghost, dev, host, false, false, (cudaStream_t)stream);
}

static int checksum_impl(const sb200_desc* desc, uint32_t ghost, const void* grid, uint64_t z_offset, uint64_t z_total,
                         cudaStream_t st, uint64_t* out);

int sb200_checksum_device(const sb200_desc* desc, uint32_t ghost, const void* grid, uint64_t z_offset,
                          uint64_t z_total, void* stream, uint64_t* out) {
    if (int rc = validate(desc)) return rc;
    if (!grid || !out) return fail(SB200_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    return checksum_impl(desc, ghost, grid, z_offset, z_total, (cudaStream_t)stream, out);
}

static int checksum_impl(const sb200_desc* desc, uint32_t ghost, const void* grid, uint64_t z_offset, uint64_t z_total,
                         cudaStream_t st, uint64_t* out) {
    const Grid g = make_grid(desc, ghost);
    long long gn[3] = {g.n[0], g.n[1], g.n[2]};
    if (z_total) gn[desc->ndim - 1] = (long long)z_total;
    CU_TRY(cudaMemsetAsync(DS().d_acc, 0, sizeof(unsigned long long), st));
    const long long total = g.n[0] * g.n[1] * g.n[2];
    long long blocks = (total + 255) / 256;
    if (blocks > DS().sm_count * 16) blocks = DS().sm_count * 16;
    if (desc->dtype == SB200_F64)
        checksum_kernel<double><<<(unsigned)blocks, 256, 0, st>>>((const double*)grid, g, gn[0], gn[1], (long long)z_offset, DS().d_acc);
    else
        checksum_kernel<float><<<(unsigned)blocks, 256, 0, st>>>((const float*)grid, g, gn[0], gn[1], (long long)z_offset, DS().d_acc);
    CU_TRY(cudaGetLastError());
    unsigned long long h = 0;
    CU_TRY(cudaMemcpyAsync(&h, DS().d_acc, sizeof h, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    *out = (uint64_t)h;
    return SB200_OK;
}

// ---- device-resident sweep over the device group: one process, N GPUs, no torch -------------
// The C++ spelling of what stencil_b200/slab.py does with one process per GPU: slabs of the
// outermost axis with ghost = r*k planes / rows, one linked launch per slab and fused block
// (peer stores over NVLink + arrival counters, slablink.cuh).  Inside one process the
// neighbours' buffers are reachable after cudaDeviceEnablePeerAccess — no IPC handles — and one
// host thread issues every launch (they are asynchronous; block b of every slab is enqueued
// before block b+1 of any, which is the order the arrival counters assume).
struct GroupSlab {
    State* st = nullptr;
    void* buf[2] = {nullptr, nullptr};
    uint32_t* flags = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    uint32_t arrivals = 0;
};

static void group_release(std::vector<GroupSlab>& slabs) {
    for (GroupSlab& gs : slabs) {
        if (!gs.st) continue;
        tls_state = gs.st;
        cudaSetDevice(gs.st->device);
        cudaDeviceSynchronize();
        for (void* p : gs.buf)
            if (p) {
                for (size_t i = 0; i < gs.st->maps.size();)
                    if (gs.st->maps[i].key.ptr == p) gs.st->maps.erase(gs.st->maps.begin() + (long)i); else ++i;
                cudaFree(p);
            }
        if (gs.flags) cudaFree(gs.flags);
        if (gs.e0) cudaEventDestroy(gs.e0);
        if (gs.e1) cudaEventDestroy(gs.e1);
    }
    tls_state = nullptr;
    if (!g_group.empty()) cudaSetDevice(g_group[0]->device);
}

static int group_sweep_impl(const sb200_desc* desc, int init, uint64_t seed, uint32_t repeats, int G,
                            std::vector<GroupSlab>& slabs, sb200_group_result* res) {
    const Choice c = choose(desc);
    if (int rc = check_choice(desc, c)) return rc;
    const int nd = desc->ndim;
    const uint64_t n_outer = desc->n[nd - 1];
    const uint32_t T = desc->iterations, r = desc->radius;
    memset(res, 0, sizeof *res);
    res->ngpus = (uint32_t)G;
    res->kernel = c.kernel;
    res->time_block = (uint32_t)c.k;
    sb200_desc sd = *desc;
    uint32_t ghost = 0;
    if (G > 1) {
        const bool linkable = c.kernel == SB200_KERNEL_WARP2D || c.kernel == SB200_KERNEL_TMA3D;
        if (!linkable)
            return fail(SB200_ERR_UNSUPPORTED, "the fused halo exchange needs the WARP2D or the TMA3D family; this "
                                               "descriptor runs on one GPU only");
        if (n_outer % (uint64_t)G) return fail(SB200_ERR_INVALID, "outermost extent %llu is not a multiple of %d GPUs",
                                               (unsigned long long)n_outer, G);
        sd.n[nd - 1] = n_outer / (uint64_t)G;
        ghost = r * (uint32_t)c.k;
        if (sd.n[nd - 1] < 2ull * ghost)
            return fail(SB200_ERR_INVALID, "slabs of %llu planes are thinner than two ghost zones (%u each)",
                        (unsigned long long)sd.n[nd - 1], ghost);
        res->exchange = 1;
    }
    const Grid g = make_grid(&sd, ghost);
    const size_t bytes = (size_t)g.elems() * esize(sd.dtype);
    // allocate, map the neighbours, fill
    for (int i = 0; i < G; ++i) {
        GroupSlab& gs = slabs[(size_t)i];
        gs.st = g_group[(size_t)i];
        tls_state = gs.st;
        CU_TRY(cudaSetDevice(gs.st->device));
        for (int nb : {i - 1, i + 1})
            if (nb >= 0 && nb < G && g_group[(size_t)nb]->device != gs.st->device) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(g_group[(size_t)nb]->device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                    return fail(SB200_ERR_CUDA, "no peer access from GPU %d to GPU %d: %s", gs.st->device,
                                g_group[(size_t)nb]->device, cudaGetErrorString(e));
                cudaGetLastError();
            }
        for (void*& p : gs.buf) {
            CU_TRY(cudaMalloc(&p, bytes));
            CU_TRY(cudaMemsetAsync(p, 0, bytes, 0));
        }
        CU_TRY(cudaMalloc((void**)&gs.flags, sizeof(uint32_t) * SB200_LINK_FLAG_WORDS));
        CU_TRY(cudaMemsetAsync(gs.flags, 0, sizeof(uint32_t) * SB200_LINK_FLAG_WORDS, 0));
        CU_TRY(cudaEventCreate(&gs.e0));
        CU_TRY(cudaEventCreate(&gs.e1));
    }
    auto for_each = [&](auto&& fn) -> int {
        for (int i = 0; i < G; ++i) {
            tls_state = slabs[(size_t)i].st;
            CU_TRY(cudaSetDevice(slabs[(size_t)i].st->device));
            if (int rc = fn(i, slabs[(size_t)i])) return rc;
        }
        return SB200_OK;
    };
    auto sync_all = [&]() { return for_each([](int, GroupSlab&) { CU_TRY(cudaDeviceSynchronize()); return (int)SB200_OK; }); };
    const uint64_t off_unit = sd.n[nd - 1], z_total = G > 1 ? n_outer : 0;
    auto fill = [&]() {
        return for_each([&](int i, GroupSlab& gs) {
            const uint64_t off = G > 1 ? off_unit * (uint64_t)i : 0;
            if (int rc = fill_impl(&sd, ghost, gs.buf[0], init, seed, off, z_total, 0)) return rc;
            return fill_impl(&sd, ghost, gs.buf[1], SB200_INIT_REF, 0, off, z_total, 0);
        });
    };
    int cur = 0;
    uint32_t launches = 0;
    auto sweep_T = [&]() -> int {
        if (G == 1) {
            tls_state = slabs[0].st;
            CU_TRY(cudaSetDevice(slabs[0].st->device));
            int in_b = 0;
            uint32_t nl = 0;
            if (int rc = sweep_impl(&sd, 0, 1, 1, slabs[0].buf[cur], slabs[0].buf[cur ^ 1], T, 0, &in_b, &nl, nullptr, nullptr))
                return rc;
            if (in_b) cur ^= 1;
            launches += nl;
            return SB200_OK;
        }
        for (uint32_t done = 0; done < T;) {
            const uint32_t k = T - done < (uint32_t)c.k ? T - done : (uint32_t)c.k;
            if (int rc = for_each([&](int i, GroupSlab& gs) {
                    sb200_slab_link lk = {};
                    lk.peer_dst_lo = i > 0 ? slabs[(size_t)i - 1].buf[cur ^ 1] : nullptr;
                    lk.peer_dst_hi = i + 1 < G ? slabs[(size_t)i + 1].buf[cur ^ 1] : nullptr;
                    lk.flags_local = gs.flags;
                    lk.flags_lo = i > 0 ? slabs[(size_t)i - 1].flags : nullptr;
                    lk.flags_hi = i + 1 < G ? slabs[(size_t)i + 1].flags : nullptr;
                    lk.wait_arrivals = gs.arrivals;
                    uint32_t arr = 0;
                    if (int rc = slab_linked_impl(&sd, ghost, i == 0, i + 1 == G, gs.buf[cur], gs.buf[cur ^ 1], k, &lk, nullptr, &arr))
                        return rc;
                    gs.arrivals += arr;
                    ++launches;
                    return (int)SB200_OK;
                }))
                return rc;
            cur ^= 1;
            done += k;
        }
        return SB200_OK;
    };
    if (int rc = fill()) return rc;
    if (int rc = sync_all()) return rc;
    // the digest after exactly T steps from the initial condition
    if (int rc = sweep_T()) return rc;
    if (int rc = sync_all()) return rc;
    uint64_t total = 0;
    if (int rc = for_each([&](int i, GroupSlab& gs) {
            uint64_t part = 0;
            if (int rc = checksum_impl(&sd, ghost, gs.buf[cur], G > 1 ? off_unit * (uint64_t)i : 0, z_total, 0, &part)) return rc;
            total += part;
            if (G > 1) {
                uint32_t words[SB200_LINK_FLAG_WORDS];
                CU_TRY(cudaMemcpy(words, gs.flags, sizeof words, cudaMemcpyDeviceToHost));
                if (words[2]) return fail(SB200_ERR_CUDA, "GPU %d: a fused halo exchange timed out waiting for a neighbour", gs.st->device);
            }
            return (int)SB200_OK;
        }))
        return rc;
    res->checksum = total;
    res->launches = launches;
    // timed: `repeats` further sweeps of T steps, device time = max over the GPUs
    if (repeats) {
        if (int rc = for_each([](int, GroupSlab& gs) { CU_TRY(cudaEventRecord(gs.e0, 0)); return (int)SB200_OK; })) return rc;
        for (uint32_t rep = 0; rep < repeats; ++rep)
            if (int rc = sweep_T()) return rc;
        if (int rc = for_each([](int, GroupSlab& gs) { CU_TRY(cudaEventRecord(gs.e1, 0)); return (int)SB200_OK; })) return rc;
        if (int rc = sync_all()) return rc;
        for (GroupSlab& gs : slabs) {
            float ms = 0;
            CU_TRY(cudaEventElapsedTime(&ms, gs.e0, gs.e1));
            if (ms > res->sweep_ms) res->sweep_ms = ms;
            if (G > 1) {
                uint32_t words[SB200_LINK_FLAG_WORDS];
                CU_TRY(cudaSetDevice(gs.st->device));
                CU_TRY(cudaMemcpy(words, gs.flags, sizeof words, cudaMemcpyDeviceToHost));
                if (words[2]) return fail(SB200_ERR_CUDA, "GPU %d: a fused halo exchange timed out waiting for a neighbour", gs.st->device);
            }
        }
        res->sweep_ms /= repeats;
    }
    return SB200_OK;
}

int sb200_group_sweep(const sb200_desc* desc, int init, uint64_t seed, uint32_t repeats, sb200_group_result* res) {
    if (int rc = validate(desc)) return rc;
    if (!res) return fail(SB200_ERR_INVALID, "null result");
    if (init != SB200_INIT_REF && init != SB200_INIT_RANDOM) return fail(SB200_ERR_INVALID, "bad init %d", init);
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    const int have = (int)g_group.size();
    const int G = desc->ngpus == 0 ? have : (int)desc->ngpus;
    if (G > have) return fail(SB200_ERR_INVALID, "ngpus = %d but %d device(s) are bound (sb200_init_multi)", G, have);
    if (desc->ndim == 1 && G > 1) return fail(SB200_ERR_INVALID, "1-D grids are not slab-decomposed");
    std::vector<GroupSlab> slabs((size_t)G);
    const int rc = group_sweep_impl(desc, init, seed, repeats, G, slabs, res);
    group_release(slabs);
    return rc;
}

int sb200_checksum_device_noalloc(const sb200_desc* desc, uint32_t ghost, const void* grid, uint64_t z_offset,
                                  uint64_t z_total, void* stream, uint64_t* out) {
    if (int rc = validate(desc)) return rc;
    if (!grid || !out) return fail(SB200_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(g_mu);
    // binding a device allocates (context, streams): this entry point must not be the one to do it
    if (!DS().ready)
        return fail(SB200_ERR_INVALID, "sb200_checksum_device_noalloc allocates nothing: call sb200_init first");
    if (desc->ndim == 1 && ghost) return fail(SB200_ERR_INVALID, "1-D grids cannot carry ghost cells");
    const Grid g = make_grid(desc, ghost);
    const long long total = g.n[0] * g.n[1] * g.n[2];
    const long long span = SB200_NOALLOC_POINTS_PER_BLOCK;
    const long long blocks = (total + span - 1) / span;
    if (blocks > SB200_NOALLOC_MAX_BLOCKS)
        return fail(SB200_ERR_INVALID,
                    "grid of %lld points exceeds the no-allocation limit of %lld points (%d partial sums of %d "
                    "points in the static scratch array); use sb200_checksum_device",
                    total, (long long)SB200_NOALLOC_MAX_BLOCKS * span, SB200_NOALLOC_MAX_BLOCKS,
                    SB200_NOALLOC_POINTS_PER_BLOCK);
    State& D = DS();
    CU_TRY(cudaSetDevice(D.device));
    if (!D.noalloc_scratch) {
        // once per device, with that device current: the device address behind the symbol
        unsigned long long* p = nullptr;
        CU_TRY(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) return fail(SB200_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        D.noalloc_scratch = p;
    }
    long long gn[3] = {g.n[0], g.n[1], g.n[2]};
    if (z_total) gn[desc->ndim - 1] = (long long)z_total;
    cudaStream_t st = (cudaStream_t)stream;
    // the array keeps the previous call's contents: ticket and result start at zero; every
    // partial this launch reads is written by it first
    CU_TRY(cudaMemsetAsync(D.noalloc_scratch, 0, 2 * sizeof(unsigned long long), st));
    if (desc->dtype == SB200_F64)
        checksum_noalloc_kernel<double><<<(unsigned)blocks, 256, 0, st>>>((const double*)grid, g, gn[0], gn[1], (long long)z_offset, span, D.noalloc_scratch);
    else
        checksum_noalloc_kernel<float><<<(unsigned)blocks, 256, 0, st>>>((const float*)grid, g, gn[0], gn[1], (long long)z_offset, span, D.noalloc_scratch);
    CU_TRY(cudaGetLastError());
    unsigned long long h = 0;
    CU_TRY(cudaMemcpyAsync(&h, D.noalloc_scratch + 1, sizeof h, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    *out = (uint64_t)h;
    return SB200_OK;
}

int sb200_host_alloc(void** ptr, size_t bytes) {
    if (!ptr) return fail(SB200_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    // portable: page-locked for every device of the group, not only the current one
    CU_TRY(cudaHostAlloc(ptr, bytes, cudaHostAllocPortable));
    return SB200_OK;
}

// ---- NUMA placement of host grids (no libnuma on the image: raw mbind) ----------------------
static int gpu_numa_node(int dev) {
    char bus[32];
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, dev) != cudaSuccess) {
        cudaGetLastError();
        return -1;
    }
    for (char* p = bus; *p; ++p) *p = (char)tolower((unsigned char)*p);
    char path[96];
    snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", bus);
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    int node = -1;
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
    return node;
}

// Prefer memory node `node` for the pages inside [p, p + bytes) (rounded inwards to pages);
// move == true also migrates pages that already exist.
static bool prefer_node(char* p, size_t bytes, int node, bool move) {
    const uintptr_t page = (uintptr_t)sysconf(_SC_PAGESIZE);
    const uintptr_t lo = ((uintptr_t)p + page - 1) / page * page, hi = ((uintptr_t)p + bytes) / page * page;
    if (node < 0 || node >= 1024 || hi <= lo) return false;
    unsigned long mask[16] = {0};
    mask[node / 64] = 1ul << (node % 64);
    return syscall(SYS_mbind, (void*)lo, (size_t)(hi - lo), 1 /* MPOL_PREFERRED */, mask, 1024ul,
                   move ? 2u /* MPOL_MF_MOVE */ : 0u) == 0;
}

static size_t dense_host_bytes(const sb200_desc* d) {
    size_t b = esize(d->dtype);
    for (int a = 0; a < d->ndim; ++a) b *= (size_t)(d->n[a] + 2 * d->radius);
    return b;
}

// Put the planes each slab owns on its GPU's memory node.  Returns the number of distinct nodes.
static int place_slabs(const sb200_desc* d, char* base, bool move) {
    const int have = (int)g_group.size();
    const int want = d->ngpus == 0 ? have : (int)d->ngpus < have ? (int)d->ngpus : have;
    sb200_slab_plan plan;
    if (want < 2 || slab_plan_impl(d, want, &plan) != SB200_OK || plan.ngpus_used < 2) return 0;
    const size_t outer_bytes = dense_host_bytes(d) / (size_t)(d->n[d->ndim - 1] + 2 * d->radius);
    int nodes[SB200_MAX_GPUS], distinct = 0;
    for (uint32_t i = 0; i < plan.ngpus_used; ++i) {
        nodes[i] = gpu_numa_node(g_group[i]->device);
        // padded index = interior index + radius; the end slabs take the halo planes along
        const size_t lo = i == 0 ? 0 : (size_t)(plan.slab[i].own_lo + d->radius) * outer_bytes;
        const size_t hi = i + 1 == plan.ngpus_used ? dense_host_bytes(d) : (size_t)(plan.slab[i].own_hi + d->radius) * outer_bytes;
        if (!prefer_node(base + lo, hi - lo, nodes[i], move)) continue;
        bool seen = false;
        for (uint32_t j = 0; j < i; ++j) seen = seen || nodes[j] == nodes[i];
        distinct += !seen;
    }
    return distinct;
}

struct PlacedAlloc {
    void* ptr;
    size_t bytes;
};
static std::vector<PlacedAlloc> g_placed;  // sb200_host_alloc_for allocations (mmap + register)

int sb200_host_alloc_for(const sb200_desc* desc, void** ptr, int* nodes_used) {
    if (int rc = validate(desc)) return rc;
    if (!ptr) return fail(SB200_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    const size_t bytes = dense_host_bytes(desc);
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED) return fail(SB200_ERR_INVALID, "mmap of %zu bytes failed", bytes);
    const int nodes = place_slabs(desc, (char*)p, false);  // policy first, pages appear at registration
    if (cudaHostRegister(p, bytes, cudaHostRegisterPortable) != cudaSuccess) {
        const cudaError_t e = cudaGetLastError();
        munmap(p, bytes);
        return fail(SB200_ERR_CUDA, "cudaHostRegister of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    g_placed.push_back({p, bytes});
    *ptr = p;
    if (nodes_used) *nodes_used = nodes;
    return SB200_OK;
}

int sb200_host_place(const sb200_desc* desc, void* ptr, int* nodes_used) {
    if (int rc = validate(desc)) return rc;
    if (!ptr) return fail(SB200_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    const int nodes = place_slabs(desc, (char*)ptr, true);
    if (nodes_used) *nodes_used = nodes;
    return SB200_OK;
}

int sb200_host_free(void* ptr) {
    if (!ptr) return SB200_OK;
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    for (size_t i = 0; i < g_placed.size(); ++i)
        if (g_placed[i].ptr == ptr) {
            const size_t bytes = g_placed[i].bytes;
            g_placed.erase(g_placed.begin() + (long)i);
            CU_TRY(cudaHostUnregister(ptr));
            munmap(ptr, bytes);
            return SB200_OK;
        }
    CU_TRY(cudaFreeHost(ptr));
    return SB200_OK;
}

int sb200_host_register(void* ptr, size_t bytes) {
    if (!ptr || bytes == 0) return fail(SB200_ERR_INVALID, "bad host range");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    CU_TRY(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return SB200_OK;
}

int sb200_host_unregister(void* ptr) {
    if (!ptr) return SB200_OK;
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    CU_TRY(cudaHostUnregister(ptr));
    return SB200_OK;
}

static int run_host_impl(const sb200_desc* desc, void* in, void* out, sb200_timing* timing);
static int run_host_multi(const sb200_desc* desc, const sb200_slab_plan& plan, void* in, void* out,
                          sb200_timing* timing);

// Both ping-pong grids of the host path on the calling thread's device (cached across calls).
static int ensure_grids(size_t bytes) {
    State& D = DS();
    if (D.buf_bytes >= bytes) return SB200_OK;
    for (int i = 0; i < 2; ++i) {
        if (D.buf[i]) cudaFree(D.buf[i]);
        D.buf[i] = nullptr;
    }
    D.buf_bytes = 0;
    D.maps.clear();
    CU_TRY(cudaMalloc(&D.buf[0], bytes));
    CU_TRY(cudaMalloc(&D.buf[1], bytes));
    D.buf_bytes = bytes;
    return SB200_OK;
}

// The precondition of the host path: `out` carries the same halo as `in`.  Checked on a
// sample of halo cells spread over every face (a full compare would read both arrays once
// more on the host).  Returns the number of sampled cells that differ.
static int halo_sample_mismatches(const sb200_desc* d, const void* in, const void* out) {
    const long long r = d->radius;
    long long ext[3] = {1, 1, 1}, stride[3] = {0, 0, 0};
    long long s = 1;
    for (int a = 0; a < d->ndim; ++a) {
        ext[a] = (long long)d->n[a] + 2 * r;
        stride[a] = s;
        s *= ext[a];
    }
    const size_t es = esize(d->dtype);
    int bad = 0;
    for (int a = 0; a < d->ndim; ++a)
        for (int side = 0; side < 2; ++side)
            for (int smp = 0; smp < 16; ++smp) {
                // a pseudo-random cell of this face: coordinate `a` inside the halo, the rest anywhere
                uint64_t h = mix64_seeded(0x5b200u + (uint64_t)a * 2 + side, (uint64_t)smp);
                long long off = 0;
                for (int b = 0; b < d->ndim; ++b) {
                    long long c;
                    if (b == a) c = (side ? ext[b] - r : 0) + (long long)(h % (uint64_t)r);
                    else c = (long long)(h % (uint64_t)ext[b]);
                    h = mix64(h);
                    off += c * stride[b];
                }
                bad += memcmp((const char*)in + (size_t)off * es, (const char*)out + (size_t)off * es, es) != 0;
            }
    return bad;
}

int sb200_run_host(const sb200_desc* desc, void* in, void* out, sb200_timing* timing) {
    if (int rc = validate(desc)) return rc;
    if (!in || !out) return fail(SB200_ERR_INVALID, "null host buffer");
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    if (const int bad = halo_sample_mismatches(desc, in, out))
        return fail(SB200_ERR_INVALID, "the halo of `out` differs from the halo of `in` (%d of the sampled cells): "
                                       "both buffers must carry the same Dirichlet halo", bad);
    tls_state = nullptr;  // the caller's thread works on the first device of the group
    const int have = (int)g_group.size();
    const int want = desc->ngpus == 0 ? have : (int)desc->ngpus;
    if (want > have)
        return fail(SB200_ERR_INVALID, "ngpus = %d but %d device(s) are bound (sb200_init_multi)", want, have);
    int rc;
    sb200_slab_plan plan;
    if (want > 1 && (rc = slab_plan_impl(desc, want, &plan)) != SB200_OK) return rc;
    if (want > 1 && plan.ngpus_used > 1) {
        rc = run_host_multi(desc, plan, in, out, timing);
    } else {
        CU_TRY(cudaSetDevice(DS().device));
        rc = run_host_impl(desc, in, out, timing);
        // an error return may leave a copy from / to the caller's host buffers in flight
        if (rc != SB200_OK) cudaDeviceSynchronize();
    }
    return rc;
}

// Untimed preparation of everything sb200_run_host(desc) needs, on every device it will use:
// ping-pong grids, the staging image, kernel attributes, and one warm-up launch of each
// kernel variant the call will take (fused depth k and the remainder T mod k).
static int prepare_on_current(const sb200_desc* sd) {
    const Grid g = make_grid(sd, 0);
    const size_t es = esize(sd->dtype);
    if (int rc = ensure_grids((size_t)g.elems() * es)) return rc;
    State& D = DS();
    const Choice c = choose(sd);
    if (int rc = check_choice(sd, c)) return rc;
    const size_t dense = (size_t)(g.n[0] + 2 * g.r) * g.pitch_y * g.planes * es;
    if (pipeline_chunks(sd, c, g, dense) > 1 && D.staging_bytes < dense) {
        if (D.staging) cudaFree(D.staging);
        D.staging = nullptr;
        D.staging_bytes = 0;
        CU_TRY(cudaMalloc(&D.staging, dense));
        D.staging_bytes = dense;
    }
    CU_TRY(cudaMemsetAsync(D.buf[0], 0, (size_t)g.elems() * es, 0));
    CU_TRY(cudaMemsetAsync(D.buf[1], 0, (size_t)g.elems() * es, 0));
    const uint32_t T = sd->iterations;
    if (T > 0) {
        const bool resident = c.kernel == SB200_KERNEL_RESIDENT2D ||
                              (sd->kernel == SB200_KERNEL_AUTO && sd->time_block == 0 && c.kernel == SB200_KERNEL_WARP2D &&
                               sd->radius == 1 && T >= 16 && resident2d_fits(g.n[0], g.n[1], (int)es));
        if (resident) {
            CU_TRY(launch_resident2d(D.buf[0], D.buf[1], g, sd->shape, (int)es, 1, 0));
        } else {
            const Domain dom = make_domain(g, 1, 1);
            const long long n = outer_n(g);
            const long long oe = n < 8LL * c.k * g.r ? n : 8LL * c.k * g.r;
            const uint32_t ks[2] = {T < (uint32_t)c.k ? T : (uint32_t)c.k, T % (uint32_t)c.k};
            for (uint32_t k : ks)
                if (k)
                    if (int rc = pass(sd, c, g, dom, D.buf[0], D.buf[1], (int)k, 0, oe, 0)) return rc;
        }
    }
    CU_TRY(cudaDeviceSynchronize());
    return SB200_OK;
}

int sb200_prepare(const sb200_desc* desc) {
    if (int rc = validate(desc)) return rc;
    std::lock_guard<std::mutex> lk(g_mu);
    if (int rc = ensure_ready()) return rc;
    const int have = (int)g_group.size();
    const int want = desc->ngpus == 0 ? have : (int)desc->ngpus;
    if (want > have)
        return fail(SB200_ERR_INVALID, "ngpus = %d but %d device(s) are bound (sb200_init_multi)", want, have);
    sb200_slab_plan plan;
    if (int rc = slab_plan_impl(desc, want, &plan)) return rc;
    int rc = SB200_OK;
    for (uint32_t i = 0; i < plan.ngpus_used && rc == SB200_OK; ++i) {
        const sb200_desc sd = plan.ngpus_used > 1
                                  ? slab_desc(desc, (long long)(plan.slab[i].sub_hi - plan.slab[i].sub_lo))
                                  : *desc;
        tls_state = g_group[i];
        rc = cudaSetDevice(g_group[i]->device) == cudaSuccess ? prepare_on_current(&sd)
                                                               : fail(SB200_ERR_CUDA, "cudaSetDevice(%d) failed", g_group[i]->device);
    }
    tls_state = nullptr;
    cudaSetDevice(g_group[0]->device);
    return rc;
}

// Host-side rendezvous of the slab worker threads (C++17: no std::barrier).
struct HostBarrier {
    std::mutex mu;
    std::condition_variable cv;
    int expected, arrived = 0;
    explicit HostBarrier(int n) : expected(n) {}
    void arrive_and_wait() {
        std::unique_lock<std::mutex> lk(mu);
        ++arrived;
        cv.notify_all();
        cv.wait(lk, [&] { return arrived >= expected; });
    }
    void abandon() {  // a worker that failed before the rendezvous must not hang the others
        std::lock_guard<std::mutex> lk(mu);
        --expected;
        cv.notify_all();
    }
};

struct SlabWorker {
    State* st = nullptr;
    sb200_desc sd;
    void* in = nullptr;
    void* out = nullptr;
    SlabRange own{0, -1};
    SlabSync sync;
    HostBarrier* barrier = nullptr;
    bool met = false;
    sb200_timing tm;
    int rc = SB200_OK;
    char err[560];
};

static void slab_barrier_cb(void* arg) {
    SlabWorker* w = (SlabWorker*)arg;
    w->met = true;
    w->barrier->arrive_and_wait();
}

static int slab_worker_body(SlabWorker& w) {
    tls_state = w.st;
    CU_TRY(cudaSetDevice(w.st->device));
    const Grid g = make_grid(&w.sd, 0);
    if (int rc = ensure_grids((size_t)g.elems() * esize(w.sd.dtype))) return rc;
    const Choice c = choose(&w.sd);
    const size_t dense = (size_t)(g.n[0] + 2 * g.r) * g.pitch_y * g.planes * esize(w.sd.dtype);
    const int C = pipeline_chunks(&w.sd, c, g, dense);
    if (C < 2) return fail(SB200_ERR_INVALID, "slab of %lld planes cannot be pipelined", outer_n(g));
    return run_host_pipelined(&w.sd, c, w.in, w.out, &w.tm, C, w.own, &w.sync);
}

static void slab_worker_main(SlabWorker* w) {
    memset(&w->tm, 0, sizeof w->tm);
    w->rc = slab_worker_body(*w);
    if (w->rc != SB200_OK) {
        snprintf(w->err, sizeof w->err, "GPU %d: %.500s", w->st->device, g_err);
        if (!w->met) w->barrier->abandon();
        cudaDeviceSynchronize();  // nothing of this slab may still touch the caller's buffers
    }
    tls_state = nullptr;
}

static int run_host_multi(const sb200_desc* desc, const sb200_slab_plan& plan, void* in, void* out,
                          sb200_timing* timing) {
    const auto w0 = std::chrono::steady_clock::now();
    const int G = (int)plan.ngpus_used;
    const long long r = desc->radius;
    const size_t es = esize(desc->dtype);
    // dense host bytes per outermost index (host layout: boundary_matrix.hpp:54-59,115-123)
    const size_t outer_bytes = (size_t)(desc->n[0] + 2 * r) * (desc->ndim == 3 ? (size_t)(desc->n[1] + 2 * r) : 1) * es;
    HostBarrier barrier(G);
    std::vector<SlabWorker> ws((size_t)G);
    Events events;  // the slabs' zones_read events; destroyed after the threads have joined
    for (int i = 0; i < G; ++i) {
        SlabWorker& w = ws[(size_t)i];
        const auto& sl = plan.slab[i];
        w.st = g_group[(size_t)i];
        w.sd = slab_desc(desc, (long long)(sl.sub_hi - sl.sub_lo));
        w.in = (char*)in + (size_t)sl.sub_lo * outer_bytes;
        w.out = (char*)out + (size_t)sl.sub_lo * outer_bytes;
        w.own = {(long long)(sl.own_lo - sl.sub_lo), (long long)(sl.own_hi - sl.sub_lo)};
        w.barrier = &barrier;
        w.sync.host_barrier = slab_barrier_cb;
        w.sync.barrier_arg = &w;
        CU_TRY(cudaSetDevice(w.st->device));
        w.sync.zones_read = events.make();
        CU_TRY(events.err);
        {   // how many leading uploads cover the deep-halo zones (same builder the worker runs)
            const Choice c = choose(&w.sd);
            const Grid g = make_grid(&w.sd, 0);
            const size_t dense = (size_t)(g.n[0] + 2 * g.r) * g.pitch_y * g.planes * es;
            const int C = pipeline_chunks(&w.sd, c, g, dense);
            std::vector<sb200_pipe_op> ops;
            if (C > 1) build_pipeline(&w.sd, c, g, C, pipeline_streams(), ops, w.own, &w.sync.zone_uploads);
        }
    }
    for (int i = 0; i < G; ++i) {
        if (i > 0) ws[(size_t)i].sync.wait[0] = ws[(size_t)i - 1].sync.zones_read;
        if (i + 1 < G) ws[(size_t)i].sync.wait[1] = ws[(size_t)i + 1].sync.zones_read;
    }
    CU_TRY(cudaSetDevice(g_group[0]->device));
    std::vector<std::thread> threads;
    for (int i = 0; i < G; ++i) threads.emplace_back(slab_worker_main, &ws[(size_t)i]);
    for (auto& t : threads) t.join();
    const auto w1 = std::chrono::steady_clock::now();
    for (const SlabWorker& w : ws)
        if (w.rc != SB200_OK) return fail(w.rc, "%s", w.err);
    if (timing) {
        memset(timing, 0, sizeof *timing);
        for (const SlabWorker& w : ws) {
            timing->h2d_ms = w.tm.h2d_ms > timing->h2d_ms ? w.tm.h2d_ms : timing->h2d_ms;
            timing->sweep_ms = w.tm.sweep_ms > timing->sweep_ms ? w.tm.sweep_ms : timing->sweep_ms;
            timing->d2h_ms = w.tm.d2h_ms > timing->d2h_ms ? w.tm.d2h_ms : timing->d2h_ms;
            timing->launches += w.tm.launches;
        }
        timing->wall_ms = std::chrono::duration<double, std::milli>(w1 - w0).count();
        timing->time_block = ws[0].tm.time_block;
        timing->kernel = ws[0].tm.kernel;
        timing->result_in_out = (int)(desc->iterations & 1u);
        timing->ngpus = (uint32_t)G;
        timing->pipelined = 1;
    }
    return SB200_OK;
}

static int run_host_impl(const sb200_desc* desc, void* in, void* out, sb200_timing* timing) {
    const auto w0 = std::chrono::steady_clock::now();
    const Grid g = make_grid(desc, 0);
    const size_t bytes = (size_t)g.elems() * esize(desc->dtype);
    if (int rc = ensure_grids(bytes)) return rc;
    {
        const Choice c = choose(desc);
        if (int rc = check_choice(desc, c)) return rc;
        const long long hx0 = g.n[0] + 2 * g.r;
        const size_t dense0 = (size_t)hx0 * g.pitch_y * g.planes * esize(desc->dtype);
        const int C = pipeline_chunks(desc, c, g, dense0);
        if (C > 1) return run_host_pipelined(desc, c, in, out, timing, C);
    }
    cudaStream_t st = 0;
    Events events;
    cudaEvent_t ev[4];
    for (auto& e : ev) e = events.make(cudaEventDefault);
    CU_TRY(events.err);
    CU_TRY(cudaEventRecord(ev[0], st));
    // Both reference buffers carry the same Dirichlet halo by construction
    // (Stencil::generate_initialized_matrix, stencil.cpp:190-207): upload `in` once and
    // clone it on the device rather than moving `out` across PCIe as well.  The padding
    // columns beyond the halo are zeroed once so no uninitialised bits are ever read.
    // One CONTIGUOUS host->device copy of the dense image into buf[1] (used as staging; a
    // strided 3-D copy reaches only ~60 % of the PCIe rate), then one kernel re-pitches it
    // into both buffers.
    {
        const long long hx = g.n[0] + 2 * g.r;
        const size_t dense_bytes = (size_t)hx * g.pitch_y * g.planes * esize(desc->dtype);
        void* staging = (char*)DS().buf[1] + (bytes - dense_bytes);  // tail of buf[1]
        CU_TRY(cudaMemcpyAsync(staging, in, dense_bytes, cudaMemcpyHostToDevice, st));
        // the staging image sits at the tail of buf[1]; unpack into buf[0], then clone
        // buf[0] over buf[1] (both reference matrices carry the same halo, stencil.cpp:190-207)
        const long long rows = g.pitch_y * g.planes;
        dim3 grid((unsigned)((g.pitch_x + 255) / 256), (unsigned)(rows < 65535 ? rows : 65535));
        if (desc->dtype == SB200_F64)
            unpack_dense_kernel<double><<<grid, 256, 0, st>>>((const double*)staging, (double*)DS().buf[0], g, hx);
        else
            unpack_dense_kernel<float><<<grid, 256, 0, st>>>((const float*)staging, (float*)DS().buf[0], g, hx);
        CU_TRY(cudaGetLastError());
        CU_TRY(cudaMemcpyAsync(DS().buf[1], DS().buf[0], bytes, cudaMemcpyDeviceToDevice, st));
    }
    CU_TRY(cudaEventRecord(ev[1], st));
    int in_b = 0;
    uint32_t nl = 0, used_k = 0;
    int used_kernel = 0;
    if (int rc = sweep_impl(desc, 0, 1, 1, DS().buf[0], DS().buf[1], desc->iterations, st, &in_b, &nl, &used_k, &used_kernel))
        return rc;
    CU_TRY(cudaEventRecord(ev[2], st));
    // Step T belongs in `out` when T is odd, in `in` when even (stencil.cpp:88-92,129-134).
    void* host_dst = (desc->iterations & 1u) ? out : in;
    if (desc->iterations > 0)
        if (int rc = copy_grid(desc, 0, DS().buf[in_b], host_dst, false, true, st)) return rc;
    CU_TRY(cudaEventRecord(ev[3], st));
    CU_TRY(cudaStreamSynchronize(st));
    const auto w1 = std::chrono::steady_clock::now();
    if (timing) {
        float ms = 0;
        cudaEventElapsedTime(&ms, ev[0], ev[1]); timing->h2d_ms = ms;
        cudaEventElapsedTime(&ms, ev[1], ev[2]); timing->sweep_ms = ms;
        cudaEventElapsedTime(&ms, ev[2], ev[3]); timing->d2h_ms = ms;
        timing->wall_ms = std::chrono::duration<double, std::milli>(w1 - w0).count();
        timing->launches = nl;
        timing->time_block = used_k;
        timing->kernel = used_kernel;
        timing->result_in_out = (int)(desc->iterations & 1u);
        timing->ngpus = 1;
        timing->pipelined = 0;
    }
    return SB200_OK;
}

}  // extern "C"
