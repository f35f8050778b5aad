// slablink.cuh — device-side primitives of the fused halo exchange between slabs (shared by the
// 3-D kernel, tma3d.cuh, and the 2-D kernel, warp2d.cuh).
#pragma once
#include "common.cuh"

namespace sb200 {

// ---- slab link: face planes stored straight into the neighbours' ghost planes ----------
// Multi-GPU slabs (one process per GPU).  No extra CTAs and no extra launch: the CTAs (3-D;
// warps in 2-D, where "plane" reads "row") of the slab's first chunk also own its low face
// (planes [0, g)), those of the last chunk its high face ([n-g, n)).  A face plane is stored twice: into the local grid and, over NVLink,
// into the ghost planes of the neighbour's destination buffer (peer-mapped pointer).
// The push happens at the END of the CTA (the plane loop itself carries no link code), which
// is why the last and the first chunk are dispatched FIRST; arrival is signalled with one
// release-add per CTA on a counter in the neighbour's memory, and the neighbour's
// face-owning CTAs of the NEXT launch acquire the counter before they touch their ghosts.  The
// same counter orders the write-after-read hazard: a CTA signals only after its last read of
// the ghost planes the neighbour's next launch overwrites.  This is the B200 counterpart of
// the reference's RMA halo puts + reply counters (stencil_rma.cpp:198-255).
struct SlabLinkDev {
    int link_lo, link_hi;     // low / high end of the slab is an internal face
    int face_planes;          // g
    int nchunks;              // z chunks of this launch
    void* peer_out_lo;        // lower / upper neighbour's destination buffer (same layout)
    void* peer_out_hi;
    long long shift_lo;       // element offset of the same plane in that buffer
    long long shift_hi;
    unsigned* wait_lo;        // local arrival counters fed by the lower / upper neighbour
    unsigned* wait_hi;
    unsigned* signal_lo;      // the neighbours' counters this rank feeds
    unsigned* signal_hi;
    unsigned wait_value;      // arrivals required before this launch may read its ghosts
    unsigned* err_flag;       // set to 1 if a wait timed out (never hangs the GPU)
};

__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_sys_add_u32(unsigned* p, unsigned v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// Bounded acquire-spin on an arrival counter: gives up after 10 s, and at once if another
// CTA already gave up — a broken link ends in a reported error (sb200_link_status), never
// in a hung GPU.
static __device__ __noinline__ void link_wait(const unsigned* flag, unsigned want, unsigned* err_flag) {
    const unsigned long long t0 = global_timer_ns();
    while ((int)(ld_acquire_sys_u32(flag) - want) < 0) {
        if (global_timer_ns() - t0 > 10000000000ull || *(volatile unsigned*)err_flag != 0u) {
            *(volatile unsigned*)err_flag = 1u;
            break;
        }
        __nanosleep(100);
    }
}

}  // namespace sb200
