/* stencil_b200.h — C ABI of the B200 (sm_100a) stencil-sweep backend.
 *
 * This is the drop-in boundary for the reference's accelerator call
 *   athread_spawn(SLAVE_FUN(stencil_iterate_X), &args); athread_join();
 * (reference: src/stencil/stencil.cpp:34-53, kernel declarations
 * src/stencil/slave/stencil_slave.hpp:26-46, argument block `struct Arguments`
 * src/stencil/slave/stencil_slave.hpp:13-24).  Plain pointers and sizes only;
 * no C++ or torch types cross this boundary.  Every function returns 0 on
 * success or a negative sb200_status; the message is in sb200_last_error().
 *
 * Grid convention (reference: include/stencil/boundary_matrix.hpp:54-59,115-123):
 * row-major, x fastest; every axis carries `radius` halo cells on both sides
 * which are never written (Dirichlet).  n[0]=x, n[1]=y, n[2]=z interior extents;
 * unused axes have n = 1 and no halo.
 */
#ifndef STENCIL_B200_H
#define STENCIL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SB200_ABI_VERSION 4 /* 2: slab link, resident2d, plan, host_register; 3: pipeline_schedule;
                               4: ngpus/flags in sb200_desc, sb200_init_multi, slab plans, prepare */

typedef enum sb200_status {
    SB200_OK = 0,
    SB200_ERR_INVALID = -1,     /* bad descriptor / argument */
    SB200_ERR_CUDA = -2,        /* a CUDA runtime/driver call failed */
    SB200_ERR_UNSUPPORTED = -3, /* valid request this backend cannot serve */
    SB200_ERR_NO_DEVICE = -4    /* no CUDA device: there is NO CPU fallback */
} sb200_status;

typedef enum sb200_shape { SB200_STAR = 0, SB200_BOX = 1 } sb200_shape;
typedef enum sb200_dtype { SB200_F32 = 0, SB200_F64 = 1 } sb200_dtype;
typedef enum sb200_init_kind { SB200_INIT_REF = 0, SB200_INIT_RANDOM = 1 } sb200_init_kind;

/* Kernel family selector (sb200_desc.kernel).  AUTO picks the fastest family
 * that supports the descriptor; the others force one (tests, experiments). */
typedef enum sb200_kernel {
    SB200_KERNEL_AUTO = 0,
    SB200_KERNEL_GENERIC = 1, /* one step per launch, any ndim/shape/radius   */
    SB200_KERNEL_WARP2D = 2,  /* 2-D r=1 (star/box) and r=2..4 (star): register/shuffle temporal blocking */
    SB200_KERNEL_TMA3D = 3,   /* 3-D r=1, TMA-staged z-streaming, temporal blk */
    SB200_KERNEL_RESIDENT2D = 4, /* small 2-D r=1 grids: register-resident in one
                                   thread-block cluster for ALL steps of a call, one
                                   launch (the analogue of the reference's RMA kernel,
                                   stencil_rma.cpp:333-458); AUTO picks it for
                                   single-GPU calls of >= 16 steps when the grid fits */
    SB200_KERNEL_LINE1D = 5     /* 1-D star/box r=1..4: a warp keeps 512 points in registers,
                                   up to 16 fused steps per launch */
} sb200_kernel;

/* Problem descriptor.  Replaces the reference's `Arguments{block_size,
 * iterations, input, output}` (stencil_slave.hpp:13-24) plus the shape data
 * held inside its BoundaryMatrixView members (boundary_matrix.hpp:221-226). */
typedef struct sb200_desc {
    int32_t ndim;        /* 1, 2 or 3 */
    int32_t shape;       /* sb200_shape */
    int32_t dtype;       /* sb200_dtype */
    uint32_t radius;     /* >= 1 */
    uint64_t n[3];       /* interior extents x,y,z */
    uint32_t iterations; /* T */
    uint32_t time_block; /* fused steps per HBM round trip; 0 = auto */
    uint32_t block_hint; /* reference -b/--block-size (program_options.cpp:16-18): the CUDA
                            backend reads it as a tile / time-block preference, see sb200_plan */
    int32_t kernel;      /* sb200_kernel */
    uint32_t ngpus;      /* sb200_run_host: GPUs of this process to spread the grid over, as
                            z-slabs (y-slabs in 2-D); 0 = every device bound by sb200_init /
                            sb200_init_multi.  Ignored by the device-resident calls. */
    uint32_t flags;      /* sb200_flags */
} sb200_desc;

typedef enum sb200_flags {
    SB200_FLAG_NONE = 0,
    /* Allow kernels that re-associate the sum of a point (separable box sums: the 27-point
     * TMA3D kernel and the 9-point WARP2D kernel have such a level): results then agree
     * with the reference order within the north-star tolerance (max relative error 1e-12
     * fp64 / 1e-5 fp32) instead of bit for bit.  Off by default; shapes without a
     * separable level (stars, radius > 1) stay bit-exact under it. */
    SB200_FLAG_REASSOCIATE = 1,
    /* Read block_hint (the reference's -b/--block-size: the side of the tile one CPE owns) as
     * this backend's tile preference: the number of rows (2-D) / planes (3-D) of the outermost
     * axis one warp (WARP2D) or CTA (TMA3D) streams, instead of the planner's own choice.
     * sb200_plan reports the resulting geometry.  Other kernel families ignore it. */
    SB200_FLAG_BLOCK_TILE = 2
} sb200_flags;

/* Timing breakdown of one call (milliseconds; device times from CUDA events).
 * Small grids run upload -> sweep -> download back to back and the three device
 * times add up.  Grids of >= 64 MB are pipelined (upload, sweep and download of
 * different (z,t) parallelograms overlap on separate streams); then h2d_ms is the
 * span of all upload chunks, d2h_ms the span from the first to the last download
 * chunk, and sweep_ms the span of the WHOLE device pipeline (copies included),
 * i.e. the three overlap and sweep_ms is the number to divide the updates by. */
typedef struct sb200_timing {
    double h2d_ms;
    double sweep_ms;
    double d2h_ms;
    double wall_ms;
    uint32_t launches;     /* kernels launched by the sweep */
    uint32_t time_block;   /* fused depth actually used */
    int32_t kernel;        /* family actually used */
    int32_t result_in_out; /* 1: step T is in `out`, 0: in `in` (T parity) */
    uint32_t ngpus;        /* GPUs the call was spread over (device times: max over them) */
    uint32_t pipelined;    /* 1: upload / sweep / download overlapped (see above) */
} sb200_timing;

/* Device-side grid layout chosen by the library for a descriptor: padded so the
 * first interior x is 128-byte aligned and the row pitch is a multiple of 128 B.
 * element (x,y,z) in interior coordinates (halo = negative / >= n) lives at
 *   base + ((z + oz) * pitch_y + (y + oy)) * pitch_x + (x + ox)   [elements]. */
typedef struct sb200_layout {
    uint64_t pitch_x;  /* elements per row */
    uint64_t pitch_y;  /* rows per plane */
    uint64_t planes;   /* planes allocated (incl. halo / ghost planes) */
    uint64_t ox, oy, oz;
    uint64_t bytes;    /* total allocation size */
} sb200_layout;

/* ---- library ---------------------------------------------------------------- */
int sb200_abi_version(void);
const char* sb200_last_error(void);
/* Number of visible CUDA devices (0 if none / driver missing). */
int sb200_device_count(void);
/* Bind the calling process to `device` (one process per GPU).  Takes the one-time costs
 * (context, module load, streams) so that the first timed call does not. */
int sb200_init(int device);
/* Bind the calling process to `n` devices at once (one process, N GPUs): the counterpart of
 * the reference's single Stencil::run driving the whole 64-CPE mesh from one host thread
 * (src/stencil/stencil.cpp:34-53; run.sh:3 `-cgsp 64`).  devices == NULL means 0..n-1.
 * A device may be listed more than once: its slabs then share it, each with its own grids
 * and streams (how the slab path is tested on a one-GPU box).
 * sb200_run_host then spreads a grid over them (sb200_desc.ngpus); the device-resident
 * calls keep operating on the first device of the group. */
int sb200_init_multi(int n, const int* devices);
/* Number of devices bound by sb200_init / sb200_init_multi (0 before). */
int sb200_group_size(void);
void sb200_finalize(void);

/* ---- the drop-in call -------------------------------------------------------
 * Host buffers `in`/`out`: padded arrays, (n[a] + 2*radius) per used axis, dense.
 * Runs T = desc->iterations double-buffered sweeps on the GPU.  On return the
 * buffer selected by T's parity holds step T (odd: out, even: in), exactly like
 * the reference kernels that swap their host views every step
 * (stencil_dma.cpp:556-557; oracle stencil.cpp:129-134).  Blocking.
 *
 * Precondition (checked on a sample, SB200_ERR_INVALID if violated): `out` carries the same
 * halo as `in` — true for the reference's generate_initialized_matrix (stencil.cpp:190-207),
 * which builds both matrices alike.  Only `in` crosses PCIe; its halo is cloned into both
 * device buffers and the interior of `out` is never read.
 *
 * With more than one device bound (sb200_init_multi) and desc->ngpus != 1 the grid is cut
 * into slabs along the outermost axis, one per GPU, see sb200_slab_plan.  The bits do not
 * depend on the number of GPUs. */
int sb200_run_host(const sb200_desc* desc, void* in, void* out, sb200_timing* timing);

/* Untimed preparation of sb200_run_host(desc): allocates the device grids and the staging
 * image on every GPU the call will use and launches each kernel variant it will take once,
 * so that a timed call right after measures the steady state (the reference does its
 * accelerator bring-up, athread_init(), before the timed runs too: src/main.cpp:54).
 * Optional: sb200_run_host does all of it on demand. */
int sb200_prepare(const sb200_desc* desc);

/* How sb200_run_host spreads `desc` over `ngpus` devices — pure host logic, no device needed.
 * Slab g OWNS outermost interior indices [own_lo, own_hi) and works on the sub-grid
 * [sub_lo, sub_hi) = its own planes plus T*radius planes of deep halo on every internal
 * side, read straight from the caller's host array: after t steps the sub-grid is exact on
 * [own_lo - (T-t)*radius, own_hi + (T-t)*radius), so the slabs never talk to each other and
 * each runs the pipelined (z,t)-parallelogram schedule on its own PCIe link.  The deep halo is
 * redundant upload and compute, so slabs thinner than 2*T*radius planes are not used:
 * ngpus_used is the largest count <= ngpus whose slabs are thick enough (1 = single-GPU path). */
#define SB200_MAX_GPUS 16
typedef struct sb200_slab_plan {
    uint32_t ngpus_used;
    uint32_t deep_halo; /* T * radius */
    struct {
        uint64_t sub_lo, sub_hi; /* sub-grid, outermost interior indices of the whole grid */
        uint64_t own_lo, own_hi; /* planes this slab downloads */
    } slab[SB200_MAX_GPUS];
} sb200_slab_plan;
int sb200_slab_plan_for(const sb200_desc* desc, int ngpus, sb200_slab_plan* plan);

/* ---- device-resident API (bench, multi-GPU slabs) ----------------------------
 * `ghost` = extra planes on each side of the outermost axis beyond `radius`
 * (slab decomposition: ghost planes hold neighbour data).  0 for single GPU. */
int sb200_layout_for(const sb200_desc* desc, uint32_t ghost, sb200_layout* layout);

/* Sweep `steps` time steps between two device buffers laid out per
 * sb200_layout_for(desc, ghost).  `lo_is_boundary`/`hi_is_boundary` say whether
 * the low/high end of the outermost axis is the global Dirichlet halo (1) or a
 * slab face with `ghost` valid neighbour planes (0).  With ghost planes, at most
 * ghost/radius steps can be taken before the ghosts must be refreshed; the call
 * fails with SB200_ERR_INVALID beyond that.  Work is enqueued on `stream`
 * (a cudaStream_t, or NULL for the legacy default stream) and NOT synchronised.
 * *result_in_b is set to 1 if the final state is in `b`, 0 if in `a`. */
int sb200_sweep_device(const sb200_desc* desc, uint32_t ghost, int lo_is_boundary,
                       int hi_is_boundary, void* a, void* b, uint32_t steps, void* stream,
                       int* result_in_b, uint32_t* launches);

/* Same, but restricted to output planes [z_begin, z_end) of the outermost axis
 * (interior coordinates of this slab).  Used to compute slab faces first and
 * overlap the halo exchange with the interior.  One fused block of `steps`
 * (<= max time block) only; a -> b. */
int sb200_sweep_device_range(const sb200_desc* desc, uint32_t ghost, int lo_is_boundary,
                             int hi_is_boundary, const void* a, void* b, uint32_t steps,
                             uint64_t z_begin, uint64_t z_end, void* stream,
                             uint32_t* launches);

/* Initialise a device grid (both halo and interior), reference init
 * (stencil.cpp:190-207: interior 0, x faces 1, other halo 0) or the counter-based
 * random init.  `z_offset`/`z_total` place a slab inside the global grid so every
 * slab can be generated independently. */
int sb200_fill_device(const sb200_desc* desc, uint32_t ghost, void* grid, int init,
                      uint64_t seed, uint64_t z_offset, uint64_t z_total, void* stream);

/* Copy between a dense host-layout padded array and the device layout. */
int sb200_upload(const sb200_desc* desc, uint32_t ghost, const void* host, void* dev,
                 void* stream);
int sb200_download(const sb200_desc* desc, uint32_t ghost, const void* dev, void* host,
                   void* stream);

/* Order-independent 64-bit checksum of the interior of a device grid:
 * sum over points of mix64(bits(value) ^ mix64(global linear index)).
 * Used for bit-exact 1-vs-N GPU and full-size parity checks.  Blocking. */
int sb200_checksum_device(const sb200_desc* desc, uint32_t ghost, const void* grid,
                          uint64_t z_offset, uint64_t z_total, void* stream, uint64_t* out);

/* The same digest as sb200_checksum_device for callers that cannot allocate (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE
 * REFERENCE.  The call performs no cudaMalloc /
 * cudaFree and calls no allocating library routine: its scratch (one partial sum per block, a
 * ticket and the result) is a file-scope `static __device__` array of
 * 8 * (32 + SB200_NOALLOC_MAX_BLOCKS) = 65 792 bytes that every GPU loading the library holds
 * until the process exits.  Limits and rules:
 *   - at most SB200_NOALLOC_MAX_BLOCKS * SB200_NOALLOC_POINTS_PER_BLOCK = 134 217 728 interior
 *     points per call (512^3); larger grids are rejected with SB200_ERR_INVALID;
 *   - the library must already be bound to a device (sb200_init / sb200_init_multi): binding
 *     allocates, so this call returns SB200_ERR_INVALID instead of doing it;
 *   - every call shares the one array (it keeps the last call's contents; the call clears what
 *     must start at zero): calls that OVERLAP on the GPU — different streams, different host
 *     threads' streams — race on it.  Issue them one after another.
 * Same arguments and result as sb200_checksum_device; the caller passes the grid.  Blocking. */
#define SB200_NOALLOC_MAX_BLOCKS 8192
#define SB200_NOALLOC_POINTS_PER_BLOCK 16384
int sb200_checksum_device_noalloc(const sb200_desc* desc, uint32_t ghost, const void* grid,
                                  uint64_t z_offset, uint64_t z_total, void* stream, uint64_t* out);

/* Pinned host memory for callers that want full-speed PCIe copies in sb200_run_host. */
int sb200_host_alloc(void** ptr, size_t bytes);
int sb200_host_free(void* ptr);
/* Page-lock caller-owned host memory in place (for callers that keep their own
 * allocations, like the reference's `make_unique<T[]>` grids,
 * boundary_matrix.hpp:59): without it the copies of sb200_run_host are staged
 * by the driver and cannot overlap with the sweep.  Costs ~0.1 ms per MB once. */
int sb200_host_register(void* ptr, size_t bytes);
int sb200_host_unregister(void* ptr);
/* NUMA placement for multi-GPU calls: the planes slab g owns are put on the memory node GPU g
 * hangs off (PCIe root of its socket), so that no upload or download crosses the socket
 * interconnect.  sb200_host_alloc_for returns page-locked memory for one dense grid of `desc`
 * placed that way (free it with sb200_host_free); sb200_host_place MOVES the pages of a
 * caller-owned grid (call it before sb200_host_register).  Both are hints: on a single-node
 * host, or where the kernel refuses (containers without CAP_SYS_NICE), they change nothing and
 * still succeed.  *nodes_used (may be NULL) = number of distinct memory nodes chosen. */
int sb200_host_alloc_for(const sb200_desc* desc, void** ptr, int* nodes_used);
int sb200_host_place(const sb200_desc* desc, void* ptr, int* nodes_used);

/* ---- multi-GPU slab link: halo exchange fused into the sweep kernel ------------
 * One process per GPU.  Replaces the reference's on-chip RMA halo puts + reply
 * counters (src/stencil/slave/stencil_rma.cpp:198-255) at GPU scale: the face
 * CTAs of a launch store their planes into the local grid AND straight into the
 * ghost planes of the neighbours' destination buffers over NVLink (peer-mapped
 * pointers obtained with the sb200_ipc_* calls), then count themselves in on an
 * arrival counter in the neighbour's memory; the neighbour's next launch acquires
 * that counter before it reads its ghost planes.  No NCCL call, no extra launch
 * and no host synchronisation on the data path.
 *
 * All ranks must own slabs of identical shape, issue the same sequence of linked
 * launches, and must not touch a buffer from the host side (fill, free) while a
 * neighbour may still be sweeping (barrier first). */
#define SB200_IPC_HANDLE_BYTES 64
#define SB200_LINK_FLAG_WORDS 4 /* [0] arrivals from the lower neighbour, [1] from the
                                   upper one, [2] set to 1 if a wait timed out, [3] spare */
typedef struct sb200_slab_link {
    void* peer_dst_lo;      /* lower neighbour's DESTINATION buffer of this launch, NULL at the global boundary */
    void* peer_dst_hi;      /* upper neighbour's */
    uint32_t* flags_local;  /* SB200_LINK_FLAG_WORDS words in local device memory, zeroed once at setup */
    uint32_t* flags_lo;     /* lower neighbour's flag words (peer-mapped) */
    uint32_t* flags_hi;     /* upper neighbour's */
    uint32_t wait_arrivals; /* sum of *arrivals returned by all previous linked launches */
} sb200_slab_link;

/* Zero-filled device memory from cudaMalloc (exportable, unlike sub-allocations). */
int sb200_device_alloc(void** ptr, size_t bytes);
int sb200_device_free(void* ptr);
/* CUDA IPC: export a sb200_device_alloc pointer / map a neighbour's export. */
int sb200_ipc_export(void* dev_ptr, unsigned char handle[SB200_IPC_HANDLE_BYTES]);
int sb200_ipc_open(const unsigned char handle[SB200_IPC_HANDLE_BYTES], void** dev_ptr);
int sb200_ipc_close(void* dev_ptr);

/* One fused block of `steps` (<= max time block) steps a -> b over the whole slab
 * in ONE launch: the CTAs of the first and last z chunk are dispatched first and
 * push the face planes to the neighbours (peer stores + arrival signal).
 * Needs ghost == radius * max-steps-between-exchanges >= radius * steps and a
 * slab of at least 2*ghost planes (rows in 2-D); the TMA3D family in 3-D, the WARP2D family
 * in 2-D (faces are rows, their owners warps — the reference's RMA kernel is the
 * 2-D case, stencil_rma.cpp:333-458).
 * *arrivals = face owners (CTAs in 3-D, warps in 2-D) per face of this launch (add it to
 * wait_arrivals). */
int sb200_sweep_slab_linked(const sb200_desc* desc, uint32_t ghost, int lo_is_boundary,
                            int hi_is_boundary, const void* a, void* b, uint32_t steps,
                            const sb200_slab_link* link, void* stream, uint32_t* arrivals);

/* Blocking read of the SB200_LINK_FLAG_WORDS flag words (after synchronising the
 * device): words[2] != 0 means a face CTA timed out (10 s) waiting for a neighbour; every
 * result since is invalid.  The flag is sticky until sb200_link_reset. */
int sb200_link_status(const uint32_t* flags_local, uint32_t words[SB200_LINK_FLAG_WORDS]);
/* Zero the flag words (arrival counters and the time-out flag).  Every rank of the chain must
 * do it at the same quiescent point (device idle, barrier on both sides) and restart its
 * wait_arrivals at 0. */
int sb200_link_reset(uint32_t* flags_local);

/* ---- device-resident sweep over the device group (one process, N GPUs) -----------------------
 * The whole of a multi-GPU device-resident run in ONE call, for C and C++ hosts that have no
 * process launcher: the grid of `desc` is cut into desc->ngpus (0 = all bound devices) equal
 * slabs along the outermost axis, generated on the devices (init / seed as sb200_fill_device),
 * swept desc->iterations steps with one linked launch per slab and fused block — the
 * neighbours' buffers are reached through CUDA peer access inside the process — and digested.
 * Then `repeats` further sweeps of the same length are timed.  This is the reference's RMA
 * method (tiles resident on the accelerator, neighbours exchanging halos directly, main memory
 * out of the loop: stencil_rma.cpp:333-458) across GPUs.  Needs a kernel family with the fused
 * exchange for ngpus > 1 (WARP2D, TMA3D), an outermost extent divisible by ngpus
 * and slabs of at least 2 * radius * time_block planes.  The digest equals
 * sb200_checksum_device's of the same grid swept on one GPU, bit for bit. */
typedef struct sb200_group_result {
    double sweep_ms;     /* device time of ONE timed sweep of `iterations` steps (max over GPUs) */
    uint64_t checksum;   /* whole-grid digest after the first `iterations` steps */
    uint32_t ngpus, time_block, launches;
    int32_t kernel;      /* sb200_kernel family */
    int32_t exchange;    /* 0: none (one GPU), 1: fused peer stores */
} sb200_group_result;
int sb200_group_sweep(const sb200_desc* desc, int init, uint64_t seed, uint32_t repeats,
                      sb200_group_result* res);

/* What the library WOULD do for `steps` steps of this descriptor on a whole grid of
 * a device with `sm_count` SMs — pure host logic, no device needed (tests, tuning,
 * the experiment driver's block-size column).  Meaning of the geometry fields:
 *   TMA3D:      tiles_x x tiles_y xy tiles, `chunks` z chunks of `chunk_extent` planes
 *   WARP2D:     tiles_x strips of 128 columns, `chunks` row chunks of `chunk_extent` rows
 *   RESIDENT2D: tiles_x x tiles_y warp tiles of 128 x chunk_extent points on `chunks` CTAs
 *   LINE1D:     tiles_x warp segments storing `chunk_extent` of their 512 points each
 *   GENERIC:    all zero */
typedef struct sb200_plan_info {
    int32_t kernel;           /* sb200_kernel family */
    uint32_t time_block;      /* fused steps per launch (RESIDENT2D: all of them) */
    uint32_t launches;        /* kernel launches for `steps` steps */
    uint32_t tiles_x, tiles_y, chunks, chunk_extent;
    uint32_t pipeline_chunks; /* (z,t) parallelograms of sb200_run_host; 1 = not pipelined */
} sb200_plan_info;
int sb200_plan(const sb200_desc* desc, uint32_t steps, int sm_count, sb200_plan_info* info);

/* The schedule sb200_run_host executes for a grid it pipelines (plan.pipeline_chunks > 1), as
 * plain data — pure host logic, no device needed.  Ops of one stream run in issue order (the
 * order of the list); `wait` adds cross-stream happens-before edges.  The device holds a dense
 * staging image of `in` plus two ping-pong grids (0, 1).  *count = 0 for a call that is not
 * pipelined.  With capacity = 0 only the op count is returned. */
typedef enum sb200_pipe_kind {
    SB200_PIPE_UPLOAD = 0,   /* host `in` -> staging, padded outermost indices [lo, hi)            */
    SB200_PIPE_UNPACK = 1,   /* staging -> grids 0 AND 1 (re-pitch), padded outermost [lo, hi)     */
    SB200_PIPE_SWEEP = 2,    /* `steps` fused steps grid src -> grid 1-src, interior outermost
                                output indices [lo, hi); reads [lo - steps*r, hi + steps*r) of src,
                                clipped to the grid incl. its Dirichlet halo                        */
    SB200_PIPE_DOWNLOAD = 3  /* grid src, interior outermost [lo, hi) -> the host result buffer     */
} sb200_pipe_kind;
typedef struct sb200_pipe_op {
    int32_t kind;          /* sb200_pipe_kind */
    int32_t stream;        /* 0 = copy-in, 1 = copy-out, 2.. = sweep streams */
    int32_t parallelogram; /* which (z,t) parallelogram the op belongs to */
    int32_t block;         /* SWEEP: index of the fused block, else -1 */
    uint32_t steps;        /* SWEEP: fused steps */
    int32_t src;           /* SWEEP / DOWNLOAD: ping-pong grid read */
    int64_t lo, hi;
    int32_t wait[2];       /* indices of ops that must complete first; -1 = none */
} sb200_pipe_op;
int sb200_pipeline_schedule(const sb200_desc* desc, sb200_pipe_op* ops, uint32_t capacity,
                            uint32_t* count);
/* The schedule of slab `slab` of sb200_slab_plan_for(desc, ngpus): indices are local to the
 * slab's sub-grid (add slab[].sub_lo for indices of the whole grid).  Differences to the
 * single-GPU schedule: sweeps and downloads are clipped to the dependency cone of the owned
 * planes, and the deep-halo zones (host planes another slab downloads into when T is even)
 * are uploaded FIRST; every slab's downloads start only after its neighbours' zone uploads
 * (`zone_uploads` = how many leading UPLOAD ops that is; a cross-device event at run time). */
int sb200_pipeline_schedule_slab(const sb200_desc* desc, int ngpus, int slab, sb200_pipe_op* ops,
                                 uint32_t capacity, uint32_t* count, uint32_t* zone_uploads);

/* Largest time block the chosen kernel family supports for this descriptor. */
int sb200_max_time_block(const sb200_desc* desc);

#ifdef __cplusplus
}
#endif
#endif /* STENCIL_B200_H */
