/* stencil_oracle.c — CPU ORACLE (test infrastructure, NOT the product).
 *
 * A plain-C restatement of the reference's naive sweep,
 *   Stencil::check_result            /root/reference/src/stencil/stencil.cpp:75-131
 *   Stencil::generate_initialized_matrix            ...stencil.cpp:190-207
 *   BoundaryMatrix layout / fill_boundary   include/stencil/boundary_matrix.hpp:54-59,115-123,129-170
 * generalised to 1-D/2-D/3-D, star/box, f32/f64 exactly as SURVEY.md §8(c) defines
 * (the d=2 star f32 case reduces to the reference loop line by line).
 *
 * Parity status:
 *   2-D star f32, any radius : PINNED — bit-compared against the reference's own
 *       check_result compiled from /root/reference (oracle/ref_shim, tests/golden/).
 *   2-D star f64, any radius : pinned against the SAME reference sources compiled with
 *       float -> double (oracle/Makefile: _ref/libref64.so; tests/golden/ref_2d_f64.npz) —
 *       the reference's loop nest and add order, at a precision it does not ship.
 *       Also held, within tolerance (another add order), against grids from the reference's own
 *       DMA / DMASlavePack / RMA kernels run on the athread emulator (oracle/ref_shim/emu/,
 *       tests/golden/ref_emu_2d_f32.npz).
 *   1-D, 3-D, box            : "parity unpinned" — the reference has no such code;
 *       this file is the definition.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (stencil_b200/) never does.
 *
 * Layout: dense row-major, x fastest, `r` halo cells on both sides of every used
 * axis: P[a] = n[a] + 2r.  Halo cells are never written (stencil.cpp:94-101 only
 * visits the interior), so Dirichlet values persist in both buffers.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_STAR 0
#define ORACLE_BOX 1
#define ORACLE_F32 0
#define ORACLE_F64 1
#define ORACLE_INIT_REF 0
#define ORACLE_INIT_RANDOM 1

/* splitmix64 finaliser; the counter-based generator of SURVEY.md §8(c). */
static inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline uint64_t mix64_seeded(uint64_t seed, uint64_t idx) {
    return mix64(seed + (idx + 1) * 0x9E3779B97F4A7C15ull);
}
uint64_t oracle_mix64(uint64_t z) { return mix64(z); }
uint64_t oracle_mix64_seeded(uint64_t seed, uint64_t idx) { return mix64_seeded(seed, idx); }

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

typedef struct dims {
    int ndim;
    size_t n[3], p[3]; /* interior / padded extents */
    size_t h[3];       /* halo per axis (0 on unused axes) */
} dims;

static void make_dims(dims* d, int ndim, const uint64_t* n, unsigned r) {
    d->ndim = ndim;
    for (int a = 0; a < 3; ++a) {
        d->n[a] = a < ndim ? (size_t)n[a] : 1;
        d->h[a] = a < ndim ? r : 0;
        d->p[a] = d->n[a] + 2 * d->h[a];
    }
}

size_t oracle_padded_elems(int ndim, const uint64_t* n, unsigned r) {
    dims d;
    make_dims(&d, ndim, n, r);
    return d.p[0] * d.p[1] * d.p[2];
}

/* ---- initialisation -----------------------------------------------------------
 * "ref": interior 0 (make_unique<T[]> value-init, boundary_matrix.hpp:59); the two
 * x faces := 1 over the FULL padded extent of the other axes (fill_boundary Left /
 * Right, boundary_matrix.hpp:132-145; stencil.cpp:199-204); everything else 0.
 * "random": interior u = (mix64(seed, padded linear index of the GLOBAL grid) >> 11)
 * * 2^-53 cast to T; halo as "ref".  z_offset/z_total place a slab (outermost used
 * axis) inside a larger global grid so slabs can be generated independently. */
#define DEFINE_FILL(NAME, T)                                                              \
    void NAME(int ndim, const uint64_t* n, unsigned r, int init, uint64_t seed,           \
              uint64_t z_offset, uint64_t z_total, T* g) {                                \
        dims d;                                                                           \
        make_dims(&d, ndim, n, r);                                                        \
        const int oa = ndim - 1; /* outermost used axis */                                \
        size_t gp[3] = {d.p[0], d.p[1], d.p[2]};                                          \
        if (z_total) gp[oa] = (size_t)z_total + 2 * d.h[oa];                              \
        for (size_t z = 0; z < d.p[2]; ++z)                                               \
            for (size_t y = 0; y < d.p[1]; ++y)                                           \
                for (size_t x = 0; x < d.p[0]; ++x) {                                     \
                    size_t c[3] = {x, y, z};                                              \
                    size_t gc[3] = {x, y, z};                                             \
                    gc[oa] += (size_t)z_offset;                                           \
                    int interior = 1;                                                     \
                    for (int a = 0; a < ndim; ++a)                                        \
                        if (gc[a] < d.h[a] || gc[a] >= gp[a] - d.h[a]) interior = 0;      \
                    T v = (T)0;                                                           \
                    if (!interior) {                                                      \
                        if (gc[0] < d.h[0] || gc[0] >= gp[0] - d.h[0]) v = (T)1;          \
                    } else if (init == ORACLE_INIT_RANDOM) {                              \
                        uint64_t lin = ((uint64_t)gc[2] * gp[1] + gc[1]) * gp[0] + gc[0]; \
                        v = (T)((double)(mix64_seeded(seed, lin) >> 11) * 0x1.0p-53);     \
                    }                                                                     \
                    g[(c[2] * d.p[1] + c[1]) * d.p[0] + c[0]] = v;                        \
                }                                                                         \
    }
DEFINE_FILL(oracle_fill_f32, float)
DEFINE_FILL(oracle_fill_f64, double)

/* ---- one sweep ------------------------------------------------------------------
 * star (stencil.cpp:103-125): sum = 0; per axis x, then y, then z: low side far->near
 * (k = r..1), then high side near->far (k = 1..r); out = sum * avg with
 * avg = 1 / T(2*d*r) (stencil.cpp:85-86: 1.f / float((bw+bh)*2)).
 * box: sum over the full (2r+1)^d cube INCLUDING the centre, z outer -> y -> x inner,
 * ascending; out = sum * (1 / T((2r+1)^d)).
 * The OpenMP split is over output rows/planes only: per-point arithmetic is identical
 * for any thread count. */
#define DEFINE_STEP(NAME, T)                                                              \
    static void NAME(const dims* d, int shape, unsigned r, const T* in, T* out,           \
                     int threads) {                                                       \
        const ptrdiff_t sx = 1, sy = (ptrdiff_t)d->p[0],                                  \
                        sz = (ptrdiff_t)(d->p[0] * d->p[1]);                              \
        const ptrdiff_t st[3] = {sx, sy, sz};                                             \
        const int nd = d->ndim;                                                           \
        const ptrdiff_t R = (ptrdiff_t)r;                                                 \
        T avg;                                                                            \
        if (shape == ORACLE_STAR) {                                                       \
            avg = (T)1 / (T)(2 * nd * r);                                                 \
        } else {                                                                          \
            unsigned cnt = 1;                                                             \
            for (int a = 0; a < nd; ++a) cnt *= (2 * r + 1);                              \
            avg = (T)1 / (T)cnt;                                                          \
        }                                                                                 \
        const ptrdiff_t rows = (ptrdiff_t)(d->n[1] * d->n[2]);                            \
        (void)threads;                                                                    \
        _Pragma("omp parallel for schedule(static) num_threads(threads) if (threads > 1)") \
        for (ptrdiff_t row = 0; row < rows; ++row) {                                      \
            const size_t z = (size_t)row / d->n[1], y = (size_t)row % d->n[1];            \
            const size_t base = ((z + d->h[2]) * d->p[1] + (y + d->h[1])) * d->p[0] + d->h[0]; \
            for (size_t x = 0; x < d->n[0]; ++x) {                                        \
                const T* c = in + base + x;                                               \
                T sum = (T)0;                                                             \
                if (shape == ORACLE_STAR) {                                               \
                    for (int a = 0; a < nd; ++a) {                                        \
                        for (ptrdiff_t k = R; k >= 1; --k) sum += c[-k * st[a]];          \
                        for (ptrdiff_t k = 1; k <= R; ++k) sum += c[k * st[a]];           \
                    }                                                                     \
                } else {                                                                  \
                    const ptrdiff_t rz = nd > 2 ? R : 0, ry = nd > 1 ? R : 0;             \
                    for (ptrdiff_t dz = -rz; dz <= rz; ++dz)                              \
                        for (ptrdiff_t dy = -ry; dy <= ry; ++dy)                          \
                            for (ptrdiff_t dx = -R; dx <= R; ++dx)                        \
                                sum += c[dz * sz + dy * sy + dx];                         \
                }                                                                         \
                out[base + x] = sum * avg;                                                \
            }                                                                             \
        }                                                                                 \
    }
DEFINE_STEP(step_f32, float)
DEFINE_STEP(step_f64, double)

/* ---- T sweeps with pointer-swap double buffering (stencil.cpp:94,129-130) ---------
 * Returns 1 if step T ended in `b` (T odd), 0 if in `a` — the `swapped` flag of
 * stencil.cpp:88-92,134. */
#define DEFINE_SWEEP(NAME, STEP, T)                                                       \
    int NAME(int ndim, const uint64_t* n, unsigned r, int shape, unsigned iterations,     \
             T* a, T* b, int threads) {                                                   \
        dims d;                                                                           \
        make_dims(&d, ndim, n, r);                                                        \
        T *in = a, *out = b;                                                              \
        int swapped = 0;                                                                  \
        for (unsigned i = 0; i != iterations; ++i) {                                      \
            STEP(&d, shape, r, in, out, threads);                                         \
            T* t = in;                                                                    \
            in = out;                                                                     \
            out = t;                                                                      \
            swapped = !swapped;                                                           \
        }                                                                                 \
        return swapped;                                                                   \
    }
DEFINE_SWEEP(oracle_sweep_f32, step_f32, float)
DEFINE_SWEEP(oracle_sweep_f64, step_f64, double)

/* ---- order-independent checksum of the interior -------------------------------------
 * sum (mod 2^64) over interior points of mix64(bits(v) ^ mix64(global interior linear
 * index)); z_offset/z_total as in fill.  Same definition as sb200_checksum_device. */
#define DEFINE_CHECKSUM(NAME, T, U)                                                       \
    uint64_t NAME(int ndim, const uint64_t* n, unsigned r, const T* g, uint64_t z_offset, \
                  uint64_t z_total) {                                                     \
        dims d;                                                                           \
        make_dims(&d, ndim, n, r);                                                        \
        const int oa = ndim - 1;                                                          \
        size_t gn[3] = {d.n[0], d.n[1], d.n[2]};                                          \
        if (z_total) gn[oa] = (size_t)z_total;                                            \
        uint64_t acc = 0;                                                                 \
        for (size_t z = 0; z < d.n[2]; ++z)                                               \
            for (size_t y = 0; y < d.n[1]; ++y)                                           \
                for (size_t x = 0; x < d.n[0]; ++x) {                                     \
                    size_t gc[3] = {x, y, z};                                             \
                    gc[oa] += (size_t)z_offset;                                           \
                    uint64_t lin = ((uint64_t)gc[2] * gn[1] + gc[1]) * gn[0] + gc[0];     \
                    T v = g[((z + d.h[2]) * d.p[1] + (y + d.h[1])) * d.p[0] + x + d.h[0]]; \
                    U bits;                                                               \
                    memcpy(&bits, &v, sizeof bits);                                       \
                    acc += mix64((uint64_t)bits ^ mix64(lin));                            \
                }                                                                         \
        return acc;                                                                       \
    }
DEFINE_CHECKSUM(oracle_checksum_f32, float, uint32_t)
DEFINE_CHECKSUM(oracle_checksum_f64, double, uint64_t)
