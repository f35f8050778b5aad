// stencil.cpp — host side: problem set-up, dispatch to the B200 backend, self-check.
//
// Mirrors the structure of the reference's src/stencil/stencil.cpp: initialize_matrix()
// (:14-21,190-207), run(method) timing only the accelerated sweep (:23-57), a name->method
// map (:59-73) and check_result() with the naive loop (:75-151).
#include <stencil/stencil.hpp>

#include <stencil_b200.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <utility>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

std::uint64_t mix64(std::uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

/// Problem definition. "ref" is Stencil::generate_initialized_matrix of the reference
/// (stencil.cpp:190-207): zero interior, left/right halo = 1 (full padded height), other
/// halo = 0.  "random" puts counter-based uniform [0,1) values in the interior.
template <class T>
void fill_initial(BoundaryMatrix<T>& m, bool random, std::uint64_t seed) {
    m.fill_boundary(BoundaryPosition::Left, T(1));
    m.fill_boundary(BoundaryPosition::Right, T(1));
    if (!random) return;
    auto const& s = m.shape();
    for (std::size_t z = 0; z < m.depth(); ++z)
        for (std::size_t y = 0; y < m.height(); ++y)
            for (std::size_t x = 0; x < m.width(); ++x) {
                std::uint64_t const lin =
                    (std::uint64_t(z + s.halo[2]) * s.padded(1) + (y + s.halo[1])) * s.padded(0) + (x + s.halo[0]);
                std::uint64_t const bits = mix64(seed + (lin + 1) * 0x9E3779B97F4A7C15ull);
                m.elem_at(z, y, x) = static_cast<T>(static_cast<double>(bits >> 11) * 0x1.0p-53);
            }
}

/// One naive sweep in the reference's summation order (stencil.cpp:103-125): per axis the
/// low side far->near then the high side near->far; box: full cube, z outer, x inner.
template <class T>
void naive_step(BoundaryMatrixView<T> in, BoundaryMatrixView<T> out, int ndim, bool box, unsigned r, bool parallel) {
    auto const& s = in.shape();
    std::ptrdiff_t const stride[3] = {1, std::ptrdiff_t(s.row_stride), std::ptrdiff_t(s.plane_stride)};
    unsigned count = 1;
    for (int a = 0; a < ndim; ++a) count *= 2 * r + 1;
    T const avg = box ? T(1) / T(count) : T(1) / T(2 * ndim * r);
    std::ptrdiff_t const R = r, rz = ndim > 2 ? R : 0, ry = ndim > 1 ? R : 0;
    std::ptrdiff_t const rows = std::ptrdiff_t(in.height() * in.depth());
    (void)parallel;
#pragma omp parallel for schedule(static) if (parallel)
    for (std::ptrdiff_t row = 0; row < rows; ++row) {
        std::size_t const z = std::size_t(row) / in.height(), y = std::size_t(row) % in.height();
        T const* src = &in.elem_at(z, y, 0);
        T* dst = &out.elem_at(z, y, 0);
        for (std::size_t x = 0; x < in.width(); ++x) {
            T const* c = src + x;
            T sum = T(0);
            if (!box) {
                for (int a = 0; a < ndim; ++a) {
                    for (std::ptrdiff_t k = R; k >= 1; --k) sum += c[-k * stride[a]];
                    for (std::ptrdiff_t k = 1; k <= R; ++k) sum += c[k * stride[a]];
                }
            } else {
                for (std::ptrdiff_t dz = -rz; dz <= rz; ++dz)
                    for (std::ptrdiff_t dy = -ry; dy <= ry; ++dy)
                        for (std::ptrdiff_t dx = -R; dx <= R; ++dx) sum += c[dz * stride[2] + dy * stride[1] + dx];
            }
            dst[x] = sum * avg;
        }
    }
}

template <class T>
bool naive_sweep(BoundaryMatrix<T>& a, BoundaryMatrix<T>& b, ProgramOptions const& o, bool parallel) {
    bool swapped = false;
    for (unsigned i = 0; i != o.iterations; ++i) {
        if (swapped) {
            naive_step<T>(b.borrow(), a.borrow(), o.ndim(), o.stencil == "box", o.radius, parallel);
        } else {
            naive_step<T>(a.borrow(), b.borrow(), o.ndim(), o.stencil == "box", o.radius, parallel);
        }
        swapped = !swapped;
    }
    return swapped;
}

sb200_desc make_desc(ProgramOptions const& o) {
    sb200_desc d;
    std::memset(&d, 0, sizeof d);
    d.ndim = o.ndim();
    d.shape = o.stencil == "box" ? SB200_BOX : SB200_STAR;
    d.dtype = o.dtype == "f64" ? SB200_F64 : SB200_F32;
    d.radius = o.radius;
    for (int a = 0; a < 3; ++a) d.n[a] = o.extent(a);
    d.iterations = o.iterations;
    d.time_block = o.time_block;
    d.block_hint = o.block_size;
    d.ngpus = o.gpus;
    d.flags = (o.block_as_tile ? SB200_FLAG_BLOCK_TILE : SB200_FLAG_NONE) | (o.reassociate ? SB200_FLAG_REASSOCIATE : SB200_FLAG_NONE);
    d.kernel = o.kernel == "generic" ? SB200_KERNEL_GENERIC
               : o.kernel == "warp2d" ? SB200_KERNEL_WARP2D
               : o.kernel == "tma3d"  ? SB200_KERNEL_TMA3D
               : o.kernel == "resident2d" ? SB200_KERNEL_RESIDENT2D
               : o.kernel == "line1d" ? SB200_KERNEL_LINE1D
                                      : SB200_KERNEL_AUTO;
    return d;
}

}  // namespace

bool Stencil::parse_method(std::string_view name, InputMethod& out) {
    static std::pair<std::string_view, InputMethod> const table[] = {
        {"CUDA", CUDA}, {"CPU", CPU}, {"CPUOMP", CPU_OMP}, {"DMA", DMA},
        {"DMAStaticUnroll", DMA_STATIC_UNROLL}, {"DMASlavePack", DMA_SLAVE_PACK}, {"RMA", RMA},
    };
    for (auto const& [n, m] : table)
        if (n == name) {
            out = m;
            return true;
        }
    return false;
}

bool Stencil::prepare_cuda(ProgramOptions const& options) {
    sb200_desc const d = make_desc(options);
    return sb200_prepare(&d) == SB200_OK;
}

template <class T>
void Stencil::initialize_typed() {
    Pair<T> p;
    std::array<std::size_t, 3> const ext {options.extent(0), options.extent(1), options.extent(2)};
    p.matrix = BoundaryMatrix<T>::make(options.ndim(), ext, options.radius);
    p.result = BoundaryMatrix<T>::make(options.ndim(), ext, options.radius);
    fill_initial(p.matrix, options.init == "random", options.seed);
    fill_initial(p.result, false, 0);
    grids = std::move(p);
}

void Stencil::initialize_matrix() {
    if (options.dtype == "f64") {
        initialize_typed<double>();
    } else {
        initialize_typed<float>();
    }
}

template <class T>
auto Stencil::run_typed(InputMethod method) -> std::chrono::steady_clock::duration {
    auto& p = std::get<Pair<T>>(grids);
    info_ = RunInfo {};
    // Part of the (untimed) initialisation, like generate_initialized_matrix in the reference
    // (stencil.cpp:25-31): page-lock large grids so the backend can overlap its copies with
    // the sweep.  Small grids are not worth the registration cost.
    std::size_t const bytes = p.matrix.width_with_boundary() * p.matrix.height_with_boundary() *
                              p.matrix.depth_with_boundary() * sizeof(T);
    if (method == CUDA && options.gpus > 1 && bytes >= (std::size_t(64) << 20)) {
        // every slab's planes onto the memory node of the GPU that sweeps them (a hint)
        sb200_desc const d = make_desc(options);
        (void)sb200_host_place(&d, p.matrix.raw_data(), nullptr);
        (void)sb200_host_place(&d, p.result.raw_data(), nullptr);
    }
    bool const pin = method == CUDA && bytes >= (std::size_t(64) << 20) &&
                     sb200_host_register(p.matrix.raw_data(), bytes) == SB200_OK;
    bool const pin2 = pin && sb200_host_register(p.result.raw_data(), bytes) == SB200_OK;
    struct Unpin {
        void* a;
        void* b;
        ~Unpin() {
            if (a) sb200_host_unregister(a);
            if (b) sb200_host_unregister(b);
        }
    } unpin {pin ? p.matrix.raw_data() : nullptr, pin2 ? p.result.raw_data() : nullptr};
    auto const start = std::chrono::steady_clock::now();
    switch (method) {
    case CUDA: {
        // The one call that replaces athread_spawn(...) + athread_join() (stencil.cpp:34-53).
        sb200_desc const d = make_desc(options);
        sb200_timing t;
        if (sb200_run_host(&d, p.matrix.raw_data(), p.result.raw_data(), &t) != SB200_OK)
            throw std::runtime_error(std::string("CUDA method failed: ") + sb200_last_error());
        info_.h2d_ms = t.h2d_ms;
        info_.sweep_ms = t.sweep_ms;
        info_.d2h_ms = t.d2h_ms;
        info_.launches = t.launches;
        info_.time_block = t.time_block;
        info_.kernel = t.kernel;
        info_.gpus = t.ngpus;
        info_.pipelined = t.pipelined != 0;
        break;
    }
    case CPU:
        naive_sweep(p.matrix, p.result, options, false);
        break;
    case CPU_OMP:
        naive_sweep(p.matrix, p.result, options, true);
#ifdef _OPENMP
        info_.threads = omp_get_max_threads();
#endif
        break;
    default:
        throw std::runtime_error("Sunway athread methods (DMA, DMAStaticUnroll, DMASlavePack, RMA) are not "
                                 "available on the B200 backend; use CUDA, CPU or CPUOMP");
    }
    return std::chrono::steady_clock::now() - start;
}

// The RMA idea at GPU scale (stencil_rma.cpp:333-458: tiles stay on the accelerator, halos go
// neighbour to neighbour, main memory is out of the loop): the grid is generated on the GPUs,
// swept there and digested; nothing crosses PCIe, so the host grids are not even allocated.
auto Stencil::run_resident() -> std::chrono::steady_clock::duration {
    info_ = RunInfo {};
    sb200_desc const d = make_desc(options);
    sb200_group_result res;
    if (sb200_group_sweep(&d, options.init == "random" ? SB200_INIT_RANDOM : SB200_INIT_REF, options.seed, 1, &res) != SB200_OK)
        throw std::runtime_error(std::string("CUDA method failed: ") + sb200_last_error());
    info_.sweep_ms = res.sweep_ms;
    info_.launches = res.launches;
    info_.time_block = res.time_block;
    info_.kernel = res.kernel;
    info_.gpus = res.ngpus;
    info_.resident = true;
    info_.checksum = res.checksum;
    return std::chrono::duration_cast<std::chrono::steady_clock::duration>(std::chrono::duration<double, std::milli>(res.sweep_ms));
}

/// Digest of the naive loop's result, the same function sb200_checksum_device computes: the sum
/// over interior points of mix64(bits(value) ^ mix64(linear interior index)).
template <class T>
unsigned long long Stencil::naive_digest() const {
    std::array<std::size_t, 3> const ext {options.extent(0), options.extent(1), options.extent(2)};
    auto a = BoundaryMatrix<T>::make(options.ndim(), ext, options.radius);
    auto b = BoundaryMatrix<T>::make(options.ndim(), ext, options.radius);
    fill_initial(a, options.init == "random", options.seed);
    fill_initial(b, false, 0);
    BoundaryMatrix<T> const& want = naive_sweep(a, b, options, true) ? b : a;
    std::uint64_t sum = 0;
    for (std::size_t z = 0; z != want.depth(); ++z)
        for (std::size_t y = 0; y != want.height(); ++y)
            for (std::size_t x = 0; x != want.width(); ++x) {
                T const v = want.elem_at(z, y, x);
                std::uint64_t bits = 0;
                std::memcpy(&bits, &v, sizeof(T));
                sum += mix64(bits ^ mix64((std::uint64_t(z) * want.height() + y) * want.width() + x));
            }
    return sum;
}

auto Stencil::run(InputMethod method) -> std::chrono::steady_clock::duration {
    if (method == CUDA && options.resident) return run_resident();
    initialize_matrix();
    return options.dtype == "f64" ? run_typed<double>(method) : run_typed<float>(method);
}

auto Stencil::run(std::string_view method_name) -> std::chrono::steady_clock::duration {
    InputMethod m;
    if (!parse_method(method_name, m)) throw std::runtime_error("unknown method name: " + std::string(method_name));
    return run(m);
}

double Stencil::updates() const {
    return double(options.extent(0)) * options.extent(1) * options.extent(2) * options.iterations;
}

template <class T>
bool Stencil::check_typed() const {
    auto const& p = std::get<Pair<T>>(grids);
    std::array<std::size_t, 3> const ext {options.extent(0), options.extent(1), options.extent(2)};
    auto a = BoundaryMatrix<T>::make(options.ndim(), ext, options.radius);
    auto b = BoundaryMatrix<T>::make(options.ndim(), ext, options.radius);
    fill_initial(a, options.init == "random", options.seed);
    fill_initial(b, false, 0);
    bool const swapped = naive_sweep(a, b, options, true);
    BoundaryMatrix<T> const& want = swapped ? b : a;
    BoundaryMatrix<T> const& got = swapped ? p.result : p.matrix;
    bit_exact_ = true;
    for (std::size_t z = 0; z != want.depth(); ++z)
        for (std::size_t row = 0; row != want.height(); ++row)
            for (std::size_t col = 0; col != want.width(); ++col) {
                T const w = want.elem_at(z, row, col), g = got.elem_at(z, row, col);
                if (std::memcmp(&w, &g, sizeof(T)) != 0) bit_exact_ = false;
                if (!(std::fabs(double(w) - double(g)) < 1e-4)) {
                    std::printf("invalid result at (%zu, %zu, %zu): %.15f vs %.15f\n", z, row, col, double(w), double(g));
                    return false;
                }
            }
    return true;
}

auto Stencil::check_result() const -> bool {
    if (info_.resident) {
        unsigned long long const want = options.dtype == "f64" ? naive_digest<double>() : naive_digest<float>();
        bit_exact_ = want == info_.checksum;
        if (!bit_exact_) std::printf("invalid result: digest %#llx vs %#llx of the naive loop\n", info_.checksum, want);
        return bit_exact_;
    }
    return options.dtype == "f64" ? check_typed<double>() : check_typed<float>();
}

auto Stencil::to_bmp() const -> BMPImage {
    // Blue -> cyan -> green -> yellow -> red heat map of the FINAL field, with the reference's
    // own colour arithmetic (stencil.cpp:153-188: the first ramp in the element type, the others
    // in double, truncating conversions) so that the file is byte-identical to the reference's
    // for the same field.  The reference reads `matrix` regardless of parity (stencil.cpp:159);
    // here the buffer that holds step T is read.  Values outside [0, 1] are clamped.
    unsigned const w = options.extent(0), h = options.extent(1);
    std::vector<BMPImage::Pixel> px;
    px.reserve(std::size_t(w) * h);
    bool const in_result = options.iterations & 1u;
    std::visit(
        [&](auto const& p) {
            using T = std::decay_t<decltype(p.matrix.elem_at(0, 0, 0))>;
            auto const& m = in_result ? p.result : p.matrix;
            for (unsigned row = 0; row < h; ++row)
                for (unsigned col = 0; col < w; ++col) {
                    T value = m.elem_at(0, row, col);
                    value = value < T(0) ? T(0) : value > T(1) ? T(1) : value;
                    BMPImage::Pixel rgb;
                    if (value < 0.25) {
                        rgb.r = 0;
                        rgb.g = static_cast<unsigned char>(4 * value * 255);
                        rgb.b = 255;
                    } else if (value < 0.5) {
                        rgb.r = 0;
                        rgb.g = 255;
                        rgb.b = static_cast<unsigned char>((1 + 4 * (0.25 - value)) * 255);
                    } else if (value < 0.75) {
                        rgb.r = static_cast<unsigned char>((4 * (value - 0.5)) * 255);
                        rgb.g = 255;
                        rgb.b = 0;
                    } else {
                        rgb.r = 255;
                        rgb.g = static_cast<unsigned char>((1 + 4 * (0.75 - value)) * 255);
                        rgb.b = 0;
                    }
                    px.push_back(rgb);
                }
        },
        grids);
    return BMPImage(std::move(px), w, h);
}
