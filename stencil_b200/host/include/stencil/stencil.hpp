// stencil.hpp — host-side stencil object (the "MPE" side of the reference).
//
// Same shape as the reference class (include/stencil/stencil.hpp:12-59): construct from
// ProgramOptions, run(method) returns the wall-clock duration of the accelerated sweep,
// check_result() recomputes with the naive loop and compares.  What changed is the
// accelerator behind run(): the four Sunway athread kernels are replaced by ONE call into
// the C ABI of include/stencil_b200.h (method "CUDA"), and the element type / rank of the
// grid are run-time options instead of a hard-coded BoundaryMatrix<float>.
#ifndef STENCIL_STENCIL_HPP
#define STENCIL_STENCIL_HPP

#include <chrono>
#include <string>
#include <string_view>
#include <variant>

#include <stencil/bmp_image.hpp>
#include <stencil/boundary_matrix.hpp>
#include <stencil/program_options.hpp>

class Stencil {
public:
    /// How the sweep is executed.
    enum InputMethod {
        /// B200 backend through the C ABI (sb200_run_host). No CPU fallback.
        CUDA,
        /// The naive reference-order loop on one host thread (what check_result runs).
        CPU,
        /// The same loop, OpenMP over rows/planes, all host cores.
        CPU_OMP,
        /// The reference's Sunway methods (stencil.hpp:15-32). Recognised, not available.
        DMA,
        RMA,
        DMA_SLAVE_PACK,
        DMA_STATIC_UNROLL,
    };

    struct RunInfo {
        double h2d_ms = 0, sweep_ms = 0, d2h_ms = 0;
        unsigned launches = 0, time_block = 0;
        int kernel = 0;
        int threads = 1;
        unsigned gpus = 1;       // GPUs the run was spread over
        bool resident = false;   // device-resident run (sb200_group_sweep): no host grids were moved
        unsigned long long checksum = 0;  // resident runs: digest of the final field
        bool pipelined = false;  // sweep_ms spans the whole overlapped upload/sweep/download pipeline
    };

    Stencil() = default;
    explicit Stencil(ProgramOptions options) : options(std::move(options)) {}

    void initialize_matrix();

    /// Runs T iterations with `method`; throws std::runtime_error if the method is not
    /// available (unknown name, Sunway method, or CUDA without a usable GPU).
    auto run(InputMethod method) -> std::chrono::steady_clock::duration;
    auto run(std::string_view method_name) -> std::chrono::steady_clock::duration;

    /// Naive recomputation + comparison (abs tolerance 1e-4 like the reference,
    /// stencil.cpp:137; additionally reports whether the match is bit-exact).
    auto check_result() const -> bool;
    bool last_check_bit_exact() const { return bit_exact_; }

    auto to_bmp() const -> BMPImage;
    RunInfo const& info() const { return info_; }
    double updates() const;  // interior points x iterations

    static bool parse_method(std::string_view name, InputMethod& out);

    /// Untimed bring-up for the CUDA method (the counterpart of the reference's athread_init(),
    /// src/main.cpp:54): device buffers, kernel attributes and one warm-up launch of every
    /// kernel variant a run with these options takes.  Returns false (message in
    /// sb200_last_error) if the backend is unusable; run() then reports the error.
    static bool prepare_cuda(ProgramOptions const& options);

private:
    template <class T> struct Pair {
        BoundaryMatrix<T> matrix;  // input of an iteration
        BoundaryMatrix<T> result;  // output of an iteration
    };
    ProgramOptions options;
    std::variant<Pair<float>, Pair<double>> grids;
    RunInfo info_;
    mutable bool bit_exact_ = false;

    template <class T> void initialize_typed();
    template <class T> auto run_typed(InputMethod method) -> std::chrono::steady_clock::duration;
    auto run_resident() -> std::chrono::steady_clock::duration;
    template <class T> unsigned long long naive_digest() const;
    template <class T> bool check_typed() const;
};

#endif  // STENCIL_STENCIL_HPP
