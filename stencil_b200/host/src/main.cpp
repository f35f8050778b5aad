// main.cpp — stencil_main: same flow and the same two report lines as the reference CLI
// (src/main.cpp:12-62), which run_expr.py scrapes with a regex (run_expr.py:9).
#include <chrono>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string_view>

#include <stencil/program_options.hpp>
#include <stencil/stencil.hpp>
#include <stencil_b200.h>

namespace {
using ms_t = std::chrono::duration<double, std::milli>;

bool run_test(std::string_view method_name, ProgramOptions const& options) {
    if (options.check_result) {
        std::cout << "Start to check the correctness of method " << method_name << ".\n";
        Stencil stencil(options);
        stencil.run(method_name);
        if (stencil.check_result()) {
            std::cout << "The results of method " << method_name << " is correct.\n";
            std::cout << "  (bit-exact vs naive loop: " << (stencil.last_check_bit_exact() ? "yes" : "no") << ")\n";
        } else {
            std::cout << "The results of method " << method_name << " is incorrect.\n";
            return false;
        }
    }

    // one-time costs (device buffers, first launch of each kernel variant) stay out of the
    // timed runs, like athread_init() in the reference (src/main.cpp:54)
    if (method_name == "CUDA" && !options.resident) (void)Stencil::prepare_cuda(options);

    std::chrono::steady_clock::duration total_duration = {};
    double best_sweep_ms = 0;
    Stencil::RunInfo last {};
    double updates = 0;
    for (unsigned i = 0; i != options.repeat_count; ++i) {
        Stencil stencil(options);
        auto const duration = stencil.run(method_name);
        total_duration += duration;
        last = stencil.info();
        updates = stencil.updates();
        double const sweep = last.sweep_ms > 0 ? last.sweep_ms : ms_t(duration).count();
        if (i == 0 || sweep < best_sweep_ms) best_sweep_ms = sweep;
        std::cout << method_name << " Method spent " << ms_t(duration).count() << "ms for " << options.iterations
                  << " iterations.\n";
        if (i + 1 == options.repeat_count && !options.bmp_path.empty() && options.ndim() == 2)
            if (!stencil.to_bmp().save(options.bmp_path)) std::cerr << "cannot write " << options.bmp_path << "\n";
    }
    double const avg_ms = ms_t(total_duration / options.repeat_count).count();
    std::cout << "The average time taken by " << method_name << " method is " << avg_ms << "ms for "
              << options.iterations << " iterations.\n";

    // Extra lines (ignored by the legacy regex): throughput of the sweep itself.
    double const gstencil = best_sweep_ms > 0 ? updates / (best_sweep_ms * 1e-3) / 1e9 : 0;
    double const bytes = options.dtype == "f64" ? 16.0 : 8.0;
    std::cout << "  " << method_name << " sweep " << best_sweep_ms << " ms (best), " << gstencil << " GStencil/s, "
              << gstencil * bytes << " GB/s algorithmic";
    if (last.launches) std::cout << ", " << last.launches << " launches, time block " << last.time_block;
    if (method_name == "CUDA" && last.resident)
        std::cout << ", " << last.gpus << " GPU" << (last.gpus == 1 ? "" : "s")
                  << ", device-resident (no host copies), digest " << std::hex << std::showbase << last.checksum << std::dec;
    else if (method_name == "CUDA")
        std::cout << ", " << last.gpus << " GPU" << (last.gpus == 1 ? "" : "s")
                  << (last.pipelined ? ", sweep time = the whole overlapped upload/sweep/download pipeline"
                                     : ", sweep time = kernels only (copies: h2d " + std::to_string(last.h2d_ms) + " ms, d2h " +
                                           std::to_string(last.d2h_ms) + " ms)");
    if (last.threads > 1) std::cout << ", " << last.threads << " threads";
    std::cout << "\n";
    if (!options.json_path.empty()) {
        std::ofstream js(options.json_path, std::ios::app);
        js << "{\"method\": \"" << method_name << "\", \"avg_ms\": " << avg_ms << ", \"sweep_ms\": " << best_sweep_ms
           << ", \"gstencil_per_s\": " << gstencil << ", \"h2d_ms\": " << last.h2d_ms << ", \"d2h_ms\": " << last.d2h_ms
           << ", \"launches\": " << last.launches << ", \"time_block\": " << last.time_block
           << ", \"gpus\": " << last.gpus << ", \"pipelined\": " << (last.pipelined ? "true" : "false")
           << ", \"threads\": " << last.threads << ", \"iterations\": " << options.iterations << "}\n";
    }
    return true;
}

int run_all_test(ProgramOptions const& options) {
    int rc = 0;
    for (std::string const& method_name : options.method_names) {
        try {
            if (!run_test(method_name, options)) rc = 2;
        } catch (std::exception const& e) {
            std::cerr << "error: " << e.what() << "\n";
            rc = 3;
        }
    }
    return rc;
}
}  // namespace

int main(int argc, char** argv) {
    // Parse program arguments.  The reference brings its accelerator runtime up before any
    // timed run (athread_init(), main.cpp:54); the counterpart here is sb200_init(), which
    // creates the CUDA context, so that the first timed CUDA run does not pay for it.  A
    // failure is not fatal at this point: the CUDA method reports it when it runs.
    if (auto options = ProgramOptions::parse(argc, argv)) {
        for (std::string const& m : options->method_names)
            if (m == "CUDA") {
                (void)(options->gpus > 1 ? sb200_init_multi(int(options->gpus), nullptr) : sb200_init(0));
                break;
            }
        int const rc = run_all_test(*options);
        sb200_finalize();
        return rc;
    }
    return ProgramOptions::last_exit_code == 0 ? 0 : 1;
}
