#include <stencil/cli.hpp>
#include <stencil/program_options.hpp>

#include <sstream>

int ProgramOptions::last_exit_code = 0;

namespace {
bool one_of(std::string const& v, std::initializer_list<char const*> allowed) {
    for (auto a : allowed)
        if (v == a) return true;
    return false;
}
}  // namespace

auto ProgramOptions::parse(int argc, char** argv) -> std::optional<ProgramOptions> {
    cli::Parser app("Iterative stencil sweep (B200 backend)");
    ProgramOptions result {};
    std::string shape_text;

    app.option("-s,--matrix-size", result.matrix_size, "The side length of the input matrix.").require();
    app.option("-i,--iteration", result.iterations, "The number of iterations.").require();
    app.option("-b,--block-size", result.block_size,
               "The side length of the block into which the matrix is divided (CUDA: a hint; with "
               "--block-as-tile the rows / planes each warp / CTA streams).")
        .require();
    app.option("-r,--radius", result.radius, "The radius of the stencil shape.", 1u);
    app.option("-R,--repeat", result.repeat_count, "The number of replicate runs for each method.", 1u);
    app.list("-m,--methods", result.method_names, "List of methods to be tested.").require();
    app.flag("-c,--check-result", result.check_result,
             "Whether to compare the result with the result of the naive implementation.");

    app.option("--shape", shape_text, "Grid extents nx[,ny[,nz]] (overrides -s; 1-D/2-D/3-D).");
    app.option("--dtype", result.dtype, "Element type: f32 or f64.", std::string("f32"));
    app.option("--stencil", result.stencil, "Stencil footprint: star or box.", std::string("star"));
    app.option("--time-block", result.time_block, "Fused time steps per HBM round trip (0 = auto).", 0u);
    app.option("--init", result.init, "Initial condition: ref or random.", std::string("ref"));
    app.option("--seed", result.seed, "Seed of the random initial condition.", 1ull);
    app.option("--kernel", result.kernel, "CUDA kernel family: auto, generic, warp2d, tma3d, resident2d, line1d.", std::string("auto"));
    app.option("--gpus", result.gpus, "GPUs of this box one CUDA run is spread over (z-slabs; y-slabs in 2-D).", 1u);
    app.flag("--resident", result.resident,
             "CUDA: keep the grid on the GPUs for the whole run (generated there, swept, digested; with --gpus N "
             "the slabs exchange halos directly over NVLink). The time is the sweep alone; -c compares digests.");
    app.flag("--reassociate", result.reassociate,
             "CUDA: allow kernels that sum a box point in another order than the naive loop (separable 9- / 27-point "
             "sums): faster, equal within 1e-5 (f32) / 1e-12 (f64) relative instead of bit for bit.");
    app.flag("--block-as-tile", result.block_as_tile,
             "CUDA: use -b as the tile preference of the warp2d / tma3d kernels instead of the planner's choice.");
    app.option("--json", result.json_path, "Append one JSON record per method to this file.");
    app.option("--bmp", result.bmp_path, "Write the final 2-D field as a BMP heat map.");

    try {
        app.parse(argc, argv);
        if (!shape_text.empty()) {
            std::stringstream ss(shape_text);
            std::string item;
            while (std::getline(ss, item, ',')) {
                std::size_t pos = 0;
                unsigned long v = 0;
                try {
                    v = std::stoul(item, &pos);
                } catch (std::exception const&) {
                    pos = 0;
                }
                if (pos != item.size() || v == 0) throw cli::ParseError("--shape: bad extent '" + item + "'", 104);
                result.shape.push_back(static_cast<unsigned>(v));
            }
            if (result.shape.empty() || result.shape.size() > 3)
                throw cli::ParseError("--shape takes 1 to 3 extents", 104);
        }
        if (!one_of(result.dtype, {"f32", "f64"})) throw cli::ParseError("--dtype must be f32 or f64", 105);
        if (!one_of(result.stencil, {"star", "box"})) throw cli::ParseError("--stencil must be star or box", 105);
        if (!one_of(result.init, {"ref", "random"})) throw cli::ParseError("--init must be ref or random", 105);
        if (!one_of(result.kernel, {"auto", "generic", "warp2d", "tma3d", "resident2d", "line1d"}))
            throw cli::ParseError("--kernel must be auto, generic, warp2d, tma3d, resident2d or line1d", 105);
        if (result.gpus == 0 || result.gpus > 16) throw cli::ParseError("--gpus must be in 1..16", 105);
        if (result.reassociate && result.resident && result.check_result)
            throw cli::ParseError("--reassociate with --resident -c: the resident check compares digests, which needs the naive order", 105);
        if (result.radius == 0) throw cli::ParseError("--radius must be at least 1", 105);
        if (result.repeat_count == 0) throw cli::ParseError("--repeat must be at least 1", 105);
        if (result.shape.empty() && result.matrix_size == 0) throw cli::ParseError("--matrix-size must be positive", 105);
    } catch (cli::ParseError const& e) {
        last_exit_code = app.exit(e);
        return std::nullopt;
    }
    return result;
}
