// program_options.hpp — the CLI surface of stencil_main.
//
// The first seven fields and their flags are the reference's, unchanged
// (include/stencil/program_options.hpp:8-29, src/program_options.cpp:13-37):
//   -s/--matrix-size (req)  -i/--iteration (req)  -b/--block-size (req)  -r/--radius [1]
//   -R/--repeat [1]  -m/--methods (req, list)  -c/--check-result
// The rest are optional extensions whose defaults reproduce the reference problem
// (2-D, square, f32, star, reference initial condition).
#ifndef STENCIL_PROGRAM_OPTIONS_HPP
#define STENCIL_PROGRAM_OPTIONS_HPP

#include <optional>
#include <string>
#include <vector>

struct ProgramOptions {
    /// Side length of the matrix.
    unsigned matrix_size = 0;
    /// Reference: side of the tile each CPE owns (program_options.hpp:11-13).  CUDA backend:
    /// passed on as sb200_desc.block_hint, a tile / time-block preference.
    unsigned block_size = 0;
    /// Number of iterations.
    unsigned iterations = 0;
    /// The radius of the stencil shape.
    unsigned radius = 1;
    /// Number of replicate runs for each method.
    unsigned repeat_count = 1;
    /// Names of the methods to run (CUDA, CPU, CPUOMP; the Sunway names are recognised).
    std::vector<std::string> method_names;
    /// Compare against the naive implementation before timing.
    bool check_result = false;

    // ---- extensions -------------------------------------------------------------------
    /// Grid extents nx[,ny[,nz]]; empty = matrix_size x matrix_size.
    std::vector<unsigned> shape;
    std::string dtype = "f32";    // f32 | f64
    std::string stencil = "star"; // star | box
    unsigned time_block = 0;      // fused steps per HBM round trip; 0 = auto
    std::string init = "ref";     // ref | random
    unsigned long long seed = 1;
    std::string kernel = "auto";  // auto | generic | warp2d | tma3d | resident2d
    unsigned gpus = 1;            // GPUs one CUDA run is spread over (slabs of the outermost axis)
    bool resident = false;        // CUDA: device-resident run over the GPU group (no host grids; RMA-style)
    bool reassociate = false;     // allow separable box sums (SB200_FLAG_REASSOCIATE): -c then passes by tolerance
    bool block_as_tile = false;   // read -b as the CUDA tile preference (rows / planes per warp / CTA)
    std::string json_path;        // append one JSON record per method here
    std::string bmp_path;         // write the final 2-D field as a BMP heat map

    /// Parse the program arguments; nullopt (after printing the reason) on error/--help.
    static auto parse(int argc, char** argv) -> std::optional<ProgramOptions>;
    /// Exit code of the last failed parse (0 for --help).
    static int last_exit_code;

    int ndim() const { return shape.empty() ? 2 : static_cast<int>(shape.size()); }
    unsigned extent(int axis) const {
        if (shape.empty()) return axis < 2 ? matrix_size : 1;
        return axis < static_cast<int>(shape.size()) ? shape[axis] : 1;
    }
};

#endif  // STENCIL_PROGRAM_OPTIONS_HPP
