// emu_driver.cpp — runs the reference's OWN accelerator kernels (src/stencil/slave/
// stencil_dma.cpp, stencil_dma_slave_pack.cpp, stencil_rma.cpp: DMA / DMASlavePack / RMA
// halo transport on the 8x8 CPE array) through the unmodified Stencil::run dispatch
// (src/stencil/stencil.cpp:23-73) on the athread emulator in emu/, and hands back the grid
// they produce together with the reference's own verdict on it (Stencil::check_result,
// abs 1e-4, stencil.cpp:137).  Built into oracle/_ref/libref_emu.so by oracle/Makefile.
// Test infrastructure only.
#include <cstring>
#include <string>

#define private public  // reach Stencil::matrix / result (stencil.hpp:46-51)
#include <stencil/stencil.hpp>
#undef private
#include "slave/stencil_slave.hpp"

// The slave translation units are compiled with -D__sw_slave__ and export the plain names
// (stencil_slave.hpp:9-11); the host side calls the slave_-prefixed ones.
extern "C" {
void stencil_iterate_dma(Arguments*);
void stencil_iterate_dma_slave_pack(Arguments*);
void stencil_iterate_rma(Arguments*);
void slave_stencil_iterate_dma(Arguments* a) { stencil_iterate_dma(a); }
void slave_stencil_iterate_dma_slave_pack(Arguments* a) { stencil_iterate_dma_slave_pack(a); }
void slave_stencil_iterate_rma(Arguments* a) { stencil_iterate_rma(a); }
// stencil_dma_static_unroll.cpp does not compile with g++ 13 (narrowing in a template
// argument, :726), so that method is not available under emulation.
void slave_stencil_iterate_dma_static_unroll(Arguments*) {}

// method: "DMA" | "DMASlavePack" | "RMA".  N must be 8 * block (every CPE owns one block and
// takes part in every barrier).  out[N*N] receives the interior after T iterations;
// *self_check = the reference's own check_result() verdict (1 = within 1e-4 of its naive loop).
// Returns 0, -1 for a bad request, -2 if the emulator could not run the kernel.
int ref_emu_run_2d_f32(const char* method, unsigned N, unsigned block, unsigned r, unsigned T, float* out,
                       int* self_check) {
    const std::string m(method ? method : "");
    if (m != "DMA" && m != "DMASlavePack" && m != "RMA") return -1;
    if (N != 8 * block || T == 0 || r == 0) return -1;
    if (m != "DMA" && r != 1) return -1;  // stencil_dma_slave_pack.cpp:380-382, stencil_rma.cpp:466-471
    if (block < 4 * r) return -1;         // stencil_dma.cpp:610-613 (unchecked in the reference)
    ProgramOptions opts{};
    opts.matrix_size = N;
    opts.block_size = block;
    opts.iterations = T;
    opts.radius = r;
    opts.repeat_count = 1;
    opts.check_result = true;
    Stencil s(opts);
    athread_init();
    s.run(std::string_view(m));
    if (sb200_emu_last_spawn_failed()) return -2;  // the emulator could not set up the CPE array
    auto& fin = (T & 1u) ? s.result : s.matrix;  // stencil.cpp:88-92,129-134
    for (unsigned row = 0; row != N; ++row)
        for (unsigned col = 0; col != N; ++col) out[(size_t)row * N + col] = fin.elem_at(row, col);
    if (self_check) *self_check = s.check_result() ? 1 : 0;
    return 0;
}
}
