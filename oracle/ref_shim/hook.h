/* Force-included (-include) in front of the reference's stencil.cpp so that the
 * mismatch report `std::printf("invalid result at ...", row, col, naive, compared)`
 * (src/stencil/stencil.cpp:138-144) is captured instead of printed: the captured
 * `naive` double is the reference's own value, exact to the bit (float -> double
 * promotion is lossless).  All std headers the translation unit uses are included
 * first so the macro below cannot touch their declarations. */
#ifndef SB200_REF_SHIM_HOOK_H
#define SB200_REF_SHIM_HOOK_H
#include <algorithm>
#include <cassert>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <memory>
#include <optional>
#include <string>
#include <string_view>
#include <tuple>
#include <type_traits>
#include <unordered_map>
#include <vector>
namespace std {
int sb200_ref_capture(const char* fmt, ...);
}
#define printf sb200_ref_capture
/* SB200_REF_F64: the same reference sources at double precision.  Every std header the
 * reference includes is already in (guards above), so this one-token substitution only
 * touches the reference's own `float` (stencil.hpp:49-51, stencil.cpp:85-86,103): the loop
 * nest, the order of the adds and `sum * avg` are the reference's; `1.f / double(...)` and
 * `0.f` promote exactly. */
#ifdef SB200_REF_F64
#define float double
#endif
#endif
