#include <stencil/bmp_image.hpp>

#include <fstream>

namespace {
void put_le(std::ofstream& f, std::uint32_t v, int bytes) {
    for (int i = 0; i < bytes; ++i) f.put(static_cast<char>((v >> (8 * i)) & 0xff));
}
}  // namespace

bool BMPImage::save(std::string const& path) const {
    std::ofstream f(path, std::ios::binary);
    if (!f) return false;
    std::uint32_t const row_bytes = (width_ * 3 + 3) & ~3u;
    std::uint32_t const data_bytes = row_bytes * height_;
    f.put('B').put('M');
    put_le(f, 54 + data_bytes, 4);
    put_le(f, 0, 4);
    put_le(f, 54, 4);
    put_le(f, 40, 4);
    put_le(f, width_, 4);
    put_le(f, height_, 4);
    put_le(f, 1, 2);
    put_le(f, 24, 2);
    // the rest of the info header stays zero and rows go out in storage order (grid row 0
    // first, which BMP viewers show at the bottom): byte for byte the file the reference's
    // writer produces (src/stencil/bmp_image.cpp:7-64)
    (void)data_bytes;
    for (int i = 0; i < 6; ++i) put_le(f, 0, 4);
    for (unsigned y = 0; y < height_; ++y) {
        for (unsigned x = 0; x < width_; ++x) {
            Pixel const& p = pixels_[static_cast<std::size_t>(y) * width_ + x];
            f.put(static_cast<char>(p.b)).put(static_cast<char>(p.g)).put(static_cast<char>(p.r));
        }
        for (std::uint32_t pad = width_ * 3; pad < row_bytes; ++pad) f.put('\0');
    }
    return static_cast<bool>(f);
}
