// bmp_image.hpp — minimal 24-bit BMP writer for the --bmp heat-map dump (the reference
// ships an equivalent debugging aid, include/stencil/bmp_image.hpp:10-31, that nothing
// calls; here it is wired to a flag).
#ifndef STENCIL_BMP_IMAGE_HPP
#define STENCIL_BMP_IMAGE_HPP

#include <cstdint>
#include <string>
#include <vector>

class BMPImage {
public:
    struct Pixel {
        unsigned char b = 0, g = 0, r = 0;
    };
    BMPImage(std::vector<Pixel> pixels, unsigned width, unsigned height)
        : pixels_(std::move(pixels)), width_(width), height_(height) {}
    /// Rows are written in the order they are stored (the reference's on-disk format,
    /// src/stencil/bmp_image.cpp:19-25); returns false if the file cannot be written.
    bool save(std::string const& path) const;
    unsigned width() const { return width_; }
    unsigned height() const { return height_; }
    std::vector<Pixel> const& pixels() const { return pixels_; }

private:
    std::vector<Pixel> pixels_;
    unsigned width_, height_;
};

#endif  // STENCIL_BMP_IMAGE_HPP
