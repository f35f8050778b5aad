// boundary_matrix.hpp — padded row-major grids (1-D / 2-D / 3-D) for the stencil host.
//
// Same role and public vocabulary as the reference container
// (include/stencil/boundary_matrix.hpp:31-245: width()/height(), boundary_width(),
// elem_at(), elem_with_boundary_at(), fill_boundary(), borrow(), block_subview()), but
// written for up to three axes and for any element type, because the B200 backend sweeps
// 1-D/2-D/3-D grids in f32 and f64.  Memory layout is the contract with the C ABI
// (include/stencil_b200.h): x fastest, `boundary` cells on both sides of every used axis,
// dense (stride == padded width) for owners.
#ifndef STENCIL_BOUNDARY_MATRIX_HPP
#define STENCIL_BOUNDARY_MATRIX_HPP

#include <algorithm>
#include <array>
#include <cassert>
#include <cstddef>
#include <memory>

enum class BoundaryPosition { Left, Right, Top, Bottom, Front, Back };

namespace detail {

/// Shape bookkeeping shared by owners and views. Axis 0 = x (columns), 1 = y (rows), 2 = z.
struct PaddedShape {
    std::array<std::size_t, 3> interior {0, 1, 1};
    std::array<unsigned, 3> halo {0, 0, 0};
    std::size_t row_stride = 0;    // elements between consecutive rows
    std::size_t plane_stride = 0;  // elements between consecutive planes

    std::size_t padded(int axis) const { return interior[axis] + 2 * std::size_t(halo[axis]); }
    std::size_t padded_count() const { return padded(0) * padded(1) * padded(2); }
    std::size_t offset(std::size_t pz, std::size_t py, std::size_t px) const {
        return pz * plane_stride + py * row_stride + px;
    }
};

template <class ElemTy, bool Owner>
class BoundaryMatrix {
public:
    using value_type = ElemTy;

    BoundaryMatrix() = default;

    /// Owning 2-D grid (the reference's constructor signature): `width` x `height` interior
    /// with `boundary_width` / `boundary_height` halo cells per side, value-initialised.
    template <bool O = Owner, std::enable_if_t<O, int> = 0>
    BoundaryMatrix(std::size_t width, std::size_t height, unsigned boundary_width, unsigned boundary_height) {
        shape_.interior = {width, height, 1};
        shape_.halo = {boundary_width, boundary_height, 0};
        allocate();
    }

    /// Owning grid of `ndim` axes with a uniform halo of `radius` on every used axis.
    template <bool O = Owner, std::enable_if_t<O, int> = 0>
    static BoundaryMatrix make(int ndim, std::array<std::size_t, 3> extent, unsigned radius) {
        BoundaryMatrix m;
        for (int a = 0; a < 3; ++a) {
            m.shape_.interior[a] = a < ndim ? extent[a] : 1;
            m.shape_.halo[a] = a < ndim ? radius : 0;
        }
        m.allocate();
        return m;
    }

    /// Non-owning view over existing storage.
    template <bool O = Owner, std::enable_if_t<!O, int> = 0>
    BoundaryMatrix(PaddedShape shape, ElemTy* data) : shape_(shape), data_(data) {}

    // --- reference vocabulary ---------------------------------------------------------
    std::size_t width() const { return shape_.interior[0]; }
    std::size_t height() const { return shape_.interior[1]; }
    std::size_t depth() const { return shape_.interior[2]; }
    std::size_t width_with_boundary() const { return shape_.padded(0); }
    std::size_t height_with_boundary() const { return shape_.padded(1); }
    std::size_t depth_with_boundary() const { return shape_.padded(2); }
    unsigned boundary_width() const { return shape_.halo[0]; }
    unsigned boundary_height() const { return shape_.halo[1]; }
    unsigned boundary_depth() const { return shape_.halo[2]; }
    std::size_t internal_data_stride() const { return shape_.row_stride; }
    std::size_t elem_count() const { return width() * height() * depth(); }
    std::size_t elem_with_boundary_count() const { return shape_.padded_count(); }
    bool empty() const { return elem_count() == 0; }
    PaddedShape const& shape() const { return shape_; }

    ElemTy& elem_with_boundary_at(std::size_t row, std::size_t col) const {
        assert(row < height_with_boundary() && col < width_with_boundary());
        return raw_data()[shape_.offset(shape_.halo[2], row, col)];
    }
    ElemTy& elem_with_boundary_at(std::size_t plane, std::size_t row, std::size_t col) const {
        return raw_data()[shape_.offset(plane, row, col)];
    }
    ElemTy& elem_at(std::size_t row, std::size_t col) const {
        return elem_with_boundary_at(row + shape_.halo[1], col + shape_.halo[0]);
    }
    ElemTy& elem_at(std::size_t plane, std::size_t row, std::size_t col) const {
        return elem_with_boundary_at(plane + shape_.halo[2], row + shape_.halo[1], col + shape_.halo[0]);
    }

    /// Set one face of the halo (over the FULL padded extent of the other axes, corners
    /// included — the reference's semantics, boundary_matrix.hpp:129-170).
    void fill_boundary(BoundaryPosition position, ElemTy value) {
        std::array<std::size_t, 3> lo {0, 0, 0};
        std::array<std::size_t, 3> hi {shape_.padded(0), shape_.padded(1), shape_.padded(2)};
        int const axis = static_cast<int>(position) / 2;
        bool const high_side = static_cast<int>(position) % 2 == 1;
        if (high_side) {
            lo[axis] = hi[axis] - shape_.halo[axis];
        } else {
            hi[axis] = shape_.halo[axis];
        }
        for (std::size_t z = lo[2]; z < hi[2]; ++z)
            for (std::size_t y = lo[1]; y < hi[1]; ++y)
                std::fill(raw_data() + shape_.offset(z, y, lo[0]), raw_data() + shape_.offset(z, y, hi[0]), value);
    }

    BoundaryMatrix<ElemTy, false> borrow() const { return {shape_, raw_data()}; }

    /// The (row_idx, col_idx) tile of a `block_width` x `block_height` tiling, clipped at the
    /// grid edge; an empty view when the tile lies outside (boundary_matrix.hpp:190-218).
    [[nodiscard]] BoundaryMatrix<ElemTy, false>
    block_subview(std::size_t block_width, std::size_t block_height, unsigned row_idx, unsigned col_idx) const {
        PaddedShape s = shape_;
        std::size_t const r0 = row_idx * block_height, c0 = col_idx * block_width;
        if (r0 >= height() || c0 >= width()) {
            s.interior = {0, 0, depth()};
            return {s, raw_data()};
        }
        s.interior = {(std::min)(block_width, width() - c0), (std::min)(block_height, height() - r0), depth()};
        return {s, raw_data() + shape_.offset(0, r0, c0)};
    }

    ElemTy* raw_data() const {
        if constexpr (Owner) {
            return data_.get();
        } else {
            return data_;
        }
    }

private:
    PaddedShape shape_;
    std::conditional_t<Owner, std::unique_ptr<ElemTy[]>, ElemTy*> data_ {};

    void allocate() {
        shape_.row_stride = shape_.padded(0);
        shape_.plane_stride = shape_.padded(0) * shape_.padded(1);
        if constexpr (Owner) data_ = std::make_unique<ElemTy[]>(shape_.padded_count());
    }

    template <class E, bool O>
    friend class BoundaryMatrix;
};
}  // namespace detail

template <class ElemTy>
using BoundaryMatrix = detail::BoundaryMatrix<ElemTy, true>;
template <class ElemTy>
using BoundaryMatrixView = detail::BoundaryMatrix<ElemTy, false>;

#endif  // STENCIL_BOUNDARY_MATRIX_HPP
