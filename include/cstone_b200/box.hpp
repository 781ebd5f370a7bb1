/*
 * cstone_b200/box.hpp — cstone::Box<T> and BoundaryType with the reference's public interface
 * [ref: sfc/box.hpp:45-50 BoundaryType, :112-200 Box<T>] and the memory layout of cstone_box_f32 / cstone_box_f64
 * (limits, lengths, inverse lengths computed once on the host as T(1)/(max-min), boundary types), so that a Box can
 * be handed to the C ABI without conversion.  Plain C++17, no CUDA headers.
 */
#pragma once

#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <type_traits>

#include "../cstone_b200.h"

namespace cstone
{

struct CpuTag
{
};
struct GpuTag
{
};

using LocalIndex    = unsigned;
using TreeNodeIndex = int;

enum class BoundaryType : char
{
    open     = 0,
    periodic = 1,
    fixed    = 2
};

template<class T>
class Box
{
    static_assert(std::is_same_v<T, float> || std::is_same_v<T, double>, "Box<T>: T must be float or double");

public:
    using CType = std::conditional_t<std::is_same_v<T, float>, cstone_box_f32, cstone_box_f64>;

    Box(T xyzMin, T xyzMax, BoundaryType b = BoundaryType::open)
        : Box(xyzMin, xyzMax, xyzMin, xyzMax, xyzMin, xyzMax, b, b, b)
    {
    }

    Box(T xmin,
        T xmax,
        T ymin,
        T ymax,
        T zmin,
        T zmax,
        BoundaryType bx = BoundaryType::open,
        BoundaryType by = BoundaryType::open,
        BoundaryType bz = BoundaryType::open)
    {
        const T lim[6] = {xmin, xmax, ymin, ymax, zmin, zmax};
        for (int i = 0; i < 6; ++i)
            c_.limits[i] = lim[i];
        for (int d = 0; d < 3; ++d)
        {
            c_.lengths[d]     = lim[2 * d + 1] - lim[2 * d];
            c_.inv_lengths[d] = T(1) / c_.lengths[d];
        }
        c_.boundaries[0] = char(bx), c_.boundaries[1] = char(by), c_.boundaries[2] = char(bz);
    }

    explicit Box(const CType& c)
        : c_(c)
    {
    }

    T xmin() const { return c_.limits[0]; }
    T xmax() const { return c_.limits[1]; }
    T ymin() const { return c_.limits[2]; }
    T ymax() const { return c_.limits[3]; }
    T zmin() const { return c_.limits[4]; }
    T zmax() const { return c_.limits[5]; }

    T lx() const { return c_.lengths[0]; }
    T ly() const { return c_.lengths[1]; }
    T lz() const { return c_.lengths[2]; }
    T ilx() const { return c_.inv_lengths[0]; }
    T ily() const { return c_.inv_lengths[1]; }
    T ilz() const { return c_.inv_lengths[2]; }

    BoundaryType boundaryX() const { return BoundaryType(c_.boundaries[0]); }
    BoundaryType boundaryY() const { return BoundaryType(c_.boundaries[1]); }
    BoundaryType boundaryZ() const { return BoundaryType(c_.boundaries[2]); }

    T minExtent() const { return lx() < ly() ? (lx() < lz() ? lx() : lz()) : (ly() < lz() ? ly() : lz()); }
    T maxExtent() const { return lx() > ly() ? (lx() > lz() ? lx() : lz()) : (ly() > lz() ? ly() : lz()); }

    const CType* c() const { return &c_; }

    friend bool operator==(const Box& a, const Box& b)
    {
        for (int i = 0; i < 6; ++i)
            if (a.c_.limits[i] != b.c_.limits[i]) return false;
        for (int d = 0; d < 3; ++d)
            if (a.c_.boundaries[d] != b.c_.boundaries[d]) return false;
        return true;
    }

private:
    CType c_;
};

namespace detail
{
/*! the reference's error behaviour: CUDA errors print and terminate [ref: cuda/errorcheck.cuh:30-40], logic errors
 *  throw std::runtime_error [ref: domain/domain.hpp:108-112, :448-453] */
inline void check(long long status, const char* what)
{
    if (status > 0)
    {
        std::fprintf(stderr, "%s: %s\n", what, cstone_b200_status_string(int(status)));
        std::exit(EXIT_FAILURE);
    }
    if (status < 0) throw std::runtime_error(std::string(what) + ": " + cstone_b200_status_string(int(status)));
}
} // namespace detail

} // namespace cstone
