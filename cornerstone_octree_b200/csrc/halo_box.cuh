/*
 * halo_box.cuh — integer box algebra of the halo search, shared by findHalos (halos.cu) and the domain engine
 * (domain.cu).  Follows traversal/boxoverlap.hpp of the reference (cited per function).
 */
#pragma once

#include "common.cuh"

namespace csb
{

// ------------------------------------------------------------------------------------------------
// integer box helpers
// ------------------------------------------------------------------------------------------------
__device__ inline bool overlapTwoRanges(int a, int b, int c, int d) { return b > c && d > a; }

//! periodic interval overlap, the three images the reference tests  [ref: traversal/boxoverlap.hpp:44-70]
template<int R>
__device__ inline bool overlapRange(int a, int b, int c, int d)
{
    return overlapTwoRanges(a, b, c, d) || overlapTwoRanges(a + R, b + R, c, d) || overlapTwoRanges(a, b, c + R, d + R);
}

template<class K>
__device__ inline bool overlapBoxes(const IBox& a, const IBox& b)
{
    constexpr int R = 1 << KeyTraits<K>::maxLevel;
    return overlapRange<R>(a.lo[0], a.hi[0], b.lo[0], b.hi[0]) && overlapRange<R>(a.lo[1], a.hi[1], b.lo[1], b.hi[1]) &&
           overlapRange<R>(a.lo[2], a.hi[2], b.lo[2], b.hi[2]);
}

//! [ref: traversal/boxoverlap.hpp:134-182 addDelta / makeHaloBox, sfc/common.hpp:80-90 toNBitIntCeil]
template<class K, class T>
__device__ inline IBox makeHaloBox(const IBox& node, float radius, const DBox<T>& box)
{
    constexpr int maxCoord = 1 << KeyTraits<K>::maxLevel;
    IBox halo;
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        T scaled       = T(radius) * box.ilen[d];
        unsigned cells = unsigned(ceil(scaled * T(unsigned(maxCoord))));
        int delta      = int(min(cells, unsigned(maxCoord - 1)));
        int lo         = node.lo[d] - delta;
        int hi         = node.hi[d] + delta;
        if (!box.pbc[d])
        {
            lo = min(max(0, lo), maxCoord);
            hi = min(max(0, hi), maxCoord);
        }
        halo.lo[d] = lo;
        halo.hi[d] = hi;
    }
    return halo;
}

//! is the box inside the key range [codeStart, codeEnd)?  [ref: traversal/boxoverlap.hpp:93-116, sfc/common.hpp:312-319]
template<class K>
__device__ inline bool containedIn(int kind, const uint16_t* enc, K codeStart, K codeEnd, const IBox& b)
{
    constexpr int R = 1 << KeyTraits<K>::maxLevel;
    if (min(min(b.lo[0], b.lo[1]), b.lo[2]) < 0 || max(max(b.hi[0], b.hi[1]), b.hi[2]) > R)
        return codeStart == 0 && codeEnd == nodeRange<K>(0);

    K lowCode   = sfcEncode<K>(kind, enc, uint32_t(b.lo[0]), uint32_t(b.lo[1]), uint32_t(b.lo[2]));
    K highCode  = sfcEncode<K>(kind, enc, uint32_t(b.hi[0] - 1), uint32_t(b.hi[1] - 1), uint32_t(b.hi[2] - 1));
    unsigned cl = unsigned(commonPrefix<K>(lowCode, highCode)) / 3u;
    K range     = nodeRange<K>(cl);
    K envStart  = lowCode & ~K(range - 1);
    K envEnd    = envStart + range;
    return envStart >= codeStart && envEnd <= codeEnd;
}

} // namespace csb
