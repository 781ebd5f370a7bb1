/*
 * domain_host.hpp — host-side (O(L_global), O(numRanks)) logic of the domain decomposition.  Plain C++, no CUDA.
 * Semantics follow the reference functions cited per function; the code is our own.
 */
#pragma once

#include <algorithm>
#include <cstdint>
#include <numeric>
#include <vector>

namespace csb
{

//! index of the first leaf whose start key is >= key  [ref: tree/csarray.hpp findNodeAbove]
template<class K>
inline int findNodeAbove(const K* leaves, int numLeafKeys, K key)
{
    return int(std::lower_bound(leaves, leaves + numLeafKeys, key) - leaves);
}

/*! bins of consecutive leaves holding (nearly) equal particle counts
 *  [ref: domain/domaindecomp.hpp:48-71 uniformBins]: bins[i] = lower_bound(scan, uint64(i * total / numBins)) */
inline void uniformBins(const uint32_t* counts, int numLeaves, int numBins, std::vector<int>& bins,
                        std::vector<uint32_t>& binCounts)
{
    std::vector<uint64_t> scan(size_t(numLeaves) + 1, 0);
    for (int i = 0; i < numLeaves; ++i)
        scan[i + 1] = scan[i] + counts[i];
    bins.assign(numBins + 1, 0);
    binCounts.assign(numBins, 0);
    const double binCount = double(scan.back()) / numBins;
    bins.front()          = 0;
    bins.back()           = numLeaves;
    for (int i = 1; i < numBins; ++i)
    {
        uint64_t target = uint64_t(i * binCount);
        bins[i]         = int(std::lower_bound(scan.begin(), scan.end(), target) - scan.begin());
    }
    for (int i = 1; i < numBins; ++i)
        binCounts[i - 1] = uint32_t(scan[bins[i]] - scan[bins[i - 1]]);
    binCounts.back() = uint32_t(scan.back() - scan[bins[numBins - 1]]);
}

/*! limit the shift of every rank boundary to the previous range of the neighbouring rank and recount if any moved
 *  [ref: domain/domaindecomp.hpp:139-166 limitBoundaryShifts]; boundaries and rankCounts are updated in place */
template<class K>
inline void limitBoundaryShifts(int numRanks, const K* oldBoundaries, K* boundaries, const K* leaves,
                                const uint32_t* counts, int numLeaves, uint32_t* rankCounts)
{
    bool recount = false;
    for (int r = 1; r < numRanks; ++r)
    {
        K nb = std::min(std::max(boundaries[r], oldBoundaries[r - 1]), oldBoundaries[r + 1]);
        if (nb != boundaries[r])
        {
            recount       = true;
            boundaries[r] = nb;
        }
    }
    if (!recount) return;
    for (int r = 0; r < numRanks; ++r)
    {
        int a    = findNodeAbove(leaves, numLeaves + 1, boundaries[r]);
        int b    = findNodeAbove(leaves, numLeaves + 1, boundaries[r + 1]);
        size_t c = 0;
        for (int i = a; i < b; ++i)
            c += counts[i];
        rankCounts[r] = uint32_t(c);
    }
}

/*! rank boundaries (SFC keys) of an equal-count split of the global leaves, shifts limited against the previous
 *  boundaries when given  [ref: domain/domaindecomp.hpp:115-124 makeSfcAssignment, domain/domain.hpp:288-290] */
template<class K>
inline void makeAssignment(int numRanks, const K* leaves, const uint32_t* counts, int numLeaves,
                           const K* oldBoundaries /* may be null */, std::vector<K>& boundaries,
                           std::vector<uint32_t>& rankCounts)
{
    std::vector<int> bins;
    uniformBins(counts, numLeaves, numRanks, bins, rankCounts);
    boundaries.resize(numRanks + 1);
    for (int r = 0; r <= numRanks; ++r)
        boundaries[r] = leaves[bins[r]];
    if (oldBoundaries)
        limitBoundaryShifts(numRanks, oldBoundaries, boundaries.data(), leaves, counts, numLeaves, rankCounts.data());
}

//! [ref: sfc/box.hpp:415-431 limitBoxShrinking] the box may lose at most 5 % of its previous extent per side
template<class T>
inline void limitBoxShrinking(const T* fit, const T* prev, T* out, T shrinkLimit = T(0.05))
{
    for (int d = 0; d < 3; ++d)
    {
        T len          = prev[2 * d + 1] - prev[2 * d];
        T lo           = prev[2 * d] + shrinkLimit * len;
        T hi           = prev[2 * d + 1] - shrinkLimit * len;
        out[2 * d]     = std::min(fit[2 * d], lo);
        out[2 * d + 1] = std::max(fit[2 * d + 1], hi);
    }
}

} // namespace csb
