/*
 * cstone_b200/find_neighbors.cuh — the device-side neighbour traversal for CLIENT kernels, header only.
 *
 * Same name, template parameters, argument list and output convention as the reference's
 *     cstone::traverseNeighbors(bodyBegin, bodyEnd, x, y, z, h, tree, box, warpNidx, ngmax, globalPool)
 * [ref: traversal/find_neighbors.cuh:435-506; launch pattern test/performance/neighbor_driver.cu:89-128]:
 * one WARP searches the neighbours of the up to 32 particles [bodyBegin, bodyEnd), lane l owns particle
 * min(bodyBegin + l, bodyEnd - 1); the k-th neighbour of lane l lands at warpNidx[k * TravConfig::targetSize + l]
 * (warp-interleaved, so that the client's loop over its neighbours reads coalesced), the full count is returned even
 * when it exceeds ngmax.  warpNidx is per-warp scratch of ngmax * 32 indices that the client reuses from group to
 * group: neighbours are consumed inside the client kernel and the particle-major list of findNeighbors (19 GB at
 * 32 M particles and ngmax 150) is never written.
 *
 * Semantics: the neighbour SET of a particle equals the one of cstone_b200_find_neighbors / the reference's CPU
 * findNeighbors (j != i, |x_j - x_i|^2 < 4 h_i^2 with the reference's distanceSq expression [ref: findneighbors.hpp:
 * 51-78], evaluated with round-to-nearest intrinsics so that a client build with -fmad=true cannot contract it); the
 * ORDER is this traversal's (leaves in depth-first order of the warp's walk), as in the reference, whose GPU and CPU
 * orders differ too [ref: test/unit_cuda/traversal/find_neighbors.cu sorts before comparing].
 *
 * Design: the walk is shared by the warp (8 lanes test the 8 children of a node against the box of the warp's search
 * cubes, scaled by tree.searchExtFactor; pending nodes wait in globalPool); the particles of a reached leaf are loaded
 * one per lane and handed round by shuffles; every lane tests every candidate for itself.
 */
#pragma once

#include <cuda_runtime.h>

#include "domain.hpp" // Box, OctreeNsView, LocalIndex, TreeNodeIndex

namespace cstone
{

struct TravConfig
{
    //! size of the per-warp traversal stack in globalPool, in node indices
    static constexpr unsigned memPerWarp = 512;
    //! particles per target group = lanes of a warp
    static constexpr unsigned targetSize = 32;
    static constexpr unsigned nwt        = 1;
    //! threads per block the pool helpers below assume
    static constexpr unsigned numThreads = 128;
    //! size of globalPool (TreeNodeIndex) for a launch of numBlocks blocks of numThreads threads
    static constexpr std::size_t poolSize(unsigned numBlocks)
    {
        return std::size_t(memPerWarp) * (numThreads / 32) * numBlocks;
    }
    //! size of the neighbour scratch (LocalIndex) for such a launch
    static constexpr std::size_t nidxSize(unsigned numBlocks, unsigned ngmax)
    {
        return std::size_t(ngmax) * numThreads * numBlocks;
    }
};

//! return type of traverseNeighbors: neighbour count(s) of the calling lane [ref: util::array<unsigned, TravConfig::nwt>]
struct NeighborCount
{
    unsigned v[TravConfig::nwt];
    __host__ __device__ unsigned operator[](int i) const { return v[i]; }
    __host__ __device__ unsigned& operator[](int i) { return v[i]; }
};

namespace detail_nb
{

template<class T>
struct BoxPod;
template<>
struct BoxPod<float>
{
    using type = cstone_box_f32;
};
template<>
struct BoxPod<double>
{
    using type = cstone_box_f64;
};

__device__ inline double mulRn(double a, double b) { return __dmul_rn(a, b); }
__device__ inline float mulRn(float a, float b) { return __fmul_rn(a, b); }
__device__ inline double addRn(double a, double b) { return __dadd_rn(a, b); }
__device__ inline float addRn(float a, float b) { return __fadd_rn(a, b); }

//! [ref: findneighbors.hpp:51-78 distanceSq]  (x1 - x2 per dimension, optional fold, dx*dx + dy*dy + dz*dz)
template<bool Pbc, class T, class B>
__device__ inline T distanceSq(T x1, T y1, T z1, T x2, T y2, T z2, const B& b)
{
    T dx = addRn(x1, -x2), dy = addRn(y1, -y2), dz = addRn(z1, -z2);
    if (Pbc)
    {
        dx = addRn(dx, -mulRn(mulRn(T(b.boundaries[0] == 1), b.lengths[0]), rint(mulRn(dx, b.inv_lengths[0]))));
        dy = addRn(dy, -mulRn(mulRn(T(b.boundaries[1] == 1), b.lengths[1]), rint(mulRn(dy, b.inv_lengths[1]))));
        dz = addRn(dz, -mulRn(mulRn(T(b.boundaries[2] == 1), b.lengths[2]), rint(mulRn(dz, b.inv_lengths[2]))));
    }
    return addRn(addRn(mulRn(dx, dx), mulRn(dy, dy)), mulRn(dz, dz));
}

//! does the node box (centre c, half size s) touch the box of the warp's search cubes (centre tc, half size ts)?
//! A superset test: a margin of a few ulp keeps every node the per-particle tests of findNeighbors would accept.
template<bool Pbc, class T, class B>
__device__ inline bool boxesTouch(const T* c, const T* s, const T* tc, const T* ts, const B& b)
{
    bool touch = true;
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        T dx = c[d] - tc[d];
        if (Pbc && b.boundaries[d] == 1) dx -= b.lengths[d] * rint(dx * b.inv_lengths[d]);
        const T reach = s[d] + ts[d];
        touch         = touch && (fabs(dx) <= reach + T(1e-6) * (reach + fabs(c[d]) + fabs(tc[d])));
    }
    return touch;
}

} // namespace detail_nb

template<class Tc, class Th, class KeyType>
__device__ NeighborCount traverseNeighbors(LocalIndex bodyBegin,
                                           LocalIndex bodyEnd,
                                           const Tc* __restrict__ x,
                                           const Tc* __restrict__ y,
                                           const Tc* __restrict__ z,
                                           const Th* __restrict__ h,
                                           const OctreeNsView<Tc, KeyType>& tree,
                                           const Box<Tc>& box,
                                           LocalIndex* warpNidx,
                                           unsigned ngmax,
                                           TreeNodeIndex* globalPool)
{
    using namespace detail_nb;
    const auto& b          = reinterpret_cast<const typename BoxPod<Tc>::type&>(box);
    const unsigned lane    = threadIdx.x & 31u;
    const unsigned full    = 0xffffffffu;
    const unsigned warpIdx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    TreeNodeIndex* stack   = globalPool + std::size_t(TravConfig::memPerWarp) * warpIdx;

    // this lane's particle [ref: find_neighbors.cuh:369-386 loadTarget]
    const LocalIndex i = min(bodyBegin + lane, bodyEnd - 1);
    const Tc xi = x[i], yi = y[i], zi = z[i];
    const Tc hi = Tc(h[i]);
    const Tc r2 = Tc(4.0) * hi * hi; // radiusSq as the CPU path forms it [ref: findneighbors.hpp:109]
    const Tc r  = Tc(2) * fabs(hi);

    // box of the warp's search cubes [ref: find_neighbors.cuh:389-413 warpBbox]
    Tc lo[3] = {xi - r, yi - r, zi - r}, up[3] = {xi + r, yi + r, zi + r};
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            lo[d] = fmin(lo[d], __shfl_xor_sync(full, lo[d], o));
            up[d] = fmax(up[d], __shfl_xor_sync(full, up[d], o));
        }
    }
    Tc tc[3], ts[3];
    bool inside = true;
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        tc[d]  = (up[d] + lo[d]) * Tc(0.5);
        ts[d]  = (up[d] - lo[d]) * Tc(0.5) * Tc(tree.searchExtFactor);
        inside = inside && (tc[d] - ts[d] >= b.limits[2 * d]) && (tc[d] + ts[d] <= b.limits[2 * d + 1]);
    }
    const bool anyPbc = b.boundaries[0] == 1 || b.boundaries[1] == 1 || b.boundaries[2] == 1;
    const bool usePbc = anyPbc && !inside; // [ref: find_neighbors.cuh:467-469]

    const Tc* centers = reinterpret_cast<const Tc*>(tree.centers);
    const Tc* sizes   = reinterpret_cast<const Tc*>(tree.sizes);
    unsigned nc       = 0;

    auto leaf = [&](TreeNodeIndex node)
    {
        const TreeNodeIndex l = tree.internalToLeaf[node];
        const LocalIndex jb = tree.layout[l], je = tree.layout[l + 1];
        for (LocalIndex base = jb; base < je; base += 32u)
        {
            const LocalIndex j = base + lane;
            Tc xj = 0, yj = 0, zj = 0;
            if (j < je) xj = x[j], yj = y[j], zj = z[j];
            const unsigned cnt = min(32u, je - base);
            for (unsigned k = 0; k < cnt; ++k)
            {
                const Tc xk = __shfl_sync(full, xj, k), yk = __shfl_sync(full, yj, k), zk = __shfl_sync(full, zj, k);
                const Tc d2 = usePbc ? distanceSq<true>(xk, yk, zk, xi, yi, zi, b) : distanceSq<false>(xk, yk, zk, xi, yi, zi, b);
                if (d2 < r2 && base + k != i)
                {
                    if (nc < ngmax) warpNidx[std::size_t(nc) * TravConfig::targetSize + lane] = base + k;
                    ++nc;
                }
            }
        }
    };

    // root
    bool rootTouch;
    {
        rootTouch = usePbc ? boxesTouch<true>(centers, sizes, tc, ts, b) : boxesTouch<false>(centers, sizes, tc, ts, b);
    }
    if (rootTouch)
    {
        if (tree.childOffsets[0] == 0) leaf(0);
        else
        {
            int sp = 0;
            if (lane == 0) stack[0] = tree.childOffsets[0];
            sp = 1;
            __syncwarp();
            while (sp > 0)
            {
                --sp;
                const TreeNodeIndex fc = stack[sp];
                __syncwarp();
                // lanes 0-7: one child each
                bool touch           = false;
                TreeNodeIndex childC = 0;
                if (lane < 8u)
                {
                    const TreeNodeIndex n = fc + TreeNodeIndex(lane);
                    touch  = usePbc ? boxesTouch<true>(centers + 3 * std::size_t(n), sizes + 3 * std::size_t(n), tc, ts, b)
                                    : boxesTouch<false>(centers + 3 * std::size_t(n), sizes + 3 * std::size_t(n), tc, ts, b);
                    childC = tree.childOffsets[n];
                }
                const unsigned leafMask = __ballot_sync(full, touch && childC == 0);
                const unsigned nodeMask = __ballot_sync(full, touch && childC != 0);
                // reached internal children wait on the stack
                if (touch && childC != 0)
                {
                    const int pos = sp + __popc(nodeMask & ((1u << lane) - 1u));
                    if (pos < int(TravConfig::memPerWarp)) stack[pos] = childC;
                }
                sp += __popc(nodeMask);
                if (sp > int(TravConfig::memPerWarp))
                {
                    // stack overflow (cannot happen for trees of depth <= 21: at most 7 * 21 + 1 entries are pending)
                    NeighborCount bad;
                    bad[0] = 0xffffffffu;
                    return bad;
                }
                __syncwarp();
                for (unsigned m = leafMask; m; m &= m - 1u)
                    leaf(fc + TreeNodeIndex(__ffs(int(m)) - 1));
            }
        }
    }
    NeighborCount out;
    out[0] = nc;
    return out;
}

} // namespace cstone
