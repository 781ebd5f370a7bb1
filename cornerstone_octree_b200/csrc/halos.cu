/*
 * halos.cu — a12-a14: halo radii per leaf, halo discovery by integer box collision, node layout.
 *
 * Replaces segmentMax + scaleGpu as used by Halos::discover [ref: halos/halos.hpp:148-160,
 * primitives/primitives_gpu.cu:240-259], findHalosGpu [ref: traversal/collisions_gpu.h:58-67,
 * traversal/collisions_gpu.cu:39-104] and computeNodeLayout [ref: domain/layout.hpp:150-164].
 * Box arithmetic follows traversal/boxoverlap.hpp (cited per function).
 */
#include "halo_box.cuh"
#include "scan.cuh"

namespace csb
{

// ------------------------------------------------------------------------------------------------
// a12: radii[i] = float(max(0, max h[seg_i])) * scale
// One warp per 32 segments: short segments (<= bucket size) are scanned by their own lane.
// ------------------------------------------------------------------------------------------------
template<class Th>
__global__ void __launch_bounds__(256) haloRadiiKernel(const Th* __restrict__ h,
                                                        const uint32_t* __restrict__ seg,
                                                        size_t numSegments,
                                                        float scale,
                                                        float* __restrict__ radii)
{
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < numSegments; i += stride)
    {
        uint32_t a = seg[i], b = seg[i + 1];
        Th m       = 0;
        for (uint32_t j = a; j < b; ++j)
            m = max(m, h[j]);
        radii[i] = __fmul_rn(float(m), scale);
    }
}

// ------------------------------------------------------------------------------------------------
// a13 findHalos.  One thread per local leaf, depth-first walk in the reference's visiting order
// [ref: traversal/traversal.hpp:68-110 singleTraversal, traversal/collisions.hpp:39-105].
// The stack cannot overflow: at most 7 siblings stay pushed per level (7*21 + 1 < kHaloStack).
// ------------------------------------------------------------------------------------------------
constexpr int kHaloStack   = 160;
constexpr int kHaloThreads = 128;

template<class K, class T>
__global__ void __launch_bounds__(kHaloThreads) findHalosKernel(int kind,
                                                                 const uint16_t* __restrict__ enc,
                                                                 const uint16_t* __restrict__ dec,
                                                                 const K* __restrict__ prefixes,
                                                                 const int* __restrict__ childOffsets,
                                                                 const int* __restrict__ internalToLeaf,
                                                                 const K* __restrict__ leaves,
                                                                 const float* __restrict__ radii,
                                                                 DBox<T> box,
                                                                 int firstNode,
                                                                 int lastNode,
                                                                 int* __restrict__ flags)
{
    const K lowest  = leaves[firstNode];
    const K highest = leaves[lastNode];

    for (int leaf = firstNode + blockIdx.x * blockDim.x + threadIdx.x; leaf < lastNode;
         leaf += gridDim.x * blockDim.x)
    {
        K start       = leaves[leaf];
        K end         = leaves[leaf + 1];
        IBox nodeBox  = sfcIBox<K>(kind, dec, start, treeLevel<K>(K(end - start)));
        IBox halo     = makeHaloBox<K, T>(nodeBox, radii[leaf], box);
        if (containedIn<K>(kind, enc, lowest, highest, halo)) continue;

        auto descend = [&](int idx) -> bool
        {
            K prefix       = prefixes[idx];
            unsigned level = decodePrefixLength<K>(prefix) / 3u;
            K nodeStart    = decodePlaceholderBit<K>(prefix);
            K nodeEnd      = nodeStart + nodeRange<K>(level);
            bool inOwn     = !(nodeStart < lowest || nodeEnd > highest);
            if (inOwn) return false;
            return overlapBoxes<K>(sfcIBox<K>(kind, dec, nodeStart, level), halo);
        };

        if (!descend(0)) continue;
        if (childOffsets[0] == 0)
        {
            flags[internalToLeaf[0]] = 1;
            continue;
        }
        int stack[kHaloStack];
        int sp      = 0;
        stack[sp++] = 0;
        int node    = 0;
        do
        {
            int firstChild = childOffsets[node];
            for (int o = 0; o < 8; ++o)
            {
                int child = firstChild + o;
                if (!descend(child)) continue;
                if (childOffsets[child] == 0) flags[internalToLeaf[child]] = 1;
                else stack[sp++] = child;
            }
            node = stack[--sp];
        } while (node != 0);
    }
}

template<class K, class T>
int findHalos(int kind, const K* prefixes, const int* childOffsets, const int* internalToLeaf, const K* leaves,
              const float* radii, const typename BoxT<T>::type* box, int firstNode, int lastNode, int* flags,
              void* stream)
{
    if (kind != CSTONE_SFC_MORTON && kind != CSTONE_SFC_HILBERT) return CSTONE_E_INVALID_ARGUMENT;
    if (lastNode <= firstNode) return 0;
    if (!prefixes || !childOffsets || !internalToLeaf || !leaves || !radii || !box || !flags || firstNode < 0)
        return CSTONE_E_INVALID_ARGUMENT;
    SfcTables tabs;
    int st = getSfcTables(&tabs);
    if (st) return st;
    // one thread per leaf, uncapped: most leaves end at the containedIn test, the hardware scheduler balances the rest
    unsigned grid = unsigned((size_t(lastNode - firstNode) + kHaloThreads - 1) / kHaloThreads);
    findHalosKernel<K, T><<<grid, kHaloThreads, 0, asStream(stream)>>>(kind, tabs.enc, tabs.dec, prefixes,
                                                                         childOffsets, internalToLeaf, leaves, radii,
                                                                         makeDBox<T>(box), firstNode, lastNode, flags);
    CSB_LAUNCHED();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// a14 layout: exclusive scan of the counts of present leaves, masking fused into the scan loader
// ------------------------------------------------------------------------------------------------
struct LoadPresentCount
{
    const uint32_t* counts;
    const int* flags;
    int first, last, numLeaves;
    __device__ uint32_t operator()(size_t i) const
    {
        int k = int(i);
        if (k >= numLeaves) return 0u;
        bool present = (first <= k && k < last) || flags[k];
        return present ? counts[k] : 0u;
    }
};

} // namespace csb

extern "C" {

int cstone_b200_halo_radii_f32(const float* h, const uint32_t* segments, size_t numSegments, float scale,
                               float* radii, void* stream)
{
    if (numSegments == 0) return 0;
    if (!h || !segments || !radii) return CSTONE_E_INVALID_ARGUMENT;
    csb::haloRadiiKernel<float><<<csb::gridFor(numSegments, 256), 256, 0, csb::asStream(stream)>>>(
        h, segments, numSegments, scale, radii);
    CSB_LAUNCHED();
    return 0;
}
int cstone_b200_halo_radii_f64(const double* h, const uint32_t* segments, size_t numSegments, float scale,
                               float* radii, void* stream)
{
    if (numSegments == 0) return 0;
    if (!h || !segments || !radii) return CSTONE_E_INVALID_ARGUMENT;
    csb::haloRadiiKernel<double><<<csb::gridFor(numSegments, 256), 256, 0, csb::asStream(stream)>>>(
        h, segments, numSegments, scale, radii);
    CSB_LAUNCHED();
    return 0;
}

#define CSB_HALO_API(KS, K, TS, T)                                                                                     \
    int cstone_b200_find_halos_##KS##_##TS(int kind, const K* prefixes, const int32_t* childOffsets,                   \
                                           const int32_t* internalToLeaf, const K* leaves, const float* radii,         \
                                           const cstone_box_##TS* box, int32_t firstNode, int32_t lastNode,            \
                                           int32_t* flags, void* stream)                                               \
    {                                                                                                                  \
        return csb::findHalos<K, T>(kind, prefixes, childOffsets, internalToLeaf, leaves, radii, box, firstNode,       \
                                    lastNode, flags, stream);                                                          \
    }
CSB_HALO_API(k32, uint32_t, f32, float)
CSB_HALO_API(k32, uint32_t, f64, double)
CSB_HALO_API(k64, uint64_t, f32, float)
CSB_HALO_API(k64, uint64_t, f64, double)

int cstone_b200_compute_node_layout(const uint32_t* counts, const int32_t* flags, int32_t firstAssigned,
                                    int32_t lastAssigned, int32_t numLeaves, uint32_t* layout, void* temp,
                                    void* stream)
{
    const bool allAssigned = firstAssigned <= 0 && lastAssigned >= numLeaves; // flags are not read then
    if (numLeaves < 0 || !layout || !temp || (numLeaves && (!counts || (!flags && !allAssigned))))
        return CSTONE_E_INVALID_ARGUMENT;
    csb::LoadPresentCount load{counts, flags, firstAssigned, lastAssigned, numLeaves};
    return csb::launchScan<false>(load, layout, size_t(numLeaves) + 1, 0u, temp, nullptr, csb::asStream(stream));
}

} // extern "C"
