/*
 * octree.cu — a10/a11: internal-tree linking, count upsweep, geometric node centres.
 *
 * Replaces buildOctreeGpu / upsweepSumGpu [ref: tree/octree_gpu.h:49-53, tree/octree_gpu.cu:55-208] and
 * computeGeoCentersGpu [ref: focus/source_center_gpu.cu:116-145]; semantics as the CPU versions
 * buildOctreeCpu [ref: tree/octree.hpp:185-214] and nodeFpCenters [ref: focus/source_center.hpp:145-157].
 */
#include "common.cuh"

extern "C" uint64_t cstone_b200_sort_by_key_temp_bytes_k32(uint64_t n);
extern "C" uint64_t cstone_b200_sort_by_key_temp_bytes_k64(uint64_t n);
extern "C" int cstone_b200_sort_by_key_k32(uint32_t*, uint32_t*, uint32_t*, uint32_t*, size_t, void*, uint64_t, int, void*);
extern "C" int cstone_b200_sort_by_key_k64(uint64_t*, uint32_t*, uint64_t*, uint32_t*, size_t, void*, uint64_t, int, void*);

namespace csb
{

inline uint64_t sortTemp(uint32_t, uint64_t n) { return cstone_b200_sort_by_key_temp_bytes_k32(n); }
inline uint64_t sortTemp(uint64_t, uint64_t n) { return cstone_b200_sort_by_key_temp_bytes_k64(n); }
inline int sortPairs(uint32_t* k, uint32_t* v, uint32_t* kb, uint32_t* vb, size_t n, void* t, uint64_t tb, void* s)
{
    return cstone_b200_sort_by_key_k32(k, v, kb, vb, n, t, tb, 1, s);
}
inline int sortPairs(uint64_t* k, uint32_t* v, uint64_t* kb, uint32_t* vb, size_t n, void* t, uint64_t tb, void* s)
{
    return cstone_b200_sort_by_key_k64(k, v, kb, vb, n, t, tb, 1, s);
}

inline size_t alignUp(size_t v, size_t a) { return (v + a - 1) / a * a; }

template<class K>
uint64_t buildOctreeTempBytes(int numLeaves)
{
    size_t M = size_t(numLeaves) + size_t(numLeaves - 1) / 7;
    return alignUp(M * sizeof(K), 256) + alignUp(M * 4, 256) + sortTemp(K(0), M) + 256;
}

//! [ref: sfc/common.hpp:288-292 digitWeight]
__host__ __device__ inline int digitWeight(int digit) { return digit >= 4 ? 7 - digit : -digit; }

/*! leaf prefixes go behind the internal part; leaf i also creates the internal node whose key range starts at
 *  leaf i whenever the common prefix of leaf i and i+1 ends on a digit boundary
 *  [ref: tree/octree.hpp:72-82 binaryKeyWeight, :95-118 createUnsortedLayoutCpu] */
template<class K>
__global__ void __launch_bounds__(256)
    unsortedLayoutKernel(const K* __restrict__ leaves, int numLeaves, int numInternal, K* __restrict__ prefixes)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numLeaves; i += gridDim.x * blockDim.x)
    {
        K key                     = leaves[i];
        K next                    = leaves[i + 1];
        unsigned level            = treeLevel<K>(K(next - key));
        prefixes[i + numInternal] = encodePlaceholderBit<K>(key, 3 * int(level));

        int cpl = commonPrefix<K>(key, next);
        if (cpl % 3 == 0 && i < numLeaves - 1)
        {
            int weight = 0;
            for (unsigned l = 1; l <= unsigned(cpl / 3) + 1u; ++l)
                weight += digitWeight(int(octalDigit<K>(key, l)));
            int slot       = (i + weight) / 7;
            prefixes[slot] = encodePlaceholderBit<K>(key, cpl);
        }
    }
}

//! both index maps from the sort permutation  [ref: tree/octree.hpp:200-209]
__global__ void __launch_bounds__(256) invertOrderKernel(int* __restrict__ internalToLeaf,
                                                          int* __restrict__ leafToInternal,
                                                          int* __restrict__ childOffsets,
                                                          int numNodes,
                                                          int numInternal)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numNodes; i += gridDim.x * blockDim.x)
    {
        int unsortedIdx              = internalToLeaf[i]; // value carried by the sort = position in the unsorted layout
        leafToInternal[unsortedIdx]  = i;
        internalToLeaf[i]            = unsortedIdx - numInternal;
        childOffsets[i]              = 0;
    }
}

template<class K>
__device__ inline int lowerBoundIdx(const K* __restrict__ a, int lo, int hi, K value)
{
    int len = hi - lo;
    while (len > 0)
    {
        int half   = len >> 1;
        bool right = a[lo + half] < value;
        lo         = right ? lo + half + 1 : lo;
        len        = right ? len - half - 1 : half;
    }
    return lo;
}

//! first node of every level  [ref: tree/octree.hpp:169-178 getLevelRangeCpu]
template<class K>
__global__ void levelRangeKernel(const K* __restrict__ prefixes, int numNodes, int* __restrict__ levelRange)
{
    unsigned level = threadIdx.x;
    if (level <= KeyTraits<K>::maxLevel)
        levelRange[level] = lowerBoundIdx<K>(prefixes, 0, numNodes, K(K(1) << (3 * level)));
    if (level == KeyTraits<K>::maxLevel + 1) levelRange[level] = numNodes;
}

//! child / parent links  [ref: tree/octree.hpp:132-166 linkTreeCpu]
template<class K>
__global__ void __launch_bounds__(256) linkKernel(const K* __restrict__ prefixes,
                                                   int numInternal,
                                                   const int* __restrict__ leafToInternal,
                                                   const int* __restrict__ levelRange,
                                                   int* __restrict__ childOffsets,
                                                   int* __restrict__ parents)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numInternal; i += gridDim.x * blockDim.x)
    {
        int node       = leafToInternal[i];
        K prefix       = prefixes[node];
        unsigned len   = decodePrefixLength<K>(prefix);
        unsigned level = len / 3;
        K childPrefix  = encodePlaceholderBit<K>(decodePlaceholderBit<K>(prefix), int(len) + 3);
        int lo         = levelRange[level + 1];
        int hi         = levelRange[level + 2];
        int child      = lowerBoundIdx<K>(prefixes, lo, hi, childPrefix);
        if (child != hi && prefixes[child] == childPrefix)
        {
            childOffsets[node]       = child;
            parents[(child - 1) / 8] = node;
        }
    }
}

template<class K>
int buildOctree(const K* leaves, int numLeaves, K* prefixes, int* childOffsets, int* parents, int* levelRange,
                int* internalToLeaf, int* leafToInternal, void* temp, uint64_t tempBytes, void* stream)
{
    if (numLeaves < 1 || !leaves || !prefixes || !childOffsets || !parents || !levelRange || !internalToLeaf ||
        !leafToInternal || !temp)
        return CSTONE_E_INVALID_ARGUMENT;
    if (tempBytes < buildOctreeTempBytes<K>(numLeaves)) return CSTONE_E_TEMP_TOO_SMALL;
    cudaStream_t s  = asStream(stream);
    int numInternal = (numLeaves - 1) / 7;
    int numNodes    = numLeaves + numInternal;

    char* p          = reinterpret_cast<char*>(alignUp(reinterpret_cast<uintptr_t>(temp), 256));
    K* keyBuf        = reinterpret_cast<K*>(p);
    p += alignUp(size_t(numNodes) * sizeof(K), 256);
    uint32_t* valBuf = reinterpret_cast<uint32_t*>(p);
    p += alignUp(size_t(numNodes) * 4, 256);
    uint64_t sortBytes = sortTemp(K(0), uint64_t(numNodes));

    unsortedLayoutKernel<K><<<gridFor(size_t(numLeaves), 256), 256, 0, s>>>(leaves, numLeaves, numInternal, prefixes);
    CSB_LAUNCHED();
    int st = sortPairs(prefixes, reinterpret_cast<uint32_t*>(internalToLeaf), keyBuf, valBuf, size_t(numNodes), p,
                       sortBytes, stream);
    if (st) return st;
    invertOrderKernel<<<gridFor(size_t(numNodes), 256), 256, 0, s>>>(internalToLeaf, leafToInternal, childOffsets,
                                                                       numNodes, numInternal);
    CSB_LAUNCHED();
    levelRangeKernel<K><<<1, 32, 0, s>>>(prefixes, numNodes, levelRange);
    CSB_LAUNCHED();
    if (numInternal > 0)
    {
        linkKernel<K><<<gridFor(size_t(numInternal), 256), 256, 0, s>>>(prefixes, numInternal, leafToInternal,
                                                                         levelRange, childOffsets, parents);
        CSB_LAUNCHED();
    }
    return 0;
}

//! saturating child sum for the nodes of one level  [ref: tree/octree.hpp:583-626, tree/octree_gpu.cu:178-208]
__global__ void __launch_bounds__(256)
    upsweepLevelKernel(int first, int last, const int* __restrict__ childOffsets, uint32_t* __restrict__ counts)
{
    for (int i = first + blockIdx.x * blockDim.x + threadIdx.x; i < last; i += gridDim.x * blockDim.x)
    {
        int c = childOffsets[i];
        if (c)
        {
            unsigned long long sum = 0;
#pragma unroll
            for (int o = 0; o < 8; ++o)
                sum += counts[c + o];
            counts[i] = uint32_t(sum > 0xFFFFFFFFull ? 0xFFFFFFFFull : sum);
        }
    }
}

//! [ref: sfc/box.hpp:334-351 centerAndSize]
template<class K, class T>
__global__ void __launch_bounds__(256) geoCentersKernel(int kind,
                                                         const uint16_t* __restrict__ decTab,
                                                         const K* __restrict__ prefixes,
                                                         int numNodes,
                                                         T* __restrict__ centers,
                                                         T* __restrict__ sizes,
                                                         DBox<T> box)
{
    __shared__ __align__(16) uint16_t tab[kHilbertTabSize];
    if (kind == CSTONE_SFC_HILBERT)
    {
        loadTable(tab, decTab);
        __syncthreads();
    }
    constexpr int maxCoord = 1 << KeyTraits<K>::maxLevel;
    const T uL             = T(1.) / T(maxCoord);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numNodes; i += gridDim.x * blockDim.x)
    {
        K prefix       = prefixes[i];
        unsigned level = decodePrefixLength<K>(prefix) / 3;
        IBox b         = sfcIBox<K>(kind, tab, decodePlaceholderBit<K>(prefix), level);
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            T halfUnit         = T(0.5) * uL * box.len[d];
            centers[3 * i + d] = box.lo[d] + T(b.hi[d] + b.lo[d]) * halfUnit;
            sizes[3 * i + d]   = T(b.hi[d] - b.lo[d]) * halfUnit;
        }
    }
}

template<class K, class T>
int geoCenters(int kind, const K* prefixes, int numNodes, T* centers, T* sizes, const typename BoxT<T>::type* box,
               void* stream)
{
    if (numNodes <= 0) return 0;
    if (!prefixes || !centers || !sizes || !box) return CSTONE_E_INVALID_ARGUMENT;
    SfcTables tabs;
    int st = getSfcTables(&tabs);
    if (st) return st;
    geoCentersKernel<K, T><<<gridFor(size_t(numNodes), 256, 4), 256, 0, asStream(stream)>>>(
        kind, tabs.dec, prefixes, numNodes, centers, sizes, makeDBox<T>(box));
    CSB_LAUNCHED();
    return 0;
}

} // namespace csb

extern "C" {

uint64_t cstone_b200_build_octree_temp_bytes_k32(int32_t n) { return csb::buildOctreeTempBytes<uint32_t>(n); }
uint64_t cstone_b200_build_octree_temp_bytes_k64(int32_t n) { return csb::buildOctreeTempBytes<uint64_t>(n); }

int cstone_b200_build_octree_k32(const uint32_t* leaves, int32_t numLeaves, uint32_t* prefixes, int32_t* childOffsets,
                                 int32_t* parents, int32_t* levelRange, int32_t* internalToLeaf,
                                 int32_t* leafToInternal, void* temp, uint64_t tempBytes, void* stream)
{
    return csb::buildOctree<uint32_t>(leaves, numLeaves, prefixes, childOffsets, parents, levelRange, internalToLeaf,
                                      leafToInternal, temp, tempBytes, stream);
}
int cstone_b200_build_octree_k64(const uint64_t* leaves, int32_t numLeaves, uint64_t* prefixes, int32_t* childOffsets,
                                 int32_t* parents, int32_t* levelRange, int32_t* internalToLeaf,
                                 int32_t* leafToInternal, void* temp, uint64_t tempBytes, void* stream)
{
    return csb::buildOctree<uint64_t>(leaves, numLeaves, prefixes, childOffsets, parents, levelRange, internalToLeaf,
                                      leafToInternal, temp, tempBytes, stream);
}

int cstone_b200_upsweep_sum(int32_t numLevels, const int32_t* levelRangeHost, const int32_t* childOffsets,
                            uint32_t* counts, void* stream)
{
    if (numLevels <= 0) return 0;
    if (!levelRangeHost || !childOffsets || !counts) return CSTONE_E_INVALID_ARGUMENT;
    for (int level = numLevels - 1; level >= 0; --level)
    {
        int first = levelRangeHost[level], last = levelRangeHost[level + 1];
        if (last <= first) continue;
        csb::upsweepLevelKernel<<<csb::gridFor(size_t(last - first), 256), 256, 0, csb::asStream(stream)>>>(
            first, last, childOffsets, counts);
        CSB_LAUNCHED();
    }
    return 0;
}

#define CSB_GEO_API(KS, K, TS, T)                                                                                      \
    int cstone_b200_compute_geo_centers_##KS##_##TS(int kind, const K* prefixes, int32_t numNodes, T* centers,         \
                                                    T* sizes, const cstone_box_##TS* box, void* stream)                \
    {                                                                                                                  \
        return csb::geoCenters<K, T>(kind, prefixes, numNodes, centers, sizes, box, stream);                           \
    }
CSB_GEO_API(k32, uint32_t, f32, float)
CSB_GEO_API(k32, uint32_t, f64, double)
CSB_GEO_API(k64, uint64_t, f32, float)
CSB_GEO_API(k64, uint64_t, f64, double)

} // extern "C"
