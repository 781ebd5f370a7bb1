/*
 * traverse_neighbors_client.cu — TEST CLIENT of include/cstone_b200/find_neighbors.cuh: a kernel of the shape of the
 * reference's neighbour driver [ref: test/performance/neighbor_driver.cu:89-128].  Warps draw target groups of 32
 * particles from a global counter, call cstone::traverseNeighbors and consume the neighbours IN PLACE from the per-warp
 * scratch (here: sum of h over the neighbours, a stand-in for a density loop); for the parity test the scratch is also
 * copied out particle-major.  Built by __graft_entry__.build() with nvcc (default flags: -fmad=true, to show that the
 * header does not depend on the client's contraction setting) into tests/cpp/_build/libtraverse_client.so.
 */
#include <cstdint>

#include "cstone_b200/find_neighbors.cuh"

using namespace cstone;

__device__ unsigned groupCounter;

template<class T, class K>
__global__ void __launch_bounds__(TravConfig::numThreads)
    clientKernel(const T* x, const T* y, const T* z, const T* h, LocalIndex first, LocalIndex last, OctreeNsView<T, K> tree,
                 Box<T> box, unsigned ngmax, LocalIndex* nidxPool, TreeNodeIndex* globalPool, unsigned* nc, LocalIndex* nbOut,
                 double* hsum)
{
    const unsigned lane    = threadIdx.x & 31u;
    const unsigned warpIdx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    LocalIndex* warpNidx   = nidxPool + std::size_t(warpIdx) * ngmax * TravConfig::targetSize;
    const unsigned groups  = (last - first + 31u) / 32u;
    for (;;)
    {
        unsigned g = 0;
        if (lane == 0) g = atomicAdd(&groupCounter, 1u);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= groups) return;
        const LocalIndex begin = first + g * 32u, end = min(begin + 32u, last);
        auto count             = traverseNeighbors(begin, end, x, y, z, h, tree, box, warpNidx, ngmax, globalPool);
        const LocalIndex i     = begin + lane;
        if (i < end)
        {
            const unsigned n = count[0], stored = min(n, ngmax);
            double s         = 0;
            for (unsigned k = 0; k < stored; ++k)
            {
                const LocalIndex j = warpNidx[std::size_t(k) * TravConfig::targetSize + lane]; // coalesced over the warp
                s += double(h[j]);
                if (nbOut) nbOut[std::size_t(i - first) * ngmax + k] = j;
            }
            nc[i - first]   = n;
            hsum[i - first] = s;
        }
        __syncwarp();
    }
}

__global__ void resetCounter() { groupCounter = 0; }

template<class T, class K>
static int run(const T* x, const T* y, const T* z, const T* h, unsigned first, unsigned last, int numLeafNodes, const K* prefixes,
               const int* childOffsets, const int* internalToLeaf, const int* levelRange, const K* leaves,
               const unsigned* layout, const T* centers, const T* sizes, const void* boxPod, unsigned ngmax, unsigned* nidxPool,
               int* globalPool, unsigned numBlocks, unsigned* nc, unsigned* nbOut, double* hsum)
{
    OctreeNsView<T, K> tree{numLeafNodes, prefixes,   childOffsets, internalToLeaf,
                            levelRange,   leaves,     layout,       reinterpret_cast<const Vec3<T>*>(centers),
                            reinterpret_cast<const Vec3<T>*>(sizes)};
    Box<T> box(*static_cast<const typename Box<T>::CType*>(boxPod));
    resetCounter<<<1, 1>>>();
    clientKernel<T, K><<<numBlocks, TravConfig::numThreads>>>(x, y, z, h, first, last, tree, box, ngmax, nidxPool, globalPool, nc,
                                                             nbOut, hsum);
    return int(cudaDeviceSynchronize());
}

extern "C" {
//! sizes of the two scratch pools (in 4-byte elements) for a launch of numBlocks blocks
void traverse_client_pool_sizes(unsigned numBlocks, unsigned ngmax, uint64_t* nidx, uint64_t* pool)
{
    *nidx = TravConfig::nidxSize(numBlocks, ngmax);
    *pool = TravConfig::poolSize(numBlocks);
}
int traverse_client_k64_f64(const double* x, const double* y, const double* z, const double* h, unsigned first, unsigned last,
                            int numLeafNodes, const uint64_t* prefixes, const int* childOffsets, const int* internalToLeaf,
                            const int* levelRange, const uint64_t* leaves, const unsigned* layout, const double* centers,
                            const double* sizes, const void* box, unsigned ngmax, unsigned* nidxPool, int* globalPool,
                            unsigned numBlocks, unsigned* nc, unsigned* nbOut, double* hsum)
{
    return run<double, uint64_t>(x, y, z, h, first, last, numLeafNodes, prefixes, childOffsets, internalToLeaf, levelRange, leaves,
                                 layout, centers, sizes, box, ngmax, nidxPool, globalPool, numBlocks, nc, nbOut, hsum);
}
int traverse_client_k32_f32(const float* x, const float* y, const float* z, const float* h, unsigned first, unsigned last,
                            int numLeafNodes, const uint32_t* prefixes, const int* childOffsets, const int* internalToLeaf,
                            const int* levelRange, const uint32_t* leaves, const unsigned* layout, const float* centers,
                            const float* sizes, const void* box, unsigned ngmax, unsigned* nidxPool, int* globalPool,
                            unsigned numBlocks, unsigned* nc, unsigned* nbOut, double* hsum)
{
    return run<float, uint32_t>(x, y, z, h, first, last, numLeafNodes, prefixes, childOffsets, internalToLeaf, levelRange, leaves,
                                layout, centers, sizes, box, ngmax, nidxPool, globalPool, numBlocks, nc, nbOut, hsum);
}
}
