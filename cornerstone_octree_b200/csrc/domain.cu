/*
 * domain.cu — the engine behind cstone::Domain<KeyType,T,GpuTag>: one rank's part of the SFC-sharded domain.
 *
 * Replaces the orchestration of Domain::sync [ref: domain/domain.hpp:196-243, :492-604], GlobalAssignmentGpu
 * [ref: domain/assignment_gpu.cuh:78-175], exchangeParticlesGpu [ref: domain/domaindecomp_mpi_gpu.cuh:85-185] and
 * Halos / haloExchangeGpu [ref: halos/halos.hpp:127-251, halos/exchange_halos_gpu.cuh:34-119].
 *
 * B200 design
 *  - all particle data stay in HBM; the engine owns SFC-ordered copies and hands them back in sync_finish;
 *  - the replicated global tree is updated on every GPU redundantly: local leaf counts -> ncclAllReduce(SUM) ->
 *    max(local, sum) guard [ref: tree/update_mpi.hpp:47-69] -> same deterministic rebalance everywhere;
 *  - particle exchange: the local particles are sorted by key, ranks own contiguous key ranges, so ONE gather puts
 *    every outgoing message in place; messages go out as one NCCL group of send/recv per array over NVLink and are
 *    received straight into their final slots (no staging, no unpack kernel);
 *  - halo discovery: every rank publishes the integer halo boxes of its surface leaves (ncclAllGather), every owner
 *    tests the foreign boxes against ITS OWN leaves (reverse collision, same integer overlap test as findHalos) and
 *    sends the particles of the leaves that are hit; received halos land left / right of the assigned range in rank
 *    order, which is SFC order.  The send lists are replayed by exchange_halos for arbitrary arrays.
 *  - the search tree handed to findNeighbors is the cornerstone leaf array over assigned + halo keys
 *    (bucketSizeFocus), linked and with node centres.
 */
#include <nccl.h>

#include <chrono>
#include <cstdio>
#include <string>
#include <memory>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/cstone_b200_domain.h"
#include "comm.hpp"
#include "let_tree.cuh"
#include "halo_box.cuh"
#include "scan.cuh"
#include "domain_host.hpp"

namespace csb
{

#define CSB_NCCL(expr)                                                                                                 \
    do                                                                                                                 \
    {                                                                                                                  \
        ncclResult_t r_ = (expr);                                                                                      \
        if (r_ != ncclSuccess) return CSTONE_E_NCCL;                                                                   \
    } while (0)


//! grow-only device buffer (capacity grows by 5 %, as the reference's reallocate policy, util/reallocate.hpp:37-66)
struct DevBuf
{
    void* p      = nullptr;
    size_t bytes = 0;
    int reserve(size_t need, bool keep = false, cudaStream_t s = 0)
    {
        if (need <= bytes) return 0;
        size_t cap = size_t(double(need) * 1.05) + 256;
        void* q    = nullptr;
        CSB_CUDA(cudaMalloc(&q, cap));
        if (keep && p && bytes)
        {
            CSB_CUDA(cudaMemcpyAsync(q, p, bytes, cudaMemcpyDeviceToDevice, s));
            CSB_CUDA(cudaStreamSynchronize(s));
        }
        if (p) cudaFree(p);
        p     = q;
        bytes = cap;
        return 0;
    }
    template<class U>
    U* as() const
    {
        return reinterpret_cast<U*>(p);
    }
    ~DevBuf()
    {
        if (p) cudaFree(p);
    }
    DevBuf()                         = default;
    DevBuf(const DevBuf&)            = delete;
    DevBuf& operator=(const DevBuf&) = delete;
};

// ------------------------------------------------------------------------------------------------
// small kernels of the engine
// ------------------------------------------------------------------------------------------------
//! out[q] = index of the first key >= queries[q]   [ref: primitives/primitives_gpu.h lowerBoundGpu]
template<class K>
__global__ void lowerBoundQueriesKernel(const K* __restrict__ keys, uint32_t n, const K* __restrict__ queries, int nq,
                                        uint32_t* __restrict__ out)
{
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    K v         = queries[q];
    uint32_t lo = 0, len = n;
    while (len > 0)
    {
        uint32_t half = len >> 1;
        bool right    = keys[lo + half] < v;
        lo            = right ? lo + half + 1 : lo;
        len           = right ? len - half - 1 : half;
    }
    out[q] = lo;
}

/* Merge of the particles that stayed on this rank (sorted run `own`) with the sorted foreign arrivals `f`:
 * ip[u] = number of own keys < f[u] (foreign first on ties), cnt[ip[u]]++.  With E = exclusive scan of cnt the merged
 * position of own element t is t + E[t+1], of foreign element u it is ip[u] + u. */
template<class K>
__global__ void insertionPointsKernel(const K* __restrict__ own, uint32_t nOwn, const K* __restrict__ f, uint32_t nF,
                                      uint32_t* __restrict__ ip, uint32_t* __restrict__ cnt)
{
    for (uint32_t u = blockIdx.x * blockDim.x + threadIdx.x; u < nF; u += gridDim.x * blockDim.x)
    {
        K v         = f[u];
        uint32_t lo = 0, len = nOwn;
        while (len > 0)
        {
            uint32_t half = len >> 1;
            bool right    = own[lo + half] < v;
            lo            = right ? lo + half + 1 : lo;
            len           = right ? len - half - 1 : half;
        }
        ip[u] = lo;
        atomicAdd(cnt + lo, 1u);
    }
}

/* Two-source map of the steady-state merge: bit 31 clear = index into the caller's arrays (a particle that stayed on
 * this rank, reached through the map of the first sort), bit 31 set = index into the receive buffers (an arrival). */
constexpr uint32_t kFromRecv = 0x80000000u;

template<class K>
__global__ void mergeScatterKernel(const K* __restrict__ own, uint32_t nOwn, const uint32_t* __restrict__ ownMap,
                                   const uint32_t* __restrict__ scanE, const K* __restrict__ f,
                                   const uint32_t* __restrict__ fv, const uint32_t* __restrict__ ip, uint32_t nF,
                                   K* __restrict__ outKeys, uint32_t* __restrict__ outOrder)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nOwn + nF; i += gridDim.x * blockDim.x)
    {
        if (i < nOwn)
        {
            uint32_t pos  = i + scanE[i + 1];
            outKeys[pos]  = own[i];
            outOrder[pos] = ownMap[i];
        }
        else
        {
            uint32_t u    = i - nOwn;
            uint32_t pos  = ip[u] + u;
            outKeys[pos]  = f[u];
            outOrder[pos] = fv[u] | kFromRecv;
        }
    }
}

struct GatherTwoArgs
{
    const void* orig[16];
    const void* recv[16];
    void* dst[16];
    int words[16]; // element size in 4-byte words
    int na;
};
//! dst[k][i] = (map[i] & kFromRecv ? recv[k] : orig[k])[map[i] & ~kFromRecv] for every array k, one read of the map
__global__ void __launch_bounds__(256) gatherTwoKernel(const uint32_t* __restrict__ map, size_t n, GatherTwoArgs a)
{
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        const uint32_t m = map[i];
        const size_t j   = m & ~kFromRecv;
        for (int k = 0; k < a.na; ++k)
        {
            const void* src = (m & kFromRecv) ? a.recv[k] : a.orig[k];
            if (a.words[k] == 2)
                reinterpret_cast<uint64_t*>(a.dst[k])[i] = reinterpret_cast<const uint64_t*>(src)[j];
            else
                for (int w = 0; w < a.words[k]; ++w)
                    reinterpret_cast<uint32_t*>(a.dst[k])[i * a.words[k] + w] =
                        reinterpret_cast<const uint32_t*>(src)[j * a.words[k] + w];
        }
    }
}

//! v[u] = u for u < nLeft, u + gap beyond: indices of the foreign arrivals in the receive buffers
__global__ void foreignIndexKernel(uint32_t* v, uint32_t nF, uint32_t nLeft, uint32_t gap)
{
    for (uint32_t u = blockIdx.x * blockDim.x + threadIdx.x; u < nF; u += gridDim.x * blockDim.x)
        v[u] = u < nLeft ? u : u + gap;
}

//! counts[i] = max(local[i], reduced[i])   [ref: tree/update_mpi.hpp:62-65]
__global__ void maxMergeKernel(const uint32_t* __restrict__ local, uint32_t* __restrict__ reduced, int n)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        reduced[i] = max(local[i], reduced[i]);
}

struct IBox6
{
    int v[6]; // xmin,xmax,ymin,ymax,zmin,zmax (integer cell coordinates, may exceed [0,2^maxLevel] for PBC)
};


// ------------------------------------------------------------------------------------------------
// type dispatch onto the C ABI of the kernels (same shared library)
// ------------------------------------------------------------------------------------------------
template<class K, class T>
struct Api
{
    static constexpr bool k64 = sizeof(K) == 8, f64 = sizeof(T) == 8;
    using CBox = typename BoxT<T>::type;

    static int keys(int kind, const T* x, const T* y, const T* z, K* keys, size_t n, const CBox* box, void* s)
    {
        if constexpr (k64 && f64) return cstone_b200_compute_sfc_keys_k64_f64(kind, x, y, z, keys, n, box, s);
        else if constexpr (k64 && !f64) return cstone_b200_compute_sfc_keys_k64_f32(kind, x, y, z, keys, n, box, s);
        else if constexpr (!k64 && f64) return cstone_b200_compute_sfc_keys_k32_f64(kind, x, y, z, keys, n, box, s);
        else return cstone_b200_compute_sfc_keys_k32_f32(kind, x, y, z, keys, n, box, s);
    }
    static uint64_t sortTemp(uint64_t n)
    {
        if constexpr (k64) return cstone_b200_sort_by_key_temp_bytes_k64(n);
        else return cstone_b200_sort_by_key_temp_bytes_k32(n);
    }
    static int sort(K* keys, uint32_t* vals, K* kb, uint32_t* vb, size_t n, void* temp, uint64_t tb, int iota, void* s)
    {
        if constexpr (k64) return cstone_b200_sort_by_key_k64(keys, vals, kb, vb, n, temp, tb, iota, s);
        else return cstone_b200_sort_by_key_k32(keys, vals, kb, vb, n, temp, tb, iota, s);
    }
    static int updateOctree(const K* f, const K* l, unsigned bucket, K* ta, K* tb, uint32_t* counts, int32_t* ops,
                            int32_t cap, int32_t* n, int32_t* conv, void* temp, void* s)
    {
        if constexpr (k64) return cstone_b200_update_octree_k64(f, l, bucket, 0xFFFFFFFFu, ta, tb, counts, ops, cap, n, conv, temp, s);
        else return cstone_b200_update_octree_k32(f, l, bucket, 0xFFFFFFFFu, ta, tb, counts, ops, cap, n, conv, temp, s);
    }
    //! rebalance decision + scan only: new node count and "every op was keep" [ref: tree/csarray_gpu.h:75-77]
    static int nodeOps(const K* tree, int32_t n, const uint32_t* counts, unsigned bucket, int32_t* ops, int32_t* newN,
                       int32_t* conv, void* temp, void* s)
    {
        if constexpr (k64) return cstone_b200_compute_node_ops_k64(tree, n, counts, bucket, ops, newN, conv, temp, s);
        else return cstone_b200_compute_node_ops_k32(tree, n, counts, bucket, ops, newN, conv, temp, s);
    }
    static int nodeCounts(const K* tree, uint32_t* counts, int32_t n, const K* f, const K* l, void* s)
    {
        if constexpr (k64) return cstone_b200_compute_node_counts_k64(tree, counts, n, f, l, 0xFFFFFFFFu, 1, s);
        else return cstone_b200_compute_node_counts_k32(tree, counts, n, f, l, 0xFFFFFFFFu, 1, s);
    }
    static uint64_t octreeTemp(int32_t L)
    {
        if constexpr (k64) return cstone_b200_build_octree_temp_bytes_k64(L);
        else return cstone_b200_build_octree_temp_bytes_k32(L);
    }
    static int buildOctree(const K* leaves, int32_t L, K* prefixes, int32_t* co, int32_t* parents, int32_t* lr,
                           int32_t* itl, int32_t* lti, void* temp, uint64_t tb, void* s)
    {
        if constexpr (k64) return cstone_b200_build_octree_k64(leaves, L, prefixes, co, parents, lr, itl, lti, temp, tb, s);
        else return cstone_b200_build_octree_k32(leaves, L, prefixes, co, parents, lr, itl, lti, temp, tb, s);
    }
    static int geoCenters(int kind, const K* prefixes, int32_t M, T* centers, T* sizes, const CBox* box, void* s)
    {
        if constexpr (k64 && f64) return cstone_b200_compute_geo_centers_k64_f64(kind, prefixes, M, centers, sizes, box, s);
        else if constexpr (k64 && !f64) return cstone_b200_compute_geo_centers_k64_f32(kind, prefixes, M, centers, sizes, box, s);
        else if constexpr (!k64 && f64) return cstone_b200_compute_geo_centers_k32_f64(kind, prefixes, M, centers, sizes, box, s);
        else return cstone_b200_compute_geo_centers_k32_f32(kind, prefixes, M, centers, sizes, box, s);
    }
    static int minmax(const T* x, const T* y, const T* z, size_t n, T* out6, void* temp, void* s)
    {
        if constexpr (f64) return cstone_b200_minmax3_f64(x, y, z, n, out6, temp, s);
        else return cstone_b200_minmax3_f32(x, y, z, n, out6, temp, s);
    }
    static int haloRadii(const T* h, const uint32_t* seg, size_t n, float scale, float* radii, void* s)
    {
        if constexpr (f64) return cstone_b200_halo_radii_f64(h, seg, n, scale, radii, s);
        else return cstone_b200_halo_radii_f32(h, seg, n, scale, radii, s);
    }
    static int findHalos(int kind, const K* prefixes, const int32_t* co, const int32_t* itl, const K* leaves, const float* radii,
                         const CBox* box, int32_t first, int32_t last, int32_t* flags, void* s)
    {
        if constexpr (k64 && f64) return cstone_b200_find_halos_k64_f64(kind, prefixes, co, itl, leaves, radii, box, first, last, flags, s);
        else if constexpr (k64 && !f64) return cstone_b200_find_halos_k64_f32(kind, prefixes, co, itl, leaves, radii, box, first, last, flags, s);
        else if constexpr (!k64 && f64) return cstone_b200_find_halos_k32_f64(kind, prefixes, co, itl, leaves, radii, box, first, last, flags, s);
        else return cstone_b200_find_halos_k32_f32(kind, prefixes, co, itl, leaves, radii, box, first, last, flags, s);
    }
    static int neighbors(const T* x, const T* y, const T* z, const T* h, uint32_t first, uint32_t last, const CBox* box,
                         const int32_t* co, const int32_t* itl, const uint32_t* layout, const T* centers, const T* sizes,
                         uint32_t ngmax, uint32_t* nb, uint32_t* nc, int32_t numLeaves, void* temp, void* s)
    {
        if (!temp)
        {
            if constexpr (f64) return cstone_b200_find_neighbors_f64(x, y, z, h, first, last, box, co, itl, layout, centers, sizes, 1.0f, ngmax, nb, nc, s);
            else return cstone_b200_find_neighbors_f32(x, y, z, h, first, last, box, co, itl, layout, centers, sizes, 1.0f, ngmax, nb, nc, s);
        }
        if constexpr (f64) return cstone_b200_find_neighbors_ws_f64(x, y, z, h, first, last, box, co, itl, layout, centers, sizes, 1.0f, ngmax, nb, nc, numLeaves, temp, s);
        else return cstone_b200_find_neighbors_ws_f32(x, y, z, h, first, last, box, co, itl, layout, centers, sizes, 1.0f, ngmax, nb, nc, numLeaves, temp, s);
    }
};

//! device-resident cornerstone leaf array with ping-pong buffers  [ref: tree/update_gpu.cuh:58-82]
template<class K, class T>
struct LeafArrayDev
{
    DevBuf treeA, treeB, counts, nodeOps, temp;
    int capacity = 0, numNodes = 0;

    int ensure(int cap, cudaStream_t s)
    {
        if (cap <= capacity) return 0;
        cap = int(cap * 1.05) + 64;
        CSB_TRY(treeA.reserve(sizeof(K) * (size_t(cap) + 1), true, s));
        CSB_TRY(treeB.reserve(sizeof(K) * (size_t(cap) + 1)));
        CSB_TRY(counts.reserve(4 * size_t(cap), true, s));
        CSB_TRY(nodeOps.reserve(4 * (size_t(cap) + 1)));
        CSB_TRY(temp.reserve(cstone_b200_scan_temp_bytes(uint64_t(cap) + 1) + 1024));
        capacity = cap;
        return 0;
    }
    //! reset to the root node holding `count` particles
    int reset(uint32_t count, cudaStream_t s)
    {
        CSB_TRY(ensure(1024, s));
        K root[2] = {K(0), nodeRange<K>(0)};
        CSB_CUDA(cudaMemcpyAsync(treeA.p, root, sizeof(root), cudaMemcpyHostToDevice, s));
        CSB_CUDA(cudaMemcpyAsync(counts.p, &count, 4, cudaMemcpyHostToDevice, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        numNodes = 1;
        return 0;
    }
    /*! one rebalance + recount step [ref: tree/csarray.hpp:429-448]; counts must hold the counts to decide on.
     *  countsAreCurrent: the counts were computed on [first, last) for THIS tree — then a decision of "keep" for every
     *  node leaves tree and counts as they are, and neither the tree is written again nor the keys are counted again */
    int update(const K* first, const K* last, unsigned bucket, bool& converged, cudaStream_t s, bool countsAreCurrent = false)
    {
        if (countsAreCurrent)
        {
            int32_t newN = 0, conv = 0;
            CSB_TRY((Api<K, T>::nodeOps(treeA.as<K>(), numNodes, counts.as<uint32_t>(), bucket, nodeOps.as<int32_t>(), &newN, &conv,
                                        temp.p, s)));
            if (conv && newN == numNodes)
            {
                converged = true;
                return 0;
            }
        }
        for (;;)
        {
            int32_t n = numNodes, conv = 0;
            int st = Api<K, T>::updateOctree(first, last, bucket, treeA.as<K>(), treeB.as<K>(), counts.as<uint32_t>(),
                                             nodeOps.as<int32_t>(), capacity, &n, &conv, temp.p, s);
            if (st == CSTONE_E_CAPACITY)
            {
                CSB_TRY(ensure(n, s));
                continue;
            }
            if (st) return st;
            numNodes  = n;
            converged = conv != 0;
            return 0;
        }
    }
};


// ------------------------------------------------------------------------------------------------
// halo discovery kernels (multi-rank only)
// ------------------------------------------------------------------------------------------------
/*! halo boxes of the local leaves that reach outside the key range [lowest, highest) of this rank
 *  [ref: traversal/collisions.hpp:78-105: same box and the same containedIn early-out as findHalos] */
template<class K, class T>
__global__ void __launch_bounds__(128) surfaceBoxesKernel(int kind,
                                                          const uint16_t* __restrict__ enc,
                                                          const uint16_t* __restrict__ dec,
                                                          const K* __restrict__ leaves,
                                                          const uint32_t* __restrict__ counts,
                                                          const float* __restrict__ radii,
                                                          int numLeaves,
                                                          K lowest,
                                                          K highest,
                                                          DBox<T> box,
                                                          IBox6* __restrict__ out,
                                                          uint32_t capacity,
                                                          uint32_t* __restrict__ numOut)
{
    for (int leaf = blockIdx.x * blockDim.x + threadIdx.x; leaf < numLeaves; leaf += gridDim.x * blockDim.x)
    {
        if (counts[leaf] == 0) continue;
        // every non-empty leaf of this tree holds particles of this rank, also one that straddles a rank boundary
        // (sparse regions, ranks with fewer particles than the bucket size): containedIn rejects interior leaves
        K start = leaves[leaf], end = leaves[leaf + 1];
        IBox nodeBox = sfcIBox<K>(kind, dec, start, treeLevel<K>(K(end - start)));
        IBox halo    = makeHaloBox<K, T>(nodeBox, radii[leaf], box);
        if (containedIn<K>(kind, enc, lowest, highest, halo)) continue;
        uint32_t slot = atomicAdd(numOut, 1u);
        if (slot < capacity)
        {
            IBox6 b;
#pragma unroll
            for (int d = 0; d < 3; ++d)
                b.v[2 * d] = halo.lo[d], b.v[2 * d + 1] = halo.hi[d];
            out[slot] = b;
        }
    }
}

/*! reverse collision: which of MY non-empty leaves does a foreign halo box overlap?  One thread per foreign box,
 *  depth-first over my linked tree, nodes outside my key range are skipped.  flags[srcRank * numLeaves + leaf] = 1 */
template<class K>
__global__ void __launch_bounds__(128) reverseCollisionKernel(int kind,
                                                              const uint16_t* __restrict__ dec,
                                                              const K* __restrict__ prefixes,
                                                              const int* __restrict__ childOffsets,
                                                              const int* __restrict__ internalToLeaf,
                                                              const uint32_t* __restrict__ counts,
                                                              int numLeaves,
                                                              K lowest,
                                                              K highest,
                                                              const IBox6* __restrict__ boxes,  // [numRanks][stride]
                                                              const uint32_t* __restrict__ numBoxes, // [numRanks]
                                                              uint32_t stride,
                                                              int numRanks,
                                                              int myRank,
                                                              unsigned char* __restrict__ flags)
{
    const int src = blockIdx.y;
    if (src == myRank) return;
    const uint32_t nb = min(numBoxes[src], stride);
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < nb; b += gridDim.x * blockDim.x)
    {
        IBox6 fb = boxes[size_t(src) * stride + b];
        IBox halo;
#pragma unroll
        for (int d = 0; d < 3; ++d)
            halo.lo[d] = fb.v[2 * d], halo.hi[d] = fb.v[2 * d + 1];
        unsigned char* myFlags = flags + size_t(src) * numLeaves;

        auto descend = [&](int idx) -> bool
        {
            K prefix       = prefixes[idx];
            unsigned level = decodePrefixLength<K>(prefix) / 3u;
            K nodeStart    = decodePlaceholderBit<K>(prefix);
            K nodeEnd      = nodeStart + nodeRange<K>(level);
            if (nodeEnd <= lowest || nodeStart >= highest) return false; // nothing of mine below this node
            return overlapBoxes<K>(sfcIBox<K>(kind, dec, nodeStart, level), halo);
        };
        auto hitLeaf = [&](int idx)
        {
            int leaf = internalToLeaf[idx];
            if (counts[leaf]) myFlags[leaf] = 1;
        };

        if (!descend(0)) continue;
        if (childOffsets[0] == 0)
        {
            hitLeaf(0);
            continue;
        }
        int stack[160];
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
                if (childOffsets[child] == 0) hitLeaf(child);
                else stack[sp++] = child;
            }
            node = stack[--sp];
        } while (node != 0);
    }
}

//! per (peer, leaf), peers concatenated: number of particles to send = flag ? count : 0 (one scan for all peers)
struct LoadFlaggedCount
{
    const unsigned char* flags; // [numRanks][numLeaves]
    const uint32_t* counts;
    size_t numLeaves, total;
    __device__ uint32_t operator()(size_t i) const
    {
        if (i >= total || !flags[i]) return 0u;
        return counts[i % numLeaves];
    }
};

//! expand the flagged leaves of all peers into particle indices (positions in the SFC-ordered assigned arrays);
//! sendScan is the exclusive scan over (peer, leaf), so the lists of the peers follow each other in rank order
__global__ void __launch_bounds__(256) expandSendListKernel(const unsigned char* __restrict__ flags,
                                                            const uint32_t* __restrict__ layout,
                                                            const uint32_t* __restrict__ sendScan,
                                                            size_t numLeaves,
                                                            size_t total,
                                                            uint32_t* __restrict__ sendIdx)
{
    for (size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += size_t(gridDim.x) * blockDim.x)
    {
        if (!flags[idx]) continue;
        const size_t leaf = idx % numLeaves;
        const uint32_t a = layout[leaf], b = layout[leaf + 1], o = sendScan[idx];
        for (uint32_t k = 0; k < b - a; ++k)
            sendIdx[o + k] = a + k;
    }
}

struct Elem12
{
    uint32_t v[3];
};
//! dst[i] = src[idx[i] + offset] for 4-, 8-, 12- or 16-byte elements (pack of halo messages)
template<class E>
__global__ void __launch_bounds__(256)
    packKernel(const uint32_t* __restrict__ idx, uint32_t n, uint32_t offset, const E* __restrict__ src, E* __restrict__ dst)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        dst[i] = src[idx[i] + offset];
}



// ------------------------------------------------------------------------------------------------
// pack-and-push halo exchange (n1): ONE kernel gathers the outgoing halo elements of all arrays and stores them
// straight into the receive windows of the destination ranks over NVLink (peer pointers from CUDA IPC), then raises
// a flag in every peer's window; ONE kernel on the receiving side waits for the flags and moves the window contents
// into the halo ranges of the arrays.  Replaces gather-per-array + grouped send/recv
// [ref: halos/exchange_halos_gpu.cuh:34-119 haloExchangeGpu: gatherRanges per array, MPI_Isend/Irecv, unpack].
//
// Window of rank q (device memory of q, mapped by every peer):
//   [0, 4096)   header: flags[2][P] (uint64: last epoch whose data from rank r is complete, per slot), ticket, status
//   slot 0/1    data of the exchanges with even / odd epoch; per array a region of align256(recvTotal[q] * elemBytes[a])
//               bytes, inside it the elements of source rank r start at element sum(count[r' -> q], r' < r).
// ------------------------------------------------------------------------------------------------
//! out[rangeOffset[r] + k] = rangeStart[r] + k for k < rangeOffset[r + 1] - rangeOffset[r]: index ranges written out
__global__ void __launch_bounds__(256) expandRangesKernel(const uint32_t* __restrict__ rangeStart, const uint32_t* __restrict__ rangeOffset,
                                                          int numRanges, uint32_t total, uint32_t* __restrict__ out)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x)
    {
        int lo = 0, hi = numRanges; // the range with rangeOffset[r] <= i < rangeOffset[r + 1]
        while (hi - lo > 1)
        {
            const int mid = (lo + hi) / 2;
            if (rangeOffset[mid] <= i) lo = mid;
            else hi = mid;
        }
        out[i] = rangeStart[lo] + (i - rangeOffset[lo]);
    }
}

//! [ref: focus/octree_focus_mpi.hpp:600-610 addMacs] leaves whose node failed the MAC become halos too
__global__ void __launch_bounds__(256)
    addMacsKernel(const char* __restrict__ macs, const int32_t* __restrict__ leafToInternal, int numLeaves, int32_t* __restrict__ flags)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < numLeaves && macs[leafToInternal[i]] && !flags[i]) flags[i] = 1;
}

constexpr int kMaxPushRanks  = 64;
constexpr int kMaxPushArrays = 16;
constexpr size_t kWinHeader  = 4096;

struct PushArgs
{
    int na, P, me;
    uint32_t nSend, start;
    const void* src[kMaxPushArrays];
    int words[kMaxPushArrays];
    uint32_t sendOffset[kMaxPushRanks + 1];
    uint32_t dstElemOff[kMaxPushRanks]; // where my elements start inside rank q's regions
    uint32_t recvTotal[kMaxPushRanks];  // elements per region of rank q
    char* peerData[kMaxPushRanks];      // data slot of this epoch in rank q's window
    unsigned long long* peerFlag[kMaxPushRanks]; // flag (slot, me) in rank q's window
    unsigned int* ticket;
    unsigned long long epoch;
};

struct UnpackArgs
{
    int na, P, me, spin;
    void* dst[kMaxPushArrays];
    int words[kMaxPushArrays];
    uint32_t recvCount[kMaxPushRanks], recvOffset[kMaxPushRanks], srcElemOff[kMaxPushRanks + 1];
    const char* myData;
    const unsigned long long* myFlags; // flags of this slot, one per source rank
    unsigned long long epoch;
    unsigned int* status;
};

__host__ __device__ inline size_t pushRegionOffset(const int* words, int a, uint32_t total)
{
    size_t off = 0;
    for (int k = 0; k < a; ++k)
        off += (size_t(total) * size_t(words[k]) * 4 + 255) / 256 * 256;
    return off;
}

template<int W>
__device__ inline void copyWords(const uint32_t* __restrict__ s, uint32_t* __restrict__ d)
{
    if (W == 1) *d = *s;
    else if (W == 2) *reinterpret_cast<uint2*>(d) = *reinterpret_cast<const uint2*>(s);
    else if (W == 4) *reinterpret_cast<uint4*>(d) = *reinterpret_cast<const uint4*>(s);
    else
    {
        d[0] = s[0], d[1] = s[1], d[2] = s[2];
    }
}
__device__ inline void copyElem(int words, const uint32_t* s, uint32_t* d)
{
    if (words == 1) copyWords<1>(s, d);
    else if (words == 2) copyWords<2>(s, d);
    else if (words == 4) copyWords<4>(s, d);
    else copyWords<3>(s, d);
}

__global__ void __launch_bounds__(256) pushHalosKernel(const uint32_t* __restrict__ sendIdx, const __grid_constant__ PushArgs p)
{
    const int a     = blockIdx.y;
    const int words = p.words[a];
    __shared__ size_t regionOff[kMaxPushRanks];
    for (int q = threadIdx.x; q < p.P; q += blockDim.x)
        regionOff[q] = pushRegionOffset(p.words, a, p.recvTotal[q]);
    __syncthreads();
    const uint32_t* src = static_cast<const uint32_t*>(p.src[a]);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < p.nSend; i += gridDim.x * blockDim.x)
    {
        int lo = 0, hi = p.P; // destination: the q with sendOffset[q] <= i < sendOffset[q + 1]
        while (hi - lo > 1)
        {
            const int mid = (lo + hi) / 2;
            if (p.sendOffset[mid] <= i) lo = mid;
            else hi = mid;
        }
        const int q         = lo;
        const size_t elem   = size_t(p.dstElemOff[q]) + (i - p.sendOffset[q]);
        uint32_t* d         = reinterpret_cast<uint32_t*>(p.peerData[q] + regionOff[q]) + elem * size_t(words);
        const uint32_t* sEl = src + (size_t(sendIdx[i]) + p.start) * size_t(words);
        copyElem(words, sEl, d);
    }
    // the last block of the grid tells every peer that my data of this epoch has arrived
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0)
    {
        const unsigned int blocks = gridDim.x * gridDim.y;
        if (atomicAdd(p.ticket, 1u) == blocks - 1)
        {
            *p.ticket = 0;
            __threadfence_system();
            for (int q = 0; q < p.P; ++q)
                if (q != p.me) *reinterpret_cast<volatile unsigned long long*>(p.peerFlag[q]) = p.epoch;
            __threadfence_system();
        }
    }
}

__global__ void __launch_bounds__(256) unpackHalosKernel(const __grid_constant__ UnpackArgs u)
{
    if (u.spin)
    {
        // every source rank has finished writing this epoch into my window (bounded wait: a lost flag must not hang the GPU)
        if (threadIdx.x < u.P && int(threadIdx.x) != u.me)
        {
            const volatile unsigned long long* f = u.myFlags + threadIdx.x;
            unsigned long long t0;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            while (*f < u.epoch)
            {
                unsigned long long t1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 20000000000ull)
                {
                    atomicExch(u.status, 1u);
                    break;
                }
            }
        }
        __threadfence_system();
        __syncthreads();
    }
    const int a     = blockIdx.y;
    const int words = u.words[a];
    const uint32_t total = u.srcElemOff[u.P];
    const uint32_t* region = reinterpret_cast<const uint32_t*>(u.myData + pushRegionOffset(u.words, a, total));
    uint32_t* dst = static_cast<uint32_t*>(u.dst[a]);
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < total; j += gridDim.x * blockDim.x)
    {
        int lo = 0, hi = u.P;
        while (hi - lo > 1)
        {
            const int mid = (lo + hi) / 2;
            if (u.srcElemOff[mid] <= j) lo = mid;
            else hi = mid;
        }
        const size_t di = size_t(u.recvOffset[lo]) + (j - u.srcElemOff[lo]);
        copyElem(words, region + size_t(j) * size_t(words), dst + di * size_t(words));
    }
}

// ------------------------------------------------------------------------------------------------
// the engine
// ------------------------------------------------------------------------------------------------
constexpr int kMaxProps = 12;

template<class K, class T>
class Engine
{
public:
    using CBox = typename BoxT<T>::type;

    Engine(int rank, int nRanks, unsigned bucket, unsigned bucketFocus, float theta, const CBox& box)
        : rank_(rank), nRanks_(nRanks), bucket_(bucket), bucketFocus_(bucketFocus), theta_(theta), box_(box)
    {
    }
    //! one-shot: the stream waits for `e` before h and the properties are first read (their upload may still run)
    void setFieldsReadyEvent(void* e) { fieldsReady_ = reinterpret_cast<cudaEvent_t>(e); }
    void setComm(void* c) { comm_.reset(new NcclComm(reinterpret_cast<ncclComm_t>(c), rank_, nRanks_)); }
    //! P ranks as P host threads of this process on one GPU (comm.hpp LocalComm): test transport
    void setLocalWorld(LocalWorld* w) { comm_.reset(new LocalComm(w, rank_)); }
    ~Engine()
    {
        for (cudaEvent_t e : ev_)
            if (e) cudaEventDestroy(e);
        /* windows that were exported to other processes are NOT freed here: a peer may still have them mapped, and
         * freeing exported memory before every importer closed it is undefined; they go with the process (an engine
         * lives as long as its process in every client we know).  The peers' mappings in THIS process are closed. */
        if (comm_ && !winPeers_.empty()) comm_->releaseShared(winPeers_);
        if (comm_ && comm_->hostSynchronous())
        {
            if (win_) cudaFree(win_);
            for (void* w : winOld_)
                cudaFree(w);
        }
    }
    //! halo exchange transport: 1 = pack-and-push over peer memory (default), 0 = pack + grouped send/recv
    int setHaloPush(int on)
    {
        if (on != 0 && on != 1) return CSTONE_E_INVALID_ARGUMENT;
        haloPush_ = on;
        return 0;
    }
    //! 1 if a receiving kernel gave up waiting for a peer's data (pack-and-push), 0 otherwise; synchronises the stream
    int pushStatus(int32_t* out)
    {
        *out = 0;
        if (!win_) return 0;
        unsigned int st = 0;
        CSB_CUDA(cudaMemcpyAsync(&st, static_cast<char*>(win_) + 3080, 4, cudaMemcpyDeviceToHost, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        *out = int32_t(st);
        return 0;
    }
    //! space-filling curve of the keys (the reference's Domain hard-wires Hilbert, sfc/sfc.hpp:54-55); before the first sync
    int setSfcKind(int kind)
    {
        if (kind != CSTONE_SFC_MORTON && kind != CSTONE_SFC_HILBERT) return CSTONE_E_INVALID_ARGUMENT;
        sfcKind_ = kind;
        return 0;
    }
    //! [ref: domain/domain.hpp:396 setEndIndex] particles appended behind the assigned range join it at the next sync
    int setEndIndex(uint64_t i)
    {
        if (i < start_ || i > nTotal_) return CSTONE_E_INVALID_ARGUMENT;
        end_ = i;
        return 0;
    }
    //! [ref: domain/domain.hpp:412 setHaloFactor, halos/halos.hpp:162] halo search radius = 2 * factor * max h of the leaf
    int setHaloFactor(float f)
    {
        if (!(f > 0.f)) return CSTONE_E_INVALID_ARGUMENT;
        haloFactor_ = f;
        return 0;
    }
    /*! how halos are found at nRanks > 1.  0 (default): surface boxes of every rank against the leaves of the assigned
     *  particles (fast path, halo set a subset of the reference's).  1: the reference's way -- the multi-rank focus tree
     *  (LET, let_tree.cuh) + findHalos on it + key-range requests between peers [ref: domain/domain.hpp:217-243]; then
     *  halo flags, layout, startIndex / endIndex and the tree handed to findNeighbors are the reference's. */
    int setHaloMode(int mode)
    {
        if ((mode != 0 && mode != 1) || !firstCall_) return CSTONE_E_INVALID_ARGUMENT;
        haloMode_ = mode;
        return 0;
    }
    //! LET mode: host copies of the focus-tree leaves, their counts, the halo flags, the layout, the peer assignment
    const LetTree<K, T>* let() const { return let_.get(); }
    const PinnedVec<int32_t>& letFlags() const { return hFlags_; }
    const PinnedVec<uint32_t>& letLayout() const { return hLayout_; }
    //! [ref: domain/domain.hpp:404-406 startCell / endCell] leaves of the search tree that hold the assigned particles
    int cells(int32_t* out2)
    {
        if (!search_.numNodes || boundaries_.empty()) return CSTONE_E_INVALID_ARGUMENT;
        CSB_TRY(small_.reserve(cstone_b200_reduce_temp_bytes() + 256));
        uint64_t r[2];
        const K b[2] = {boundaries_[rank_], boundaries_[rank_ + 1]};
        for (int k = 0; k < 2; ++k)
        {
            if constexpr (sizeof(K) == 8)
                CSB_TRY(cstone_b200_lower_bound_u64(search_.treeA.template as<uint64_t>(), size_t(search_.numNodes) + 1, b[k], &r[k], small_.p, s_));
            else
                CSB_TRY(cstone_b200_lower_bound_u32(search_.treeA.template as<uint32_t>(), size_t(search_.numNodes) + 1, b[k], &r[k], small_.p, s_));
        }
        out2[0] = int32_t(r[0]), out2[1] = int32_t(r[1]);
        return 0;
    }
    //! device time between the stage markers of the last sync (ms); kNumMarks - 1 values
    void stageMs(float* ms) const
    {
        for (int k = 0; k + 1 < kNumMarks; ++k)
        {
            ms[k] = 0.f;
            if (ev_[k] && ev_[k + 1]) cudaEventElapsedTime(&ms[k], ev_[k], ev_[k + 1]);
        }
    }
    static constexpr int kNumMarks = 11;

    int64_t syncBegin(uint64_t nIn, const T* x, const T* y, const T* z, const T* h, int nprops, const void* const* props,
                      const int* propBytes);
    /*! [ref: domain/domain.hpp:245-325 syncGrav] sync with the focus tree refined by the vector MAC around the
     *  expansion (mass) centres; props[0] must be the masses (element type T).  Needs halo mode 1 and nRanks > 1. */
    int64_t syncGravBegin(uint64_t nIn, const T* x, const T* y, const T* z, const T* h, int nprops, const void* const* props,
                          const int* propBytes)
    {
        if (haloMode_ != 1 || nprops < 1 || propBytes[0] != int(sizeof(T))) return CSTONE_E_INVALID_ARGUMENT;
        grav_           = true;
        const int64_t r = syncBegin(nIn, x, y, z, h, nprops, props, propBytes);
        grav_           = false;
        return r;
    }
    // ---- globalFocusExchange for client quantities [ref: focus/octree_focus_mpi.hpp:276-300 peerExchange, :762-784]
    static bool elemOk(int eb) { return eb == 4 || eb == 8 || eb == 12 || eb == 16 || eb == 24 || eb == 32; }
    //! nodeQ: one element per node of the focus tree (device).  Cells in a peer's range take the owner's values.
    int focusPeerExchange(void* nodeQ, int eb)
    {
        if (haloMode_ != 1 || !let_ || !nodeQ || !elemOk(eb)) return CSTONE_E_INVALID_ARGUMENT;
        CSB_TRY(let_->peerExchange(comm_.get(), nodeQ, eb, s_));
        return int(cudaStreamSynchronize(s_));
    }
    //! gLeafQ: one element per GLOBAL leaf (device): my leaves from nodeQ, everybody else's from their owners
    int focusGlobalGather(const void* nodeQ, int eb, void* gLeafQ)
    {
        if (haloMode_ != 1 || !let_ || !nodeQ || !gLeafQ || !elemOk(eb)) return CSTONE_E_INVALID_ARGUMENT;
        CSB_TRY(let_->linkGlobal(hGlobalLeaves_, s_));
        CSB_TRY(let_->populateGlobal(nodeQ, eb, hGlobalLeaves_, gLeafQ, nullptr, s_));
        CSB_TRY(let_->gatherGlobalLeaves(comm_.get(), gLeafQ, eb, hGlobalLeaves_, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        return let_->takeError() ? CSTONE_E_LET : 0;
    }
    //! gNodeQ: one element per node of the global octree (globalOctree()); cells of my tree that belong to no peer read theirs
    int focusGlobalExtract(const void* gNodeQ, int eb, void* nodeQ)
    {
        if (haloMode_ != 1 || !let_ || !nodeQ || !gNodeQ || !elemOk(eb)) return CSTONE_E_INVALID_ARGUMENT;
        CSB_TRY(let_->extractGlobal(gNodeQ, eb, nodeQ, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        return let_->takeError() ? CSTONE_E_LET : 0;
    }
    //! the replicated global tree, linked on the device (valid after focusGlobalGather or syncGrav, until the next sync)
    int globalOctree(int32_t* numLeaves, int32_t* numNodes, const void** p6, int32_t* levelRangeHost) const
    {
        if (!let_ || let_->globalTree().numLeaves_ == 0) return CSTONE_E_INVALID_ARGUMENT;
        const FocusTree<K>& g = let_->globalTree();
        *numLeaves = g.numLeaves_, *numNodes = g.numNodes_;
        p6[0] = g.prefixes_.p, p6[1] = g.childOffsets_.p, p6[2] = g.parents_.p, p6[3] = g.levelRangeDev_.p, p6[4] = g.i2l_.p, p6[5] = g.l2i_.p;
        for (unsigned l = 0; l < KeyTraits<K>::maxLevel + 2; ++l)
            levelRangeHost[l] = g.levelRange_[l];
        return 0;
    }
    //! syncGrav: expansion centres (x, y, z, MAC radius^2) per node of the focus tree and its MAC marks (device)
    void letCenters(const void** centers4, const char** macs, int32_t* numNodes) const
    {
        *centers4 = let_ ? let_->t.centers_.p : nullptr;
        *macs     = let_ ? let_->t.macs_.template as<char>() : nullptr;
        *numNodes = let_ ? let_->t.numNodes_ : 0;
    }
    int syncFinish(K* keys, T* x, T* y, T* z, T* h, void* const* props);
    int exchangeHalos(int narrays, void* const* arrays, const int* elemBytes);
    int reapplySync(int narrays, const void* const* in, const int* elemBytes, void* const* out);
    int findNeighbors(const T* x, const T* y, const T* z, const T* h, uint32_t ngmax, uint32_t* nb, uint32_t* nc) const;

    void info(int64_t* o) const
    {
        o[0] = start_, o[1] = end_, o[2] = nTotal_, o[3] = search_.numNodes;
        o[4] = search_.numNodes + (search_.numNodes - 1) / 7, o[5] = global_.numNodes;
        o[6] = nHaloSend_, o[7] = nTotal_ - (end_ - start_);
    }
    const CBox& box() const { return box_; }
    void treeView(const void** p) const
    {
        p[0] = search_.treeA.p, p[1] = search_.counts.p, p[2] = layout_.p, p[3] = prefixes_.p, p[4] = childOffsets_.p;
        p[5] = internalToLeaf_.p, p[6] = leafToInternal_.p, p[7] = levelRange_.p, p[8] = centers_.p, p[9] = sizes_.p;
    }
    void globalTree(const K** l, const uint32_t** c, const K** a) const
    {
        *l = hGlobalLeaves_.data(), *c = hGlobalCounts_.data(), *a = boundaries_.data();
    }

private:
    void mark(int k)
    {
        if (!ev_[k]) cudaEventCreate(&ev_[k]);
        cudaEventRecord(ev_[k], s_);
    }
    cudaEvent_t ev_[kNumMarks] = {nullptr};
    int updateBox(const T* x, const T* y, const T* z, size_t n);
    int updateGlobalTree(const K* keys, size_t n);
    int allreduceCounts();
    //! internal tree + particle layout of a leaf array; one set per tree, so that a tree whose leaves did not change
    //! between two syncs is not linked again (the leaf COUNTS change every step, the leaves rarely do)
    struct Linked
    {
        DevBuf prefixes, childOffsets, parents, levelRange, internalToLeaf, leafToInternal, layout;
        bool valid = false;
    };
    int linkTree(LeafArrayDev<K, T>& tree, Linked& ln, bool withCenters, bool leavesChanged);
    int convergeTree(LeafArrayDev<K, T>& tree, const K* keys, size_t n, bool& fresh, int maxSteadyIter = 64,
                     bool* leavesChanged = nullptr);
    int discoverHalos();
    int discoverHalosLet();
    int exchangeHalosPush(int narrays, void* const* arrays, const int* elemBytes);
    int ensureWindow(size_t slotBytes);
    int hostCounts(const uint32_t* dev, int n, std::vector<uint32_t>& out)
    {
        out.resize(n);
        CSB_CUDA(cudaMemcpyAsync(out.data(), dev, size_t(n) * 4, cudaMemcpyDeviceToHost, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        return 0;
    }
    int gatherMulti(const uint32_t* map, size_t n, int na, const void* const* src, void* const* dst, const int* eb)
    {
        return cstone_b200_gather_multi(map, n, na, src, dst, eb, s_);
    }
    //! the assigned set in SFC order: from the receive buffers through order_, or from two sources through map2_
    int gatherAssigned(size_t n, int na, const int* which, void* const* dst, const int* eb)
    {
        if (n == 0) return 0;
        if (!map2_)
        {
            const void* src[4 + kMaxProps];
            int ebs[4 + kMaxProps];
            for (int k = 0; k < na; ++k)
                src[k] = recv_[which[k]].p, ebs[k] = eb[which[k]];
            return gatherMulti(order_.template as<uint32_t>(), n, na, src, dst, ebs);
        }
        GatherTwoArgs a{};
        a.na = na;
        for (int k = 0; k < na; ++k)
        {
            a.orig[k] = origSrc_[which[k]], a.recv[k] = recv_[which[k]].p, a.dst[k] = dst[k];
            a.words[k] = eb[which[k]] / 4;
        }
        gatherTwoKernel<<<gridFor(n, 256, 8), 256, 0, s_>>>(map2_, n, a);
        CSB_LAUNCHED();
        return 0;
    }

    int rank_, nRanks_;
    unsigned bucket_, bucketFocus_;
    float theta_;
    CBox box_;
    bool firstCall_   = true;
    std::unique_ptr<Comm> comm_;
    cudaEvent_t fieldsReady_ = nullptr;
    int sfcKind_      = CSTONE_SFC_HILBERT;
    float haloFactor_ = 1.0f;
    int haloMode_     = 0;
    //! the multi-rank code path (assignment, exchange, focus tree, halos) runs, possibly over a world of one rank
    bool multi() const { return nRanks_ > 1 || haloMode_ == 1; }
    std::unique_ptr<LocalWorld> selfWorld_;
    bool grav_        = false; // this sync is a syncGrav: property 0 is the mass
    bool gravReady_   = false; // the expansion centres belong to the current tree (the last sync was a syncGrav)
    float centerDriftTol_ = 1.05f; // [ref: domain/domain.hpp:665]
    DevBuf sgx_, sgy_, sgz_, sgm_; // x, y, z, m of the assigned particles in SFC order (syncGrav: expansion centres)
    std::unique_ptr<LetTree<K, T>> let_;
    PinnedVec<int32_t> hFlags_;   // host copies in page-locked memory
    PinnedVec<uint32_t> hLayout_;
    DevBuf letLayoutDev_;
    cudaStream_t s_   = nullptr; // legacy default stream: ordered with the caller's default-stream work

    // particle state
    uint64_t start_ = 0, end_ = 0, nTotal_ = 0, nAssigned_ = 0, nHaloSend_ = 0;
    int nprops_ = 0;
    int propBytes_[kMaxProps] = {0};

    // scratch + SFC-ordered arrays
    DevBuf keys_, keyBuf_, order_, valBuf_, sortTemp_, mmTemp_, small_;
    // steady-state merge: the assigned set is gathered from two sources through map2_ (see mergeScatterKernel)
    const uint32_t* map2_ = nullptr;
    const void* origSrc_[4 + 12] = {nullptr};
    // exchange pattern of the last sync, kept for reapplySync [ref: domain/domain.hpp:334-378]
    struct Pattern
    {
        bool valid = false, steady = false;
        uint64_t inStart = 0, np = 0, nAssigned = 0, start = 0;
        std::vector<uint32_t> sendStart, sendCount, recvOff, recvCount;
    } pat_;
    DevBuf order1Keep_; // map of the first sort when the second sort overwrites order_ (first sync / large exchanges)
    DevBuf pack_[4 + kMaxProps], recv_[4 + kMaxProps], fin_[4 + kMaxProps];
    DevBuf skeys_, fkeys_, shAssigned_, merge_;

    // trees
    LeafArrayDev<K, T> global_, focus_, search_;
    bool globalFresh_ = true, focusFresh_ = true, searchFresh_ = true;
    std::vector<K> hGlobalLeaves_, boundaries_, oldBoundaries_;
    std::vector<uint32_t> hGlobalCounts_, rankCounts_;
    Linked lnS_, lnF_; // of search_ (assigned + halos) and of focus_ (assigned, halo discovery)
    DevBuf &prefixes_ = lnS_.prefixes, &childOffsets_ = lnS_.childOffsets, &parents_ = lnS_.parents,
           &levelRange_ = lnS_.levelRange, &internalToLeaf_ = lnS_.internalToLeaf, &leafToInternal_ = lnS_.leafToInternal,
           &layout_ = lnS_.layout;
    DevBuf centers_, sizes_;
    DevBuf octTemp_, scanTemp_, radii_;
    mutable DevBuf nbInfo_; // node records of findNeighbors

    // halo exchange pattern
    DevBuf surf_, allSurf_, nSurf_, flags_, sendScan_, sendIdx_, haloPack_;
    std::vector<uint32_t> haloSendCount_, haloSendOffset_, haloRecvCount_, haloRecvOffset_;
    std::vector<uint32_t> haloMatrix_; // [src * P + dst]: halo particles src sends to dst (the same on every rank)
    // pack-and-push windows (see pushHalosKernel)
    int haloPush_ = -1; // -1: decide at first use (on unless CSB_HALO_PUSH=0), 0 / 1
    void* win_    = nullptr;
    size_t winSlot_ = 0;
    std::vector<void*> winPeers_, winOld_;
    unsigned long long winEpoch_ = 0;
};

template<class K, class T>
int Engine<K, T>::updateBox(const T* x, const T* y, const T* z, size_t n)
{
    // [ref: sfc/box_mpi.hpp:85-119 makeGlobalBox, domain/assignment.hpp:80-82]
    T mm[6];
    CSB_TRY(mmTemp_.reserve(cstone_b200_minmax_temp_bytes() + 64));
    if (n) CSB_TRY((Api<K, T>::minmax(x, y, z, n, mm, mmTemp_.p, s_)));
    else
        for (int d = 0; d < 3; ++d)
            mm[2 * d] = T(INFINITY), mm[2 * d + 1] = -T(INFINITY);
    bool pbc[3], anyOpen = false;
    for (int d = 0; d < 3; ++d)
    {
        pbc[d] = box_.boundaries[d] == CSTONE_BOUNDARY_PERIODIC;
        if (pbc[d]) mm[2 * d] = box_.limits[2 * d], mm[2 * d + 1] = box_.limits[2 * d + 1];
        anyOpen |= !pbc[d];
    }
    if (anyOpen && nRanks_ > 1)
    {
        T v[6] = {mm[0], -mm[1], mm[2], -mm[3], mm[4], -mm[5]};
        CSB_TRY(small_.reserve(4096));
        CSB_CUDA(cudaMemcpyAsync(small_.p, v, sizeof(v), cudaMemcpyHostToDevice, s_));
        CSB_TRY(comm_->allReduce(small_.p, 6, sizeof(T) == 8 ? CommType::f64 : CommType::f32, CommOp::min, s_));
        CSB_CUDA(cudaMemcpyAsync(v, small_.p, sizeof(v), cudaMemcpyDeviceToHost, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        for (int d = 0; d < 3; ++d)
            mm[2 * d] = v[2 * d], mm[2 * d + 1] = -v[2 * d + 1];
    }
    T lim[6];
    if (firstCall_)
        for (int k = 0; k < 6; ++k)
            lim[k] = mm[k];
    else limitBoxShrinking<T>(mm, box_.limits, lim);
    for (int d = 0; d < 3; ++d)
    {
        box_.limits[2 * d]     = lim[2 * d];
        box_.limits[2 * d + 1] = lim[2 * d + 1];
        box_.lengths[d]        = lim[2 * d + 1] - lim[2 * d];
        box_.inv_lengths[d]    = T(1) / box_.lengths[d];
    }
    return 0;
}

template<class K, class T>
int Engine<K, T>::allreduceCounts()
{
    if (nRanks_ == 1) return 0;
    // [ref: tree/update_mpi.hpp:58-66] counts = max(local, Allreduce(SUM)(local))
    int n = global_.numNodes;
    CSB_TRY(small_.reserve(size_t(n) * 4 + 4096));
    CSB_CUDA(cudaMemcpyAsync(small_.p, global_.counts.p, size_t(n) * 4, cudaMemcpyDeviceToDevice, s_));
    CSB_TRY(comm_->allReduce(small_.p, size_t(n), CommType::u32, CommOp::sum, s_));
    maxMergeKernel<<<gridFor(size_t(n), 256), 256, 0, s_>>>(global_.counts.template as<uint32_t>(),
                                                          small_.template as<uint32_t>(), n);
    CSB_LAUNCHED();
    CSB_CUDA(cudaMemcpyAsync(global_.counts.p, small_.p, size_t(n) * 4, cudaMemcpyDeviceToDevice, s_));
    return 0;
}

template<class K, class T>
int Engine<K, T>::updateGlobalTree(const K* keys, size_t n)
{
    // [ref: domain/assignment.hpp:92-98] one update step per sync, iterated to convergence on the first call
    if (globalFresh_)
    {
        // the root count must be the GLOBAL particle count so that every rank takes the same decisions
        CSB_TRY(global_.reset(uint32_t(n), s_));
        CSB_TRY(allreduceCounts());
        globalFresh_ = false;
        bool conv    = false;
        for (int it = 0; it < 64 && !conv; ++it)
        {
            CSB_TRY(global_.update(keys, keys + n, bucket_, conv, s_));
            CSB_TRY(allreduceCounts());
        }
    }
    else
    {
        bool conv = false;
        CSB_TRY(global_.update(keys, keys + n, bucket_, conv, s_));
        CSB_TRY(allreduceCounts());
    }
    int L = global_.numNodes;
    hGlobalLeaves_.resize(size_t(L) + 1);
    hGlobalCounts_.resize(L);
    CSB_CUDA(cudaMemcpyAsync(hGlobalLeaves_.data(), global_.treeA.p, sizeof(K) * (size_t(L) + 1), cudaMemcpyDeviceToHost, s_));
    CSB_CUDA(cudaMemcpyAsync(hGlobalCounts_.data(), global_.counts.p, 4 * size_t(L), cudaMemcpyDeviceToHost, s_));
    CSB_CUDA(cudaStreamSynchronize(s_));
    return 0;
}

//! iterate a local tree to convergence on the given sorted keys (from the root on the first use)
template<class K, class T>
int Engine<K, T>::convergeTree(LeafArrayDev<K, T>& tree, const K* keys, size_t n, bool& fresh, int maxSteadyIter,
                               bool* leavesChanged)
{
    const bool wasFresh = fresh;
    if (fresh)
    {
        CSB_TRY(tree.reset(uint32_t(n), s_));
        fresh = false;
    }
    else
    {
        // counts of the previous tree on the NEW keys steer the first rebalance step
        CSB_TRY((Api<K, T>::nodeCounts(tree.treeA.template as<K>(), tree.counts.template as<uint32_t>(), tree.numNodes, keys,
                                       keys + n, s_)));
    }
    bool conv = false;
    const int iters = wasFresh ? 64 : maxSteadyIter;
    bool changed    = wasFresh;
    for (int it = 0; it < iters && !conv; ++it)
    {
        // steady state, first step: the counts were just taken on these keys with this tree
        CSB_TRY(tree.update(keys, keys + n, bucketFocus_, conv, s_, !wasFresh && it == 0));
        changed = changed || !conv; // converged == every node op was "keep": the leaves are those of the call before
    }
    if (leavesChanged) *leavesChanged = changed;
    return 0;
}

template<class K, class T>
int Engine<K, T>::linkTree(LeafArrayDev<K, T>& tree, typename Engine<K, T>::Linked& ln, bool withCenters, bool leavesChanged)
{
    int L = tree.numNodes, M = L + (L - 1) / 7;
    if (leavesChanged || !ln.valid)
    {
        CSB_TRY(ln.prefixes.reserve(sizeof(K) * size_t(M)));
        CSB_TRY(ln.childOffsets.reserve(4 * (size_t(M) + 1)));
        CSB_TRY(ln.parents.reserve(4 * size_t(std::max(1, (M - 1) / 8))));
        CSB_TRY(ln.levelRange.reserve(4 * 32));
        CSB_TRY(ln.internalToLeaf.reserve(4 * size_t(M)));
        CSB_TRY(ln.leafToInternal.reserve(4 * size_t(M)));
        uint64_t tb = Api<K, T>::octreeTemp(L);
        CSB_TRY(octTemp_.reserve(tb + 512));
        CSB_TRY((Api<K, T>::buildOctree(tree.treeA.template as<K>(), L, ln.prefixes.template as<K>(), ln.childOffsets.template as<int32_t>(),
                                        ln.parents.template as<int32_t>(), ln.levelRange.template as<int32_t>(),
                                        ln.internalToLeaf.template as<int32_t>(), ln.leafToInternal.template as<int32_t>(), octTemp_.p,
                                        octTemp_.bytes, s_)));
        ln.valid = true;
    }
    // layout = exclusive scan of the leaf counts (every leaf of this tree is present)
    CSB_TRY(ln.layout.reserve(4 * (size_t(L) + 1)));
    CSB_TRY(scanTemp_.reserve(cstone_b200_scan_temp_bytes(uint64_t(L) + 1) + 512));
    CSB_TRY(cstone_b200_compute_node_layout(tree.counts.template as<uint32_t>(), nullptr, 0, L, L, ln.layout.template as<uint32_t>(),
                                            scanTemp_.p, s_));
    if (withCenters)
    {
        CSB_TRY(centers_.reserve(sizeof(T) * 3 * size_t(M)));
        CSB_TRY(sizes_.reserve(sizeof(T) * 3 * size_t(M)));
        // the box moves with the particles: centres and sizes are recomputed in every sync
        CSB_TRY((Api<K, T>::geoCenters(sfcKind_, ln.prefixes.template as<K>(), M, centers_.template as<T>(), sizes_.template as<T>(), &box_, s_)));
    }
    return 0;
}


template<class K, class T>
int64_t Engine<K, T>::syncBegin(uint64_t nIn, const T* x, const T* y, const T* z, const T* h, int nprops,
                                const void* const* props, const int* propBytes)
{
    if (nprops < 0 || nprops > kMaxProps || (nIn && (!x || !y || !z || !h))) return CSTONE_E_INVALID_ARGUMENT;
    if (nRanks_ > 1 && !comm_) return CSTONE_E_INVALID_ARGUMENT; // set_comm or set_local_world first
    if (nRanks_ == 1 && haloMode_ == 1 && !comm_)
    {
        // one rank in the reference halo mode runs the multi-rank code path (focus tree and all) over a world of one
        selfWorld_.reset(new LocalWorld(1));
        comm_.reset(new LocalComm(selfWorld_.get(), 0));
    }
    if (nIn >= (uint64_t(1) << 30)) return CSTONE_E_TOO_LARGE;
    uint64_t inStart = 0, inEnd = nIn;
    if (!firstCall_)
    {
        if (nIn != nTotal_) return CSTONE_E_SIZE_MISMATCH; // [ref: domain/domain.hpp:130-133]
        inStart = start_, inEnd = end_;
    }
    for (int a = 0; a < nprops; ++a)
        if ((propBytes[a] != 4 && propBytes[a] != 8 && propBytes[a] != 12 && propBytes[a] != 16) || !props[a])
            return CSTONE_E_INVALID_ARGUMENT; // [ref: domain/buffer_description.hpp:98-110: elements up to 16 bytes]
    const size_t np = size_t(inEnd - inStart);
    const int na    = 4 + nprops;
    int eb[4 + kMaxProps];
    const void* src[4 + kMaxProps];
    eb[0] = eb[1] = eb[2] = eb[3] = int(sizeof(T));
    src[0] = x + inStart, src[1] = y + inStart, src[2] = z + inStart, src[3] = h + inStart;
    nprops_ = nprops;
    for (int a = 0; a < nprops; ++a)
    {
        eb[4 + a]     = propBytes[a];
        propBytes_[a] = propBytes[a];
        src[4 + a]    = reinterpret_cast<const char*>(props[a]) + inStart * size_t(propBytes[a]);
    }

    // 1. global box
    mark(0);
    CSB_TRY(updateBox(x + inStart, y + inStart, z + inStart, np));
    mark(1);

    // 2. keys of the particles this rank holds, sorted, with the permutation [ref: domain/assignment.hpp:84-90]
    auto reserveSort = [&](size_t n) -> int
    {
        CSB_TRY(keyBuf_.reserve(sizeof(K) * n + 256));
        CSB_TRY(order_.reserve(4 * n + 256, true, s_)); // (keeps the map of the first sort: the merge reads it)
        CSB_TRY(valBuf_.reserve(4 * n + 256));
        CSB_TRY(sortTemp_.reserve((Api<K, T>::sortTemp(n)) + 256));
        return 0;
    };
    CSB_TRY(keys_.reserve(sizeof(K) * np + 256));
    CSB_TRY(reserveSort(np));
    K* keys = keys_.template as<K>();
    if (np)
    {
        CSB_CUDA(cudaMemsetAsync(keys, 0, sizeof(K) * np, s_));
        CSB_TRY((Api<K, T>::keys(sfcKind_, x + inStart, y + inStart, z + inStart, keys, np, &box_, s_)));
        mark(2);
        CSB_TRY((Api<K, T>::sort(keys, order_.template as<uint32_t>(), keyBuf_.template as<K>(), valBuf_.template as<uint32_t>(), np,
                                 sortTemp_.p, sortTemp_.bytes, 1, s_)));
    }
    else mark(2);
    mark(3);

    // 3. replicated global tree and the assignment of its leaves to ranks
    CSB_TRY(updateGlobalTree(keys, np));
    makeAssignment<K>(nRanks_, hGlobalLeaves_.data(), hGlobalCounts_.data(), global_.numNodes,
                      oldBoundaries_.empty() ? nullptr : oldBoundaries_.data(), boundaries_, rankCounts_);
    oldBoundaries_ = boundaries_;
    mark(4);

    void* dst[4 + kMaxProps];
    if (!multi())
    {
        // every particle stays: one gather into the final arrays, no halos
        nAssigned_ = np, start_ = 0, end_ = np, nTotal_ = np, nHaloSend_ = 0;
        pat_.valid = true, pat_.steady = false, pat_.inStart = inStart, pat_.np = np, pat_.nAssigned = np, pat_.start = 0;
        for (int a = 0; a < na; ++a)
        {
            CSB_TRY(fin_[a].reserve(size_t(eb[a]) * np + 256));
            dst[a] = fin_[a].p;
        }
        if (fieldsReady_)
        {
            CSB_CUDA(cudaStreamWaitEvent(s_, fieldsReady_, 0));
            fieldsReady_ = nullptr;
        }
        if (np) CSB_TRY(gatherMulti(order_.template as<uint32_t>(), np, na, src, dst, eb));
        CSB_TRY(fkeys_.reserve(sizeof(K) * np + 256));
        CSB_CUDA(cudaMemcpyAsync(fkeys_.p, keys, sizeof(K) * np, cudaMemcpyDeviceToDevice, s_));
        mark(5), mark(6), mark(7), mark(8);
    }
    else
    {
        const int P = nRanks_;
        // 4. send ranges = positions of the rank boundaries in the sorted keys [ref: domain/domaindecomp.hpp:217-230]
        CSB_TRY(small_.reserve(size_t(P + 1) * (sizeof(K) + 4) + size_t(P) * P * 4 + 4096));
        K* dBound      = small_.template as<K>();
        uint32_t* dPos = reinterpret_cast<uint32_t*>(dBound + P + 1);
        CSB_CUDA(cudaMemcpyAsync(dBound, boundaries_.data(), sizeof(K) * (P + 1), cudaMemcpyHostToDevice, s_));
        lowerBoundQueriesKernel<K><<<1, 64, 0, s_>>>(keys, uint32_t(np), dBound, P + 1, dPos);
        CSB_LAUNCHED();
        std::vector<uint32_t> sendStart(P + 1), sendCount(P);
        CSB_CUDA(cudaMemcpyAsync(sendStart.data(), dPos, 4 * (P + 1), cudaMemcpyDeviceToHost, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        sendStart[0] = 0, sendStart[P] = uint32_t(np); // keys below / above the tree range cannot occur
        for (int q = 0; q < P; ++q)
            sendCount[q] = sendStart[q + 1] - sendStart[q];

        // 5. who sends how much to whom
        uint32_t* dMine = dPos + P + 1;
        uint32_t* dAll  = dMine + P;
        CSB_CUDA(cudaMemcpyAsync(dMine, sendCount.data(), 4 * P, cudaMemcpyHostToDevice, s_));
        CSB_TRY(comm_->allGather(dMine, dAll, 4 * size_t(P), s_));
        std::vector<uint32_t> matrix(size_t(P) * P);
        CSB_CUDA(cudaMemcpyAsync(matrix.data(), dAll, 4 * size_t(P) * P, cudaMemcpyDeviceToHost, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        std::vector<uint32_t> recvCount(P), recvOff(P + 1, 0);
        for (int q = 0; q < P; ++q)
        {
            recvCount[q]   = matrix[size_t(q) * P + rank_];
            recvOff[q + 1] = recvOff[q] + recvCount[q];
        }
        nAssigned_ = recvOff[P];
        if (nAssigned_ >= (uint64_t(1) << 30)) return CSTONE_E_TOO_LARGE;
        /* steady state = nearly everything stays here: the particles that stay are neither packed nor copied, the
         * assigned set is later gathered from two sources (caller arrays through the sort map, receive buffers) */
        const bool steady = nAssigned_ && size_t(uint32_t(nAssigned_) - sendCount[rank_]) * 4 <= size_t(nAssigned_);

        pat_.valid = true, pat_.steady = steady, pat_.inStart = inStart, pat_.np = np, pat_.nAssigned = nAssigned_;
        pat_.sendStart = sendStart, pat_.sendCount = sendCount, pat_.recvOff = recvOff, pat_.recvCount = recvCount;
        if (!steady && np)
        {
            // the second sort below reuses order_: keep the map of the first one for reapplySync
            CSB_TRY(order1Keep_.reserve(4 * np + 256));
            CSB_CUDA(cudaMemcpyAsync(order1Keep_.p, order_.p, 4 * np, cudaMemcpyDeviceToDevice, s_));
        }

        // 6. ONE gather orders the outgoing messages (sorted by key = grouped by destination rank)
        for (int a = 0; a < na; ++a)
        {
            CSB_TRY(pack_[a].reserve(size_t(eb[a]) * np + 256));
            dst[a]      = pack_[a].p;
            origSrc_[a] = src[a];
        }
        if (fieldsReady_)
        {
            CSB_CUDA(cudaStreamWaitEvent(s_, fieldsReady_, 0));
            fieldsReady_ = nullptr;
        }
        if (!steady)
        {
            if (np) CSB_TRY(gatherMulti(order_.template as<uint32_t>(), np, na, src, dst, eb));
        }
        else
        {
            // only what leaves: the sorted ranges below and above the part that stays
            const size_t lo = sendStart[rank_], hiStart = size_t(sendStart[rank_]) + sendCount[rank_];
            void* d2[4 + kMaxProps];
            if (lo) CSB_TRY(gatherMulti(order_.template as<uint32_t>(), lo, na, src, dst, eb));
            if (np > hiStart)
            {
                for (int a = 0; a < na; ++a)
                    d2[a] = pack_[a].template as<char>() + hiStart * size_t(eb[a]);
                CSB_TRY(gatherMulti(order_.template as<uint32_t>() + hiStart, np - hiStart, na, src, d2, eb));
            }
        }

        // 7. exchange: grouped NCCL send/recv, received straight into place [ref: domaindecomp_mpi_gpu.cuh:85-185]
        for (int a = 0; a < na; ++a)
            CSB_TRY(recv_[a].reserve(size_t(eb[a]) * nAssigned_ + 256));
        {
            std::vector<CommMsg> sends, recvs;
            for (int q = 0; q < P; ++q)
            {
                if (q == rank_) continue;
                for (int a = 0; a < na; ++a)
                {
                    if (sendCount[q])
                        sends.push_back({q, pack_[a].template as<char>() + size_t(sendStart[q]) * eb[a], size_t(sendCount[q]) * eb[a]});
                    if (recvCount[q])
                        recvs.push_back({q, recv_[a].template as<char>() + size_t(recvOff[q]) * eb[a], size_t(recvCount[q]) * eb[a]});
                }
            }
            CSB_TRY(comm_->exchange(sends, recvs, s_));
        }
        if (!steady)
            for (int a = 0; a < na; ++a)
                if (sendCount[rank_])
                    CSB_CUDA(cudaMemcpyAsync(recv_[a].template as<char>() + size_t(recvOff[rank_]) * eb[a],
                                             pack_[a].template as<char>() + size_t(sendStart[rank_]) * eb[a],
                                             size_t(sendCount[rank_]) * eb[a], cudaMemcpyDeviceToDevice, s_));

        mark(5);
        // 8. keys + order of the assigned set (P sorted runs) [ref: domain/assignment.hpp:141-150]
        const size_t n = size_t(nAssigned_);
        CSB_TRY(skeys_.reserve(sizeof(K) * n + 256));
        CSB_TRY(reserveSort(n));
        K* skeys = skeys_.template as<K>();
        const uint32_t nOwn = sendCount[rank_], nLeft = recvOff[rank_], nF = uint32_t(n) - nOwn;
        map2_ = nullptr;
        if (steady)
        {
            /* steady state: nearly everything stayed here.  The own run is already sorted (keys are those of step 2),
             * so only the foreign arrivals are keyed and sorted, then merged in:  O(n) streaming instead of a second
             * 8-pass sort of all n keys. */
            const K* ownKeys       = keys + sendStart[rank_];
            const uint32_t* ownMap = order_.template as<uint32_t>() + sendStart[rank_]; // caller index of own particle i
            uint32_t* order        = valBuf_.template as<uint32_t>();                   // two-source map (see kFromRecv)
            if (nF == 0)
            {
                CSB_CUDA(cudaMemcpyAsync(skeys, ownKeys, sizeof(K) * n, cudaMemcpyDeviceToDevice, s_));
                map2_ = ownMap;
            }
            else
            {
                const size_t scanBytes = cstone_b200_scan_temp_bytes(uint64_t(nOwn) + 1) + 512;
                auto up = [](size_t b) { return (b + 255) / 256 * 256; };
                const size_t oKb = up(sizeof(K) * nF), oV = oKb + up(sizeof(K) * nF), oVb = oV + up(4 * size_t(nF)),
                             oIp = oVb + up(4 * size_t(nF)), oCnt = oIp + up(4 * size_t(nF)),
                             oScan = oCnt + up(4 * (size_t(nOwn) + 1));
                CSB_TRY(merge_.reserve(oScan + scanBytes));
                char* mb      = merge_.template as<char>();
                K* fk         = reinterpret_cast<K*>(mb);
                K* fkb        = reinterpret_cast<K*>(mb + oKb);
                uint32_t* fv  = reinterpret_cast<uint32_t*>(mb + oV);
                uint32_t* fvb = reinterpret_cast<uint32_t*>(mb + oVb);
                uint32_t* ip  = reinterpret_cast<uint32_t*>(mb + oIp);
                uint32_t* cnt = reinterpret_cast<uint32_t*>(mb + oCnt);
                void* scanTmp = mb + oScan;
                CSB_CUDA(cudaMemsetAsync(fk, 0, sizeof(K) * nF, s_));
                const T *rx = recv_[0].template as<T>(), *ry = recv_[1].template as<T>(), *rz = recv_[2].template as<T>();
                if (nLeft) CSB_TRY((Api<K, T>::keys(sfcKind_, rx, ry, rz, fk, nLeft, &box_, s_)));
                if (nF > nLeft)
                    CSB_TRY((Api<K, T>::keys(sfcKind_, rx + nLeft + nOwn, ry + nLeft + nOwn, rz + nLeft + nOwn, fk + nLeft, nF - nLeft, &box_, s_)));
                foreignIndexKernel<<<gridFor(nF, 256), 256, 0, s_>>>(fv, nF, nLeft, nOwn);
                CSB_LAUNCHED();
                CSB_TRY((Api<K, T>::sort(fk, fv, fkb, fvb, nF, sortTemp_.p, sortTemp_.bytes, 0, s_)));
                CSB_CUDA(cudaMemsetAsync(cnt, 0, 4 * (size_t(nOwn) + 1), s_));
                insertionPointsKernel<K><<<gridFor(nF, 256), 256, 0, s_>>>(ownKeys, nOwn, fk, nF, ip, cnt);
                CSB_LAUNCHED();
                CSB_TRY(cstone_b200_exclusive_scan_u32(cnt, cnt, size_t(nOwn) + 1, 0u, scanTmp, s_));
                mergeScatterKernel<K><<<gridFor(n, 256), 256, 0, s_>>>(ownKeys, nOwn, ownMap, cnt, fk, fv, ip, nF, skeys, order);
                CSB_LAUNCHED();
                map2_ = order;
            }
        }
        else if (n)
        {
            CSB_CUDA(cudaMemsetAsync(skeys, 0, sizeof(K) * n, s_));
            CSB_TRY((Api<K, T>::keys(sfcKind_, recv_[0].template as<T>(), recv_[1].template as<T>(), recv_[2].template as<T>(), skeys, n, &box_, s_)));
            CSB_TRY((Api<K, T>::sort(skeys, order_.template as<uint32_t>(), keyBuf_.template as<K>(), valBuf_.template as<uint32_t>(), n,
                                     sortTemp_.p, sortTemp_.bytes, 1, s_)));
        }
        // h in SFC order for the halo search radii
        CSB_TRY(shAssigned_.reserve(sizeof(T) * n + 256));
        {
            const int which[1] = {3};
            void* d1[1]        = {shAssigned_.p};
            CSB_TRY(gatherAssigned(n, 1, which, d1, eb));
        }

        if (grav_ && n)
        {
            // [ref: domain/domain.hpp:262-264] x, y, z, m of the assigned particles in SFC order for the expansion centres
            DevBuf* gb[4]      = {&sgx_, &sgy_, &sgz_, &sgm_};
            const int which[4] = {0, 1, 2, 4};
            void* d4[4];
            for (int k = 0; k < 4; ++k)
            {
                CSB_TRY(gb[k]->reserve(sizeof(T) * n + 256));
                d4[k] = gb[k]->p;
            }
            CSB_TRY(gatherAssigned(n, 4, which, d4, eb));
        }
        mark(6);
        // 9. halos
        CSB_TRY(haloMode_ == 1 ? discoverHalosLet() : discoverHalos());
        mark(8);
    }

    // 10. search tree over assigned + halo particles, linked, with node centres and layout
    if (haloMode_ == 1)
    {
        // the LET is the search tree [ref: domain/domain.hpp:423-436 octreeProperties]; layout: present leaves only
        mark(9);
        CSB_TRY(linkTree(search_, lnS_, true, true));
        CSB_CUDA(cudaMemcpyAsync(lnS_.layout.p, letLayoutDev_.p, 4 * hLayout_.size(), cudaMemcpyDeviceToDevice, s_));
        mark(10);
    }
    else
    {
        bool searchChanged = true;
        CSB_TRY(convergeTree(search_, fkeys_.template as<K>(), size_t(nTotal_), searchFresh_, 64, &searchChanged));
        mark(9);
        CSB_TRY(linkTree(search_, lnS_, true, searchChanged));
        mark(10);
    }
    CSB_CUDA(cudaStreamSynchronize(s_));
    firstCall_ = false;
    if (win_)
    {
        int32_t st = 0;
        CSB_TRY(pushStatus(&st));
        if (st) return CSTONE_E_PEER_TIMEOUT;
    }
    return int64_t(nTotal_);
}


template<class K, class T>
int Engine<K, T>::discoverHalos()
{
    const int P     = nRanks_;
    const size_t n  = size_t(nAssigned_);
    const K* skeys  = skeys_.template as<K>();
    const K lowest  = boundaries_[rank_];
    const K highest = boundaries_[rank_ + 1];
    const int na    = 4 + nprops_;
    int eb[4 + kMaxProps];
    for (int a = 0; a < na; ++a)
        eb[a] = a < 4 ? int(sizeof(T)) : propBytes_[a - 4];

    // optional sub-stage timing (CSB_HALO_TIMING=1): one line per sync on stderr
    static const bool lapOn = std::getenv("CSB_HALO_TIMING") != nullptr;
    cudaEvent_t lapEv[8];
    int nLap = 0;
    auto lap = [&]()
    {
        if (!lapOn || nLap >= 8) return;
        cudaEventCreate(&lapEv[nLap]);
        cudaEventRecord(lapEv[nLap++], s_);
    };
    lap();
    /* a. tree over the assigned particles, linked (for the reverse collision) and with its particle layout.  Halo
     * discovery needs A valid leaf array with exact counts, not THE converged one: after the first sync the tree of the
     * previous step gets one rebalance + recount step per sync (it follows the particles one step behind) instead of
     * a convergence loop, which cost the ranks whose tree changed a second iteration and everybody else the wait. */
    bool focusChanged = true;
    CSB_TRY(convergeTree(focus_, skeys, n, focusFresh_, 1, &focusChanged));
    CSB_TRY(linkTree(focus_, lnF_, false, focusChanged));
    const int L = focus_.numNodes;

    lap();
    // b. search radius per leaf [ref: halos/halos.hpp:148-160]
    CSB_TRY(radii_.reserve(4 * size_t(L) + 256));
    CSB_TRY((Api<K, T>::haloRadii(shAssigned_.template as<T>(), lnF_.layout.template as<uint32_t>(), size_t(L), 2.0f * haloFactor_,
                                  radii_.template as<float>(), s_)));

    // c. halo boxes of my surface leaves
    SfcTables tabs;
    CSB_TRY(getSfcTables(&tabs));
    CSB_TRY(surf_.reserve(sizeof(IBox6) * size_t(L) + 256));
    CSB_TRY(nSurf_.reserve(4 * size_t(2 * P + 2) + size_t(P) * P * 4 + 256));
    uint32_t* dCount = nSurf_.template as<uint32_t>();      // [0] my count, [1..P] all counts
    CSB_CUDA(cudaMemsetAsync(dCount, 0, 4, s_));
    surfaceBoxesKernel<K, T><<<gridFor(size_t(L), 128), 128, 0, s_>>>(
        sfcKind_,         tabs.enc, tabs.dec, focus_.treeA.template as<K>(), focus_.counts.template as<uint32_t>(), radii_.template as<float>(), L, lowest,
        highest, makeDBox<T>(&box_), surf_.template as<IBox6>(), uint32_t(L), dCount);
    CSB_LAUNCHED();
    CSB_TRY(comm_->allGather(dCount, dCount + 1, 4, s_));
    std::vector<uint32_t> nS(P);
    CSB_CUDA(cudaMemcpyAsync(nS.data(), dCount + 1, 4 * P, cudaMemcpyDeviceToHost, s_));
    CSB_CUDA(cudaStreamSynchronize(s_));
    uint32_t maxS = 1;
    for (int q = 0; q < P; ++q)
        maxS = std::max(maxS, nS[q]);

    lap();
    // d. everybody gets everybody's surface boxes (tens of MB at most: surface ~ N^(2/3))
    CSB_TRY(surf_.reserve(sizeof(IBox6) * size_t(maxS) + 256, true, s_));
    CSB_TRY(allSurf_.reserve(sizeof(IBox6) * size_t(maxS) * P + 256));
    CSB_TRY(comm_->allGather(surf_.p, allSurf_.p, sizeof(IBox6) * size_t(maxS), s_));

    lap();
    // e. reverse collision against my own leaves
    CSB_TRY(flags_.reserve(size_t(P) * L + 256));
    CSB_CUDA(cudaMemsetAsync(flags_.p, 0, size_t(P) * L, s_));
    {
        dim3 grid(gridFor(size_t(maxS), 128, 4), P);
        reverseCollisionKernel<K><<<grid, 128, 0, s_>>>(sfcKind_, tabs.dec, lnF_.prefixes.template as<K>(), lnF_.childOffsets.template as<int32_t>(),
                                                       lnF_.internalToLeaf.template as<int32_t>(), focus_.counts.template as<uint32_t>(), L,
                                                       lowest, highest, allSurf_.template as<IBox6>(), dCount + 1, maxS, P, rank_,
                                                       flags_.template as<unsigned char>());
        CSB_LAUNCHED();
    }

    lap();
    // f. send lists: per peer the particles of my flagged leaves, in leaf (= SFC) order.  One scan and one expansion
    //    over all (peer, leaf) pairs; the per-peer offsets are the scan values at the start of each peer's row.
    const size_t PL = size_t(P) * L;
    CSB_TRY(sendScan_.reserve(4 * (PL + 1) + 256));
    CSB_TRY(scanTemp_.reserve(cstone_b200_scan_temp_bytes(uint64_t(PL) + 1) + 512));
    haloSendCount_.assign(P, 0), haloSendOffset_.assign(P + 1, 0);
    {
        LoadFlaggedCount load{flags_.template as<unsigned char>(), focus_.counts.template as<uint32_t>(), size_t(L), PL};
        uint32_t* scan = sendScan_.template as<uint32_t>();
        CSB_TRY((launchScan<false>(load, scan, PL + 1, 0u, scanTemp_.p, nullptr, s_)));
        CSB_CUDA(cudaMemcpy2DAsync(haloSendOffset_.data(), 4, scan, 4 * size_t(L), 4, size_t(P) + 1, cudaMemcpyDeviceToHost, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
    }
    for (int q = 0; q < P; ++q)
        haloSendCount_[q] = haloSendOffset_[q + 1] - haloSendOffset_[q];
    nHaloSend_ = haloSendOffset_[P];
    CSB_TRY(sendIdx_.reserve(4 * size_t(nHaloSend_) + 256));
    if (nHaloSend_)
    {
        expandSendListKernel<<<gridFor(PL, 256), 256, 0, s_>>>(flags_.template as<unsigned char>(), lnF_.layout.template as<uint32_t>(),
                                                              sendScan_.template as<uint32_t>(), size_t(L), PL,
                                                              sendIdx_.template as<uint32_t>());
        CSB_LAUNCHED();
    }

    lap();
    // g. receive counts, final layout: [halos of ranks < me][assigned][halos of ranks > me], all in SFC order
    uint32_t* dMine = dCount + 1 + P;
    uint32_t* dAll  = dMine + P;
    CSB_CUDA(cudaMemcpyAsync(dMine, haloSendCount_.data(), 4 * P, cudaMemcpyHostToDevice, s_));
    CSB_TRY(comm_->allGather(dMine, dAll, 4 * size_t(P), s_));
    std::vector<uint32_t> matrix(size_t(P) * P);
    CSB_CUDA(cudaMemcpyAsync(matrix.data(), dAll, 4 * size_t(P) * P, cudaMemcpyDeviceToHost, s_));
    CSB_CUDA(cudaStreamSynchronize(s_));
    haloMatrix_ = matrix;
    for (int q = 0; q < P; ++q)
        haloMatrix_[size_t(q) * P + q] = 0;
    haloRecvCount_.assign(P, 0), haloRecvOffset_.assign(P, 0);
    uint64_t left = 0, right = 0;
    for (int q = 0; q < P; ++q)
    {
        haloRecvCount_[q] = q == rank_ ? 0u : matrix[size_t(q) * P + rank_];
        if (q < rank_) left += haloRecvCount_[q];
        if (q > rank_) right += haloRecvCount_[q];
    }
    start_  = left;
    pat_.start = left;
    end_    = left + n;
    nTotal_ = end_ + right;
    if (nTotal_ >= (uint64_t(1) << 30)) return CSTONE_E_TOO_LARGE;
    {
        uint64_t l = 0, r = end_;
        for (int q = 0; q < P; ++q)
        {
            if (q < rank_) haloRecvOffset_[q] = uint32_t(l), l += haloRecvCount_[q];
            if (q > rank_) haloRecvOffset_[q] = uint32_t(r), r += haloRecvCount_[q];
        }
    }

    lap();
    if (lapOn && nLap != 7)
        for (int k = 0; k < nLap; ++k)
            cudaEventDestroy(lapEv[k]);
    if (lapOn && nLap == 7)
    {
        cudaEventSynchronize(lapEv[6]);
        float ms[6];
        for (int k = 0; k < 6; ++k)
            cudaEventElapsedTime(&ms[k], lapEv[k], lapEv[k + 1]);
        std::fprintf(stderr, "[csb rank %d] halo discovery ms: tree %.3f radii+surface %.3f allgather %.3f collide %.3f lists %.3f counts %.3f "
                             "(L %d, surface max %u)\n", rank_, ms[0], ms[1], ms[2], ms[3], ms[4], ms[5], L, maxS);
        for (int k = 0; k < 7; ++k)
            cudaEventDestroy(lapEv[k]);
    }
    // h. assigned particles straight to their final position, then the halos of x,y,z,h
    mark(7);
    void* dst[4 + kMaxProps];
    int which[4 + kMaxProps];
    for (int a = 0; a < na; ++a)
    {
        CSB_TRY(fin_[a].reserve(size_t(eb[a]) * nTotal_ + 256));
        which[a] = a;
        dst[a]   = fin_[a].template as<char>() + size_t(start_) * eb[a];
    }
    CSB_TRY(gatherAssigned(n, na, which, dst, eb));
    void* xyzh[4] = {fin_[0].p, fin_[1].p, fin_[2].p, fin_[3].p};
    CSB_TRY(exchangeHalos(4, xyzh, eb));

    // i. keys: assigned ones are known, halo keys are computed [ref: domain/domain.hpp:530-532]
    CSB_TRY(fkeys_.reserve(sizeof(K) * nTotal_ + 256));
    K* fk = fkeys_.template as<K>();
    CSB_CUDA(cudaMemsetAsync(fk, 0, sizeof(K) * nTotal_, s_));
    CSB_CUDA(cudaMemcpyAsync(fk + start_, skeys, sizeof(K) * n, cudaMemcpyDeviceToDevice, s_));
    const T *fx = fin_[0].template as<T>(), *fy = fin_[1].template as<T>(), *fz = fin_[2].template as<T>();
    if (start_) CSB_TRY((Api<K, T>::keys(sfcKind_, fx, fy, fz, fk, size_t(start_), &box_, s_)));
    if (nTotal_ > end_)
        CSB_TRY((Api<K, T>::keys(sfcKind_, fx + end_, fy + end_, fz + end_, fk + end_, size_t(nTotal_ - end_), &box_, s_)));
    return 0;
}

/*! halo discovery the reference's way [ref: domain/domain.hpp:217-243]: peers from the global tree, the multi-rank focus
 *  tree brought up to date, findHalos on it, layout of the present leaves, key-range requests between peers. */
template<class K, class T>
int Engine<K, T>::discoverHalosLet()
{
    const int P    = nRanks_;
    const size_t n = size_t(nAssigned_);
    const K* skeys = skeys_.template as<K>();
    const int na   = 4 + nprops_;
    int eb[4 + kMaxProps];
    for (int a = 0; a < na; ++a)
        eb[a] = a < 4 ? int(sizeof(T)) : propBytes_[a - 4];
    if (!let_) let_.reset(new LetTree<K, T>(rank_, P, bucketFocus_, sfcKind_));
    LetTree<K, T>& let = *let_;

    // optional section timing (CSB_HALO_TIMING=1): wall time between laps, stream drained at each lap, rank 0 prints
    static const bool lapOn = std::getenv("CSB_HALO_TIMING") != nullptr;
    std::string lapLog;
    auto lapT = std::chrono::steady_clock::now();
    auto lap = [&](const char* name)
    {
        if (!lapOn) return;
        cudaStreamSynchronize(s_);
        auto now = std::chrono::steady_clock::now();
        char buf[64];
        std::snprintf(buf, sizeof(buf), " %s %.2f", name, std::chrono::duration<double, std::milli>(now - lapT).count());
        lapLog += buf;
        lapT = now;
    };
    // a. peers [ref: traversal/peers.hpp:62-117], focus tree [ref: domain/domain.hpp:219-228]
    HostOctree<K> gtree;
    gtree.build(hGlobalLeaves_.data(), global_.numNodes);
    const bool grav = grav_;
    // [ref: traversal/macs.hpp:45-48 invThetaMinMac / invThetaVecMac]
    const float invThetaEff = grav ? 1.0f / theta_ + std::sqrt(3.0f) : 1.0f / theta_ + 0.5f;
    const std::vector<int> peers = findPeersMacHost<K, T>(sfcKind_, rank_, boundaries_, gtree, box_, invThetaEff);
    lap("peers");
    CSB_TRY(small_.reserve(size_t(P + 1) * (sizeof(K) + 4) + size_t(P) * P * 4 + 4096));
    auto allSum = [&](int& v) -> int
    {
        int32_t* dv = small_.template as<int32_t>();
        CSB_CUDA(cudaMemcpyAsync(dv, &v, 4, cudaMemcpyHostToDevice, s_));
        CSB_TRY(comm_->allReduce(dv, 1, CommType::i32, CommOp::sum, s_));
        CSB_CUDA(cudaMemcpyAsync(&v, dv, 4, cudaMemcpyDeviceToHost, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        return 0;
    };
    auto treeAndCounts = [&](int* converged) -> int
    {
        CSB_TRY(let.updateTree(comm_.get(), peers, boundaries_, box_, converged, s_));
        return let.updateCounts(comm_.get(), skeys, n, hGlobalLeaves_, hGlobalCounts_, s_);
    };
    auto centers = [&]() -> int
    {
        return let.updateCenters(comm_.get(), skeys, n, sgx_.template as<T>(), sgy_.template as<T>(), sgz_.template as<T>(),
                                 sgm_.template as<T>(), hGlobalLeaves_, s_);
    };
    /* The reference converges the tree on the first call only; its syncGrav after plain syncs would read expansion
     * centres of an older tree (undefined there).  Here a syncGrav that does not follow a syncGrav converges again. */
    if (firstCall_ || (grav && !gravReady_))
    {
        // [ref: focus/octree_focus_mpi.hpp:528-547 converge] with the min-distance MAC
        const float invThetaConv = grav ? 1.0f : invThetaEff;
        for (int converged = 0, it = 0; converged != P; ++it)
        {
            if (it > 200) return CSTONE_E_LET;
            CSB_TRY(let.updateMinMac(boundaries_, invThetaConv, s_));
            CSB_TRY(treeAndCounts(&converged));
            CSB_TRY(allSum(converged));
        }
        if (grav)
        {
            // [ref: domain/domain.hpp:272-286] then with the vector MAC around the expansion centres, twice at least
            CSB_TRY(let.updateMinMac(boundaries_, 1.0f, s_));
            for (int converged = 0, reps = 0; converged != P || reps < 2; ++reps)
            {
                if (reps > 200) return CSTONE_E_LET;
                CSB_TRY(treeAndCounts(&converged));
                CSB_TRY(centers());
                CSB_TRY(let.updateMacs(boundaries_, float(1.0 / theta_), s_));
                CSB_TRY(allSum(converged));
            }
        }
    }
    int haloFail = 0;
    bool consistent = true;
    do
    {
    haloFail = 0;
    if (grav)
    {
        // [ref: domain/domain.hpp:289-296]
        CSB_TRY(let.updateMacs(boundaries_, centerDriftTol_ / theta_, s_));
        CSB_TRY(treeAndCounts(nullptr));
        CSB_TRY(centers());
        CSB_TRY(let.updateMacs(boundaries_, float(1.0 / theta_), s_));
    }
    else
    {
        CSB_TRY(let.updateMinMac(boundaries_, invThetaEff, s_));
        lap("minMac");
        CSB_TRY(let.updateTree(comm_.get(), peers, boundaries_, box_, nullptr, s_));
        lap("updateTree");
        CSB_TRY(let.updateCounts(comm_.get(), skeys, n, hGlobalLeaves_, hGlobalCounts_, s_));
        lap("updateCounts");
    }

    const FocusTree<K>& ft           = let.t;
    const int L                      = ft.numLeaves_;
    const PinnedVec<K>& leaves     = let.leaves();
    const PinnedVec<uint32_t>& cnt = let.leafCounts();
    const auto& assign               = let.assignment();
    const int firstNode = assign[rank_][0], lastNode = assign[rank_][1], numSearch = lastNode - firstNode;

    // b. search radius of my leaves, collisions with the rest of the tree [ref: halos/halos.hpp:127-189 discover]
    CSB_TRY(radii_.reserve(4 * size_t(L) + 256));
    CSB_TRY(flags_.reserve(4 * size_t(L) + 256));
    CSB_TRY(sendScan_.reserve(4 * (size_t(numSearch) + 2) + 256));
    CSB_TRY(scanTemp_.reserve(cstone_b200_scan_temp_bytes(uint64_t(L) + 2) + 512));
    CSB_CUDA(cudaMemsetAsync(radii_.p, 0, 4 * size_t(L), s_));
    CSB_CUDA(cudaMemsetAsync(flags_.p, 0, 4 * size_t(L), s_));
    if (numSearch > 0)
    {
        uint32_t* seg = sendScan_.template as<uint32_t>();
        CSB_TRY(cstone_b200_exclusive_scan_u32(ft.leafCounts_.template as<uint32_t>() + firstNode, seg, size_t(numSearch) + 1, 0u,
                                               scanTemp_.p, s_));
        CSB_TRY((Api<K, T>::haloRadii(shAssigned_.template as<T>(), seg, size_t(numSearch), 2.0f * haloFactor_,
                                      radii_.template as<float>() + firstNode, s_)));
        CSB_TRY((Api<K, T>::findHalos(sfcKind_, ft.prefixes_.template as<K>(), ft.childOffsets_.template as<int32_t>(),
                                      ft.i2l_.template as<int32_t>(), ft.leaves_.template as<K>(), radii_.template as<float>(), &box_,
                                      firstNode, lastNode, flags_.template as<int32_t>(), s_)));
    }
    if (grav)
    {
        addMacsKernel<<<gridFor(size_t(L), 256), 256, 0, s_>>>(ft.macs_.template as<char>(), ft.l2i_.template as<int32_t>() + ft.numInternal_,
                                                              L, flags_.template as<int32_t>());
        CSB_LAUNCHED();
    }
    hFlags_.resize(size_t(L));
    CSB_CUDA(cudaMemcpyAsync(hFlags_.data(), flags_.p, 4 * size_t(L), cudaMemcpyDeviceToHost, s_));
    CSB_CUDA(cudaStreamSynchronize(s_));

    lap("findHalos");
    // c. layout of the present leaves [ref: domain/layout.hpp:150-164 computeNodeLayout]
    hLayout_.resize(size_t(L) + 1);
    CSB_TRY(letLayoutDev_.reserve(4 * (size_t(L) + 1) + 256));
    CSB_TRY(cstone_b200_compute_node_layout(ft.leafCounts_.template as<uint32_t>(), flags_.template as<int32_t>(), firstNode, lastNode, L,
                                            letLayoutDev_.template as<uint32_t>(), scanTemp_.p, s_));
    CSB_CUDA(cudaMemcpyAsync(hLayout_.data(), letLayoutDev_.p, 4 * (size_t(L) + 1), cudaMemcpyDeviceToHost, s_));
    CSB_CUDA(cudaStreamSynchronize(s_));
    start_ = hLayout_[firstNode], end_ = hLayout_[lastNode], nTotal_ = hLayout_[L];
    pat_.start = start_;
    const bool tooLarge = nTotal_ >= (uint64_t(1) << 30); // reported with the other flags in step f, on all ranks

    // d. per peer: the key ranges of its leaves that I flagged; it answers with particle ranges of ITS layout
    //    [ref: domain/exchange_keys.hpp:62-118 exchangeRequestKeys, domain/layout.hpp:93-123 extractMarkedElements]
    std::vector<std::vector<char>> requests(P), requested;
    for (int peer : peers)
    {
        std::vector<K> req;
        int i = assign[peer][0];
        const int e = assign[peer][1];
        while (i != e)
        {
            while (i < e && hFlags_[i] == 0)
                ++i;
            if (i != e)
            {
                req.push_back(leaves[i]);
                while (i < e && hFlags_[i] == 1)
                    ++i;
                req.push_back(leaves[i]);
            }
        }
        requests[peer].resize(req.size() * sizeof(K));
        if (!req.empty()) std::memcpy(requests[peer].data(), req.data(), requests[peer].size());
    }
    CSB_TRY(let.exchangeBytes(comm_.get(), requests, requested, s_));
    haloSendCount_.assign(P, 0), haloSendOffset_.assign(P + 1, 0);
    std::vector<uint32_t> rangeStart, rangeOffset;
    uint32_t numSend = 0;
    consistent = uint64_t(end_ - start_) == uint64_t(n) && !tooLarge;
    for (int q = 0; q < P; ++q)
    {
        haloSendOffset_[q] = numSend;
        const size_t nk    = requested[q].size() / sizeof(K);
        const K* rk        = reinterpret_cast<const K*>(requested[q].data());
        int pos            = 0;
        for (size_t k = 0; k + 1 < nk; k += 2)
        {
            // (the requested key ranges of one peer ascend)
            pos               = findNodeAboveFrom(leaves.data(), L + 1, pos, rk[k]);
            const uint32_t lo = hLayout_[pos];
            pos               = findNodeAboveFrom(leaves.data(), L + 1, pos, rk[k + 1]);
            const uint32_t hi = hLayout_[pos];
            // [ref: halos/halos.hpp:40-55 checkIndices] only assigned particles travel as halos
            if (lo < start_ || hi > end_ || hi < lo)
            {
                consistent = false;
                continue;
            }
            if (hi == lo) continue;
            rangeStart.push_back(lo - uint32_t(start_));
            rangeOffset.push_back(numSend);
            numSend += hi - lo;
        }
        haloSendCount_[q] = numSend - haloSendOffset_[q];
    }
    rangeOffset.push_back(numSend);
    haloSendOffset_[P] = numSend;
    nHaloSend_         = numSend;
    CSB_TRY(sendIdx_.reserve(4 * size_t(numSend) + 256));
    if (numSend)
    {
        // the index list of the ranges is written on the device
        const size_t nr = rangeStart.size();
        CSB_TRY(sendScan_.reserve(4 * (2 * nr + 2) + 512));
        uint32_t* dStart  = sendScan_.template as<uint32_t>();
        uint32_t* dOffset = dStart + nr;
        CSB_CUDA(cudaMemcpyAsync(dStart, rangeStart.data(), 4 * nr, cudaMemcpyHostToDevice, s_));
        CSB_CUDA(cudaMemcpyAsync(dOffset, rangeOffset.data(), 4 * (nr + 1), cudaMemcpyHostToDevice, s_));
        expandRangesKernel<<<gridFor(size_t(numSend), 256), 256, 0, s_>>>(dStart, dOffset, int(nr), numSend, sendIdx_.template as<uint32_t>());
        CSB_LAUNCHED();
    }

    // e. where the halos of each peer arrive [ref: domain/layout.hpp:166-183 computeHaloRecvList]
    haloRecvCount_.assign(P, 0), haloRecvOffset_.assign(P, 0);
    for (int peer : peers)
    {
        int firstNz = assign[peer][0], lastNz = assign[peer][1];
        while (firstNz < assign[peer][1] && hFlags_[firstNz] != 1)
            ++firstNz;
        while (lastNz > firstNz && hFlags_[lastNz - 1] != 1)
            --lastNz;
        haloRecvOffset_[peer] = hLayout_[firstNz];
        haloRecvCount_[peer]  = hLayout_[lastNz] - hLayout_[firstNz];
    }
    // [ref: halos/halos.hpp:58-100 checkHalos] every flagged leaf outside of my range must belong to a peer
    for (uint32_t i : let.nonPeerLeaves()) // (the leaves outside of my range and of every peer's range)
        if (hFlags_[i]) haloFail = 1;
    if (grav)
    {
        // [ref: domain/domain.hpp:309-317] a halo leaf outside of every peer: the tree lags the moving centres
        CSB_TRY(allSum(haloFail));
        if (haloFail) centerDriftTol_ += 0.05f;
        if (centerDriftTol_ > 10.f) return CSTONE_E_LET;
    }
    else if (haloFail) consistent = false;
    } while (grav && haloFail);

    lap("layout+requests");
    /* f. every rank sees every send and expected receive count: a mismatch (counts of a peer's range that differ from
     *    the owner's) ends the sync on ALL ranks with an error instead of a hanging exchange */
    {
        CSB_TRY(nSurf_.reserve(size_t(2 * P + 1) * 4 * size_t(P + 1) + 1024));
        uint32_t* dMine = nSurf_.template as<uint32_t>();
        uint32_t* dAll  = dMine + (2 * P + 1 + 63) / 64 * 64;
        std::vector<uint32_t> mine(2 * P + 1), all(size_t(2 * P + 1) * P);
        for (int q = 0; q < P; ++q)
            mine[q] = haloSendCount_[q], mine[P + q] = haloRecvCount_[q];
        const bool letError = let.takeError();
        mine[2 * P]         = (consistent && !letError) ? 0u : 1u;
        if (mine[2 * P] && std::getenv("CSB_TRACE_ERRORS"))
            std::fprintf(stderr, "[cstone_b200] rank %d: inconsistent halo bookkeeping: assigned %llu, layout range %llu, halo leaf "
                                 "outside of all peers %d, total %llu, tree step failed %d\n", rank_, (unsigned long long)n,
                         (unsigned long long)(end_ - start_), haloFail, (unsigned long long)nTotal_, int(letError));
        CSB_CUDA(cudaMemcpyAsync(dMine, mine.data(), 4 * mine.size(), cudaMemcpyHostToDevice, s_));
        CSB_TRY(comm_->allGather(dMine, dAll, 4 * mine.size(), s_));
        CSB_CUDA(cudaMemcpyAsync(all.data(), dAll, 4 * all.size(), cudaMemcpyDeviceToHost, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        bool ok = true;
        for (int a = 0; a < P; ++a)
        {
            ok = ok && all[size_t(a) * (2 * P + 1) + 2 * P] == 0;
            for (int b = 0; b < P; ++b)
            {
                const bool same = all[size_t(a) * (2 * P + 1) + b] == all[size_t(b) * (2 * P + 1) + P + a];
                if (!same && rank_ == 0 && std::getenv("CSB_TRACE_ERRORS"))
                    std::fprintf(stderr, "[cstone_b200] halo counts disagree: rank %d sends %u to rank %d, which expects %u\n", a,
                                 all[size_t(a) * (2 * P + 1) + b], b, all[size_t(b) * (2 * P + 1) + P + a]);
                ok = ok && same;
            }
        }
        if (!ok) return CSTONE_E_LET;
        haloMatrix_.assign(size_t(P) * P, 0);
        for (int a = 0; a < P; ++a)
            for (int b = 0; b < P; ++b)
                haloMatrix_[size_t(a) * P + b] = all[size_t(a) * (2 * P + 1) + b];
    }

    lap("agree");
    // g. assigned particles to their final position, halos of x,y,z,h, keys of the halos
    mark(7);
    void* dst[4 + kMaxProps];
    int which[4 + kMaxProps];
    for (int a = 0; a < na; ++a)
    {
        CSB_TRY(fin_[a].reserve(size_t(eb[a]) * nTotal_ + 256));
        which[a] = a;
        dst[a]   = fin_[a].template as<char>() + size_t(start_) * eb[a];
    }
    CSB_TRY(gatherAssigned(n, na, which, dst, eb));
    void* xyzh[4] = {fin_[0].p, fin_[1].p, fin_[2].p, fin_[3].p};
    CSB_TRY(exchangeHalos(4, xyzh, eb));
    CSB_TRY(fkeys_.reserve(sizeof(K) * nTotal_ + 256));
    K* fk = fkeys_.template as<K>();
    CSB_CUDA(cudaMemsetAsync(fk, 0, sizeof(K) * nTotal_, s_));
    CSB_CUDA(cudaMemcpyAsync(fk + start_, skeys, sizeof(K) * n, cudaMemcpyDeviceToDevice, s_));
    const T *fx = fin_[0].template as<T>(), *fy = fin_[1].template as<T>(), *fz = fin_[2].template as<T>();
    if (start_) CSB_TRY((Api<K, T>::keys(sfcKind_, fx, fy, fz, fk, size_t(start_), &box_, s_)));
    if (nTotal_ > end_)
        CSB_TRY((Api<K, T>::keys(sfcKind_, fx + end_, fy + end_, fz + end_, fk + end_, size_t(nTotal_ - end_), &box_, s_)));

    lap("gather+halos");
    if (lapOn && rank_ == 0) std::fprintf(stderr, "[csb rank 0] reference halo mode ms:%s\n", lapLog.c_str());
    gravReady_ = grav;
    // h. the LET becomes the search tree (leaves + counts; linked in syncBegin)
    const FocusTree<K>& ft = let.t;
    const int L            = ft.numLeaves_;
    CSB_TRY(search_.ensure(L, s_));
    CSB_CUDA(cudaMemcpyAsync(search_.treeA.p, ft.leaves_.p, sizeof(K) * (size_t(L) + 1), cudaMemcpyDeviceToDevice, s_));
    CSB_CUDA(cudaMemcpyAsync(search_.counts.p, ft.leafCounts_.p, 4 * size_t(L), cudaMemcpyDeviceToDevice, s_));
    search_.numNodes = L;
    return 0;
}

template<class K, class T>
int Engine<K, T>::exchangeHalos(int narrays, void* const* arrays, const int* elemBytes)
{
    if (nRanks_ == 1 || narrays == 0) return 0;
    if (narrays < 0 || !arrays || !elemBytes) return CSTONE_E_INVALID_ARGUMENT;
    const int P = nRanks_;
    size_t total = 0;
    std::vector<size_t> off(narrays);
    for (int a = 0; a < narrays; ++a)
    {
        if ((elemBytes[a] != 4 && elemBytes[a] != 8 && elemBytes[a] != 12 && elemBytes[a] != 16) || !arrays[a])
            return CSTONE_E_INVALID_ARGUMENT;
        off[a] = total;
        total += (size_t(nHaloSend_) * elemBytes[a] + 255) / 256 * 256;
    }
    if (haloPush_ < 0)
    {
        const char* env = std::getenv("CSB_HALO_PUSH");
        haloPush_       = (env && env[0] == '0') ? 0 : 1;
    }
    if (haloPush_ == 1 && P <= kMaxPushRanks && narrays <= kMaxPushArrays && haloMatrix_.size() == size_t(P) * P)
        return exchangeHalosPush(narrays, arrays, elemBytes);
    CSB_TRY(haloPack_.reserve(total + 256));
    // pack [ref: halos/gather_halos_gpu.cu:25-75 gatherRanges]
    if (nHaloSend_)
        for (int a = 0; a < narrays; ++a)
        {
            unsigned grid = gridFor(size_t(nHaloSend_), 256);
            if (elemBytes[a] == 16)
                packKernel<uint4><<<grid, 256, 0, s_>>>(sendIdx_.template as<uint32_t>(), uint32_t(nHaloSend_), uint32_t(start_),
                                                       reinterpret_cast<const uint4*>(arrays[a]),
                                                       reinterpret_cast<uint4*>(haloPack_.template as<char>() + off[a]));
            else if (elemBytes[a] == 12)
                packKernel<Elem12><<<grid, 256, 0, s_>>>(sendIdx_.template as<uint32_t>(), uint32_t(nHaloSend_), uint32_t(start_),
                                                        reinterpret_cast<const Elem12*>(arrays[a]),
                                                        reinterpret_cast<Elem12*>(haloPack_.template as<char>() + off[a]));
            else if (elemBytes[a] == 8)
                packKernel<uint64_t><<<grid, 256, 0, s_>>>(sendIdx_.template as<uint32_t>(), uint32_t(nHaloSend_), uint32_t(start_),
                                                          reinterpret_cast<const uint64_t*>(arrays[a]),
                                                          reinterpret_cast<uint64_t*>(haloPack_.template as<char>() + off[a]));
            else
                packKernel<uint32_t><<<grid, 256, 0, s_>>>(sendIdx_.template as<uint32_t>(), uint32_t(nHaloSend_), uint32_t(start_),
                                                          reinterpret_cast<const uint32_t*>(arrays[a]),
                                                          reinterpret_cast<uint32_t*>(haloPack_.template as<char>() + off[a]));
            CSB_LAUNCHED();
        }
    // one NCCL group; every message is received directly in its halo range [ref: halos/exchange_halos_gpu.cuh:34-119]
    std::vector<CommMsg> sends, recvs;
    for (int q = 0; q < P; ++q)
    {
        if (q == rank_) continue;
        for (int a = 0; a < narrays; ++a)
        {
            if (haloSendCount_[q])
                sends.push_back({q, haloPack_.template as<char>() + off[a] + size_t(haloSendOffset_[q]) * elemBytes[a],
                                 size_t(haloSendCount_[q]) * elemBytes[a]});
            if (haloRecvCount_[q])
                recvs.push_back({q, reinterpret_cast<char*>(arrays[a]) + size_t(haloRecvOffset_[q]) * elemBytes[a],
                                 size_t(haloRecvCount_[q]) * elemBytes[a]});
        }
    }
    return comm_->exchange(sends, recvs, s_);
}

//! the window of this rank holds two slots of slotBytes; a collective when it has to grow (every rank sees the same need)
template<class K, class T>
int Engine<K, T>::ensureWindow(size_t slotBytes)
{
    if (win_ && slotBytes <= winSlot_) return 0;
    const size_t slot = (size_t(double(slotBytes) * 1.5) + (size_t(1) << 20) + 255) / 256 * 256;
    // the peers stop using the old window before its mappings go away (it is freed with the engine)
    CSB_CUDA(cudaStreamSynchronize(s_));
    if (win_)
    {
        CSB_TRY(small_.reserve(4096));
        CSB_CUDA(cudaMemsetAsync(small_.p, 0, 4, s_));
        CSB_TRY(comm_->allReduce(small_.p, 1, CommType::i32, CommOp::sum, s_));
        CSB_CUDA(cudaStreamSynchronize(s_));
        comm_->releaseShared(winPeers_);
        winOld_.push_back(win_);
        win_ = nullptr;
    }
    CSB_CUDA(cudaMalloc(&win_, kWinHeader + 2 * slot));
    CSB_CUDA(cudaMemsetAsync(win_, 0, kWinHeader, s_));
    winSlot_ = slot;
    // peer memory may be unavailable (no P2P between two GPUs): all ranks agree, and fall back to send/recv together
    int32_t failed = comm_->shareBuffer(win_, winPeers_, s_) != 0 ? 1 : 0;
    CSB_TRY(small_.reserve(4096));
    CSB_CUDA(cudaMemcpyAsync(small_.p, &failed, 4, cudaMemcpyHostToDevice, s_));
    CSB_TRY(comm_->allReduce(small_.p, 1, CommType::i32, CommOp::max, s_));
    CSB_CUDA(cudaMemcpyAsync(&failed, small_.p, 4, cudaMemcpyDeviceToHost, s_));
    CSB_CUDA(cudaStreamSynchronize(s_));
    if (failed)
    {
        if (!winPeers_.empty()) comm_->releaseShared(winPeers_);
        winOld_.push_back(win_);
        win_ = nullptr, winSlot_ = 0, haloPush_ = 0;
        return CSTONE_E_PEER_TIMEOUT; // caller switches to the send/recv path
    }
    return 0;
}

/*! pack-and-push exchangeHalos: see pushHalosKernel.  Same arguments and result as the send/recv path. */
template<class K, class T>
int Engine<K, T>::exchangeHalosPush(int narrays, void* const* arrays, const int* elemBytes)
{
    const int P = nRanks_;
    PushArgs p{};
    UnpackArgs u{};
    p.na = u.na = narrays, p.P = u.P = P, p.me = u.me = rank_;
    for (int a = 0; a < narrays; ++a)
    {
        p.src[a] = arrays[a], u.dst[a] = arrays[a];
        p.words[a] = u.words[a] = elemBytes[a] / 4;
    }
    // geometry of every rank's window for this call: the same numbers on all ranks
    size_t need = 0;
    for (int q = 0; q < P; ++q)
    {
        uint32_t total = 0;
        for (int r = 0; r < P; ++r)
        {
            if (r == rank_) p.dstElemOff[q] = total;
            if (q == rank_) u.srcElemOff[r] = total;
            total += haloMatrix_[size_t(r) * P + q];
        }
        p.recvTotal[q] = total;
        if (q == rank_) u.srcElemOff[P] = total;
        need = std::max(need, pushRegionOffset(p.words, narrays, total));
    }
    {
        const int st = ensureWindow(need);
        if (st == CSTONE_E_PEER_TIMEOUT && haloPush_ == 0) return exchangeHalos(narrays, arrays, elemBytes);
        if (st) return st;
    }
    const unsigned long long epoch = ++winEpoch_;
    const int slot                 = int(epoch & 1);
    for (int q = 0; q < P; ++q)
    {
        char* base        = static_cast<char*>(winPeers_[q]);
        p.peerData[q]     = base + kWinHeader + size_t(slot) * winSlot_;
        p.peerFlag[q]     = reinterpret_cast<unsigned long long*>(base) + size_t(slot) * kMaxPushRanks + rank_;
        p.sendOffset[q]   = haloSendOffset_[q];
        u.recvCount[q]    = haloRecvCount_[q];
        u.recvOffset[q]   = haloRecvOffset_[q];
    }
    p.sendOffset[P] = haloSendOffset_[P];
    p.nSend = uint32_t(nHaloSend_), p.start = uint32_t(start_), p.epoch = u.epoch = epoch;
    char* mine = static_cast<char*>(win_);
    p.ticket   = reinterpret_cast<unsigned int*>(mine + 3072);
    u.status   = reinterpret_cast<unsigned int*>(mine + 3080);
    u.myData   = mine + kWinHeader + size_t(slot) * winSlot_;
    u.myFlags  = reinterpret_cast<const unsigned long long*>(mine) + size_t(slot) * kMaxPushRanks;
    u.spin     = comm_->hostSynchronous() ? 0 : 1;
    {
        dim3 grid(std::max(1u, std::min(gridFor(size_t(nHaloSend_), 256, 2), 148u * 4u)), unsigned(narrays));
        pushHalosKernel<<<grid, 256, 0, s_>>>(sendIdx_.template as<uint32_t>(), p);
        CSB_LAUNCHED();
    }
    CSB_TRY(comm_->hostBarrier(s_)); // thread-ranks on one device: pushes complete before anybody unpacks
    {
        dim3 grid(std::max(1u, std::min(gridFor(size_t(u.srcElemOff[P]), 256, 2), 148u * 4u)), unsigned(narrays));
        unpackHalosKernel<<<grid, 256, 0, s_>>>(u);
        CSB_LAUNCHED();
    }
    return 0;
}

/*! [ref: domain/domain.hpp:334-378 reapplySync] the particle exchange and re-ordering of the LAST sync applied to further
 *  arrays.  in[a]: the array in the layout the caller's arrays had BEFORE that sync (device pointer, at least
 *  inStart + np elements); out[a]: nParticlesWithHalos elements of the new layout, of which the assigned range
 *  [startIndex, endIndex) is written (halos: exchangeHalos).  The reference does this for host arrays only. */
template<class K, class T>
int Engine<K, T>::reapplySync(int narrays, const void* const* in, const int* eb, void* const* out)
{
    if (!pat_.valid) return CSTONE_E_INVALID_ARGUMENT;
    if (narrays == 0) return 0;
    if (narrays < 0 || narrays > 4 + kMaxProps || !in || !eb || !out) return CSTONE_E_INVALID_ARGUMENT;
    for (int a = 0; a < narrays; ++a)
        if ((eb[a] != 4 && eb[a] != 8 && eb[a] != 12 && eb[a] != 16) || !in[a] || !out[a]) return CSTONE_E_INVALID_ARGUMENT;
    const size_t np = size_t(pat_.np), n = size_t(pat_.nAssigned);
    const void* src[4 + kMaxProps];
    void* dst[4 + kMaxProps];
    for (int a = 0; a < narrays; ++a)
    {
        src[a] = static_cast<const char*>(in[a]) + size_t(pat_.inStart) * eb[a];
        dst[a] = static_cast<char*>(out[a]) + size_t(pat_.start) * eb[a];
    }
    if (!multi())
    {
        if (np) CSB_TRY(gatherMulti(order_.template as<uint32_t>(), np, narrays, src, dst, eb));
        return int(cudaStreamSynchronize(s_));
    }
    const int P = nRanks_;
    const auto &sendStart = pat_.sendStart, &sendCount = pat_.sendCount, &recvOff = pat_.recvOff, &recvCount = pat_.recvCount;
    const uint32_t* order1 = pat_.steady ? order_.template as<uint32_t>() : order1Keep_.template as<uint32_t>();
    void* pk[4 + kMaxProps];
    for (int a = 0; a < narrays; ++a)
    {
        CSB_TRY(pack_[a].reserve(size_t(eb[a]) * np + 256));
        CSB_TRY(recv_[a].reserve(size_t(eb[a]) * n + 256));
        pk[a] = pack_[a].p;
    }
    if (!pat_.steady)
    {
        if (np) CSB_TRY(gatherMulti(order1, np, narrays, src, pk, eb));
    }
    else
    {
        const size_t lo = sendStart[rank_], hiStart = size_t(sendStart[rank_]) + sendCount[rank_];
        void* d2[4 + kMaxProps];
        if (lo) CSB_TRY(gatherMulti(order1, lo, narrays, src, pk, eb));
        if (np > hiStart)
        {
            for (int a = 0; a < narrays; ++a)
                d2[a] = pack_[a].template as<char>() + hiStart * size_t(eb[a]);
            CSB_TRY(gatherMulti(order1 + hiStart, np - hiStart, narrays, src, d2, eb));
        }
    }
    {
        std::vector<CommMsg> sends, recvs;
        for (int q = 0; q < P; ++q)
        {
            if (q == rank_) continue;
            for (int a = 0; a < narrays; ++a)
            {
                if (sendCount[q])
                    sends.push_back({q, pack_[a].template as<char>() + size_t(sendStart[q]) * eb[a], size_t(sendCount[q]) * eb[a]});
                if (recvCount[q])
                    recvs.push_back({q, recv_[a].template as<char>() + size_t(recvOff[q]) * eb[a], size_t(recvCount[q]) * eb[a]});
            }
        }
        CSB_TRY(comm_->exchange(sends, recvs, s_));
    }
    if (!pat_.steady)
    {
        for (int a = 0; a < narrays; ++a)
            if (sendCount[rank_])
                CSB_CUDA(cudaMemcpyAsync(recv_[a].template as<char>() + size_t(recvOff[rank_]) * eb[a],
                                         pack_[a].template as<char>() + size_t(sendStart[rank_]) * eb[a],
                                         size_t(sendCount[rank_]) * eb[a], cudaMemcpyDeviceToDevice, s_));
        const void* rsrc[4 + kMaxProps];
        for (int a = 0; a < narrays; ++a)
            rsrc[a] = recv_[a].p;
        if (n) CSB_TRY(gatherMulti(order_.template as<uint32_t>(), n, narrays, rsrc, dst, eb));
    }
    else if (n)
    {
        GatherTwoArgs g{};
        g.na = narrays;
        for (int a = 0; a < narrays; ++a)
            g.orig[a] = src[a], g.recv[a] = recv_[a].p, g.dst[a] = dst[a], g.words[a] = eb[a] / 4;
        gatherTwoKernel<<<gridFor(n, 256, 8), 256, 0, s_>>>(map2_, n, g);
        CSB_LAUNCHED();
    }
    return int(cudaStreamSynchronize(s_));
}

template<class K, class T>
int Engine<K, T>::syncFinish(K* keys, T* x, T* y, T* z, T* h, void* const* props)
{
    if (nTotal_ && (!keys || !x || !y || !z || !h)) return CSTONE_E_INVALID_ARGUMENT;
    void* out[4] = {x, y, z, h};
    CSB_CUDA(cudaMemcpyAsync(keys, fkeys_.p, sizeof(K) * nTotal_, cudaMemcpyDeviceToDevice, s_));
    for (int a = 0; a < 4; ++a)
        CSB_CUDA(cudaMemcpyAsync(out[a], fin_[a].p, sizeof(T) * nTotal_, cudaMemcpyDeviceToDevice, s_));
    for (int a = 0; a < nprops_; ++a)
    {
        if (!props || !props[a]) return CSTONE_E_INVALID_ARGUMENT;
        size_t b = size_t(propBytes_[a]);
        CSB_CUDA(cudaMemcpyAsync(reinterpret_cast<char*>(props[a]) + start_ * b, fin_[4 + a].template as<char>() + start_ * b,
                                 (end_ - start_) * b, cudaMemcpyDeviceToDevice, s_));
    }
    CSB_CUDA(cudaStreamSynchronize(s_));
    return 0;
}

template<class K, class T>
int Engine<K, T>::findNeighbors(const T* x, const T* y, const T* z, const T* h, uint32_t ngmax, uint32_t* nb,
                                uint32_t* nc) const
{
    static const bool noInfo = std::getenv("CSB_NB_NOINFO") != nullptr; // measurement switch: walk without node records
    void* temp = nullptr;
    if (!noInfo)
    {
        CSB_TRY(nbInfo_.reserve(cstone_b200_find_neighbors_temp_bytes(search_.numNodes)));
        temp = nbInfo_.p;
    }
    return Api<K, T>::neighbors(x, y, z, h, uint32_t(start_), uint32_t(end_), &box_, childOffsets_.template as<int32_t>(),
                                internalToLeaf_.template as<int32_t>(), layout_.template as<uint32_t>(), centers_.template as<T>(),
                                sizes_.template as<T>(), ngmax, nb, nc, search_.numNodes, temp, s_);
}

} // namespace csb

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

int cstone_b200_nccl_unique_id(void* id128)
{
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (ncclGetUniqueId(&id) != ncclSuccess) return CSTONE_E_NCCL;
    std::memcpy(id128, &id, 128);
    return 0;
}
int cstone_b200_nccl_comm_init(const void* id128, int nranks, int rank, void** commOut)
{
    ncclUniqueId id;
    std::memcpy(&id, id128, 128);
    ncclComm_t comm;
    if (ncclCommInitRank(&comm, nranks, id, rank) != ncclSuccess) return CSTONE_E_NCCL;
    *commOut = comm;
    return 0;
}
int cstone_b200_nccl_comm_destroy(void* comm)
{
    return ncclCommDestroy(reinterpret_cast<ncclComm_t>(comm)) == ncclSuccess ? 0 : CSTONE_E_NCCL;
}

/* in-process transport (comm.hpp LocalComm): one world per group of thread-ranks that share a GPU */
void* cstone_b200_local_world_create(int nranks) { return nranks >= 1 ? new (std::nothrow) csb::LocalWorld(nranks) : nullptr; }
void cstone_b200_local_world_destroy(void* world) { delete static_cast<csb::LocalWorld*>(world); }

#define CSB_ENGINE(K, T, d) reinterpret_cast<csb::Engine<K, T>*>(d)
#define CSB_CENGINE(K, T, d) reinterpret_cast<const csb::Engine<K, T>*>(d)

#define CSB_DOMAIN_DEF(KS, K, TS, T)                                                                                   \
    cstone_b200_domain* cstone_b200_domain_create_##KS##_##TS(int rank, int nranks, unsigned bucket, unsigned bucketFocus,\
                                                              float theta, const cstone_box_##TS* box)                 \
    {                                                                                                                  \
        if (!box || nranks < 1 || rank < 0 || rank >= nranks || bucket < bucketFocus || bucketFocus == 0) return nullptr;\
        return reinterpret_cast<cstone_b200_domain*>(new csb::Engine<K, T>(rank, nranks, bucket, bucketFocus, theta, *box));\
    }                                                                                                                  \
    void cstone_b200_domain_destroy_##KS##_##TS(cstone_b200_domain* d) { delete CSB_ENGINE(K, T, d); }                 \
    int cstone_b200_domain_set_comm_##KS##_##TS(cstone_b200_domain* d, void* comm)                                     \
    {                                                                                                                  \
        if (!d || !comm) return CSTONE_E_INVALID_ARGUMENT;                                                             \
        CSB_ENGINE(K, T, d)->setComm(comm);                                                                            \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_set_local_world_##KS##_##TS(cstone_b200_domain* d, void* world)                             \
    {                                                                                                                  \
        if (!d || !world) return CSTONE_E_INVALID_ARGUMENT;                                                            \
        CSB_ENGINE(K, T, d)->setLocalWorld(static_cast<csb::LocalWorld*>(world));                                      \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_set_sfc_kind_##KS##_##TS(cstone_b200_domain* d, int kind)                                   \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->setSfcKind(kind);                                                                  \
    }                                                                                                                  \
    int cstone_b200_domain_set_end_index_##KS##_##TS(cstone_b200_domain* d, uint64_t end)                              \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->setEndIndex(end);                                                                  \
    }                                                                                                                  \
    int cstone_b200_domain_set_halo_factor_##KS##_##TS(cstone_b200_domain* d, float factor)                            \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->setHaloFactor(factor);                                                             \
    }                                                                                                                  \
    int cstone_b200_domain_set_halo_push_##KS##_##TS(cstone_b200_domain* d, int on)                                    \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->setHaloPush(on);                                                                   \
    }                                                                                                                  \
    int cstone_b200_domain_set_halo_mode_##KS##_##TS(cstone_b200_domain* d, int mode)                                  \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->setHaloMode(mode);                                                                 \
    }                                                                                                                  \
    int cstone_b200_domain_let_view_##KS##_##TS(const cstone_b200_domain* d, int32_t* numLeaves, const K** leaves,     \
                                                const uint32_t** counts, const int32_t** flags,                        \
                                                const uint32_t** layout, int32_t* assignment2P, int32_t* peers,        \
                                                int32_t* numPeers)                                                     \
    {                                                                                                                  \
        if (!d || !numLeaves || !leaves || !counts || !flags || !layout || !assignment2P || !peers || !numPeers)       \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        auto* e = reinterpret_cast<const csb::Engine<K, T>*>(d);                                                      \
        if (!e->let() || e->letLayout().empty()) return CSTONE_E_INVALID_ARGUMENT;                                     \
        *numLeaves = e->let()->numLeaves();                                                                            \
        *leaves    = e->let()->leaves().data();                                                                        \
        *counts    = e->let()->leafCounts().data();                                                                    \
        *flags     = e->letFlags().data();                                                                             \
        *layout    = e->letLayout().data();                                                                            \
        int q      = 0;                                                                                                \
        for (const auto& a : e->let()->assignment())                                                                   \
            assignment2P[2 * q] = a[0], assignment2P[2 * q + 1] = a[1], ++q;                                           \
        *numPeers = int32_t(e->let()->peers().size());                                                                 \
        for (int k = 0; k < *numPeers; ++k)                                                                            \
            peers[k] = e->let()->peers()[k];                                                                           \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_cells_##KS##_##TS(cstone_b200_domain* d, int32_t* startEnd2)                                \
    {                                                                                                                  \
        if (!d || !startEnd2) return CSTONE_E_INVALID_ARGUMENT;                                                        \
        return CSB_ENGINE(K, T, d)->cells(startEnd2);                                                                  \
    }                                                                                                                  \
    int cstone_b200_domain_set_fields_ready_event_##KS##_##TS(cstone_b200_domain* d, void* event)                      \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        CSB_ENGINE(K, T, d)->setFieldsReadyEvent(event);                                                               \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int64_t cstone_b200_domain_sync_begin_##KS##_##TS(cstone_b200_domain* d, uint64_t n, const T* x, const T* y, const T* z,\
                                                      const T* h, int nprops, const void* const* props, const int* pb) \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->syncBegin(n, x, y, z, h, nprops, props, pb);                                       \
    }                                                                                                                  \
    int64_t cstone_b200_domain_sync_grav_begin_##KS##_##TS(cstone_b200_domain* d, uint64_t n, const T* x, const T* y, const T* z,\
                                                      const T* h, int nprops, const void* const* props, const int* pb) \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->syncGravBegin(n, x, y, z, h, nprops, props, pb);                                       \
    }                                                                                                                  \
    int cstone_b200_domain_focus_peer_exchange_##KS##_##TS(cstone_b200_domain* d, void* nodeQ, int eb)                 \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->focusPeerExchange(nodeQ, eb);                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_focus_global_gather_##KS##_##TS(cstone_b200_domain* d, const void* nodeQ, int eb,           \
                                                           void* gLeafQ)                                               \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->focusGlobalGather(nodeQ, eb, gLeafQ);                                              \
    }                                                                                                                  \
    int cstone_b200_domain_focus_global_extract_##KS##_##TS(cstone_b200_domain* d, const void* gNodeQ, int eb,         \
                                                            void* nodeQ)                                               \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->focusGlobalExtract(gNodeQ, eb, nodeQ);                                             \
    }                                                                                                                  \
    int cstone_b200_domain_global_octree_##KS##_##TS(const cstone_b200_domain* d, int32_t* numLeaves,                  \
                                                     int32_t* numNodes, const void** ptrs6, int32_t* levelRangeHost)   \
    {                                                                                                                  \
        if (!d || !numLeaves || !numNodes || !ptrs6 || !levelRangeHost) return CSTONE_E_INVALID_ARGUMENT;              \
        return reinterpret_cast<const csb::Engine<K, T>*>(d)->globalOctree(numLeaves, numNodes, ptrs6, levelRangeHost);\
    }                                                                                                                  \
    int cstone_b200_domain_let_centers_##KS##_##TS(const cstone_b200_domain* d, const T** centers4, const char** macs, \
                                                   int32_t* numNodes)                                                  \
    {                                                                                                                  \
        if (!d || !centers4 || !macs || !numNodes) return CSTONE_E_INVALID_ARGUMENT;                                   \
        const void* c = nullptr;                                                                                       \
        reinterpret_cast<const csb::Engine<K, T>*>(d)->letCenters(&c, macs, numNodes);                                 \
        *centers4 = static_cast<const T*>(c);                                                                          \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_sync_finish_##KS##_##TS(cstone_b200_domain* d, K* keys, T* x, T* y, T* z, T* h,             \
                                                   void* const* props)                                                 \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->syncFinish(keys, x, y, z, h, props);                                               \
    }                                                                                                                  \
    int cstone_b200_domain_exchange_halos_##KS##_##TS(cstone_b200_domain* d, int na, void* const* arrays, const int* eb)\
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        int st = CSB_ENGINE(K, T, d)->exchangeHalos(na, arrays, eb);                                                   \
        if (st) return st;                                                                                             \
        return int(cudaStreamSynchronize(nullptr));                                                                    \
    }                                                                                                                  \
    int cstone_b200_domain_reapply_sync_##KS##_##TS(cstone_b200_domain* d, int na, const void* const* in,              \
                                                    const int* eb, void* const* out)                                   \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_ENGINE(K, T, d)->reapplySync(na, in, eb, out);                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_info_##KS##_##TS(const cstone_b200_domain* d, int64_t* info8)                               \
    {                                                                                                                  \
        if (!d || !info8) return CSTONE_E_INVALID_ARGUMENT;                                                            \
        CSB_CENGINE(K, T, d)->info(info8);                                                                             \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_stage_ms_##KS##_##TS(const cstone_b200_domain* d, float* ms10)                              \
    {                                                                                                                  \
        if (!d || !ms10) return CSTONE_E_INVALID_ARGUMENT;                                                             \
        CSB_CENGINE(K, T, d)->stageMs(ms10);                                                                           \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_box_##KS##_##TS(const cstone_b200_domain* d, cstone_box_##TS* out)                          \
    {                                                                                                                  \
        if (!d || !out) return CSTONE_E_INVALID_ARGUMENT;                                                              \
        *out = CSB_CENGINE(K, T, d)->box();                                                                            \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_tree_view_##KS##_##TS(const cstone_b200_domain* d, const void** ptrs10)                     \
    {                                                                                                                  \
        if (!d || !ptrs10) return CSTONE_E_INVALID_ARGUMENT;                                                           \
        CSB_CENGINE(K, T, d)->treeView(ptrs10);                                                                        \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_global_tree_##KS##_##TS(const cstone_b200_domain* d, const K** leaves, const uint32_t** counts,\
                                                   const K** assignment)                                               \
    {                                                                                                                  \
        if (!d || !leaves || !counts || !assignment) return CSTONE_E_INVALID_ARGUMENT;                                 \
        CSB_CENGINE(K, T, d)->globalTree(leaves, counts, assignment);                                                  \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_domain_find_neighbors_##KS##_##TS(const cstone_b200_domain* d, const T* x, const T* y, const T* z, \
                                                      const T* h, uint32_t ngmax, uint32_t* nb, uint32_t* nc)          \
    {                                                                                                                  \
        if (!d) return CSTONE_E_INVALID_ARGUMENT;                                                                      \
        return CSB_CENGINE(K, T, d)->findNeighbors(x, y, z, h, ngmax, nb, nc);                                         \
    }

CSB_DOMAIN_DEF(k64, uint64_t, f64, double)
CSB_DOMAIN_DEF(k64, uint64_t, f32, float)
CSB_DOMAIN_DEF(k32, uint32_t, f64, double)
CSB_DOMAIN_DEF(k32, uint32_t, f32, float)

int cstone_b200_host_assignment_k64(int nranks, const uint64_t* leaves, const uint32_t* counts, int32_t numLeaves,
                                    const uint64_t* oldB, uint64_t* boundaries, uint32_t* rankCounts)
{
    if (nranks < 1 || numLeaves < 1 || !leaves || !counts || !boundaries || !rankCounts) return CSTONE_E_INVALID_ARGUMENT;
    std::vector<uint64_t> b;
    std::vector<uint32_t> c;
    csb::makeAssignment<uint64_t>(nranks, leaves, counts, numLeaves, oldB, b, c);
    std::copy(b.begin(), b.end(), boundaries);
    std::copy(c.begin(), c.end(), rankCounts);
    return 0;
}
int cstone_b200_host_assignment_k32(int nranks, const uint32_t* leaves, const uint32_t* counts, int32_t numLeaves,
                                    const uint32_t* oldB, uint32_t* boundaries, uint32_t* rankCounts)
{
    if (nranks < 1 || numLeaves < 1 || !leaves || !counts || !boundaries || !rankCounts) return CSTONE_E_INVALID_ARGUMENT;
    std::vector<uint32_t> b;
    std::vector<uint32_t> c;
    csb::makeAssignment<uint32_t>(nranks, leaves, counts, numLeaves, oldB, b, c);
    std::copy(b.begin(), b.end(), boundaries);
    std::copy(c.begin(), c.end(), rankCounts);
    return 0;
}
#define CSB_HOST_TREELET(KS, K)                                                                                        \
    int cstone_b200_host_update_treelet_##KS(const K* leaves, const uint32_t* counts, int32_t first, int32_t last,     \
                                             uint32_t bucket, K* out, int32_t capacity)                                \
    {                                                                                                                  \
        if (!leaves || !counts || !out || first < 0 || last < first) return CSTONE_E_INVALID_ARGUMENT;                 \
        std::vector<int> ops(size_t(last - first));                                                                    \
        for (int i = first; i < last; ++i)                                                                             \
            ops[i - first] = csb::hostNodeOp<K>(leaves, i, counts, bucket);                                            \
        std::vector<K> tl = csb::hostApplyOps<K>(leaves, first, last, ops);                                            \
        if (int64_t(tl.size()) > int64_t(capacity)) return CSTONE_E_CAPACITY;                                          \
        std::memcpy(out, tl.data(), tl.size() * sizeof(K));                                                            \
        return int(tl.size());                                                                                         \
    }
CSB_HOST_TREELET(k32, uint32_t)
CSB_HOST_TREELET(k64, uint64_t)

#define CSB_HOST_PEERS(KS, K, TS, T)                                                                                   \
    int cstone_b200_host_find_peers_##KS##_##TS(int sfcKind, int myRank, int nranks, const K* boundaries,              \
                                                const K* leaves, int32_t numLeaves, const cstone_box_##TS* box,        \
                                                float invThetaEff, int32_t* isPeer)                                    \
    {                                                                                                                  \
        if (!boundaries || !leaves || !box || !isPeer || numLeaves < 1 || nranks < 1 || myRank < 0 || myRank >= nranks) \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        csb::HostOctree<K> tree;                                                                                       \
        tree.build(leaves, numLeaves);                                                                                 \
        std::vector<K> bnd(boundaries, boundaries + nranks + 1);                                                       \
        std::vector<int> peers = csb::findPeersMacHost<K, T>(sfcKind, myRank, bnd, tree, *box, invThetaEff);           \
        for (int r = 0; r < nranks; ++r)                                                                               \
            isPeer[r] = 0;                                                                                             \
        for (int p : peers)                                                                                            \
            isPeer[p] = 1;                                                                                             \
        return 0;                                                                                                      \
    }
CSB_HOST_PEERS(k32, uint32_t, f32, float)
CSB_HOST_PEERS(k32, uint32_t, f64, double)
CSB_HOST_PEERS(k64, uint64_t, f32, float)
CSB_HOST_PEERS(k64, uint64_t, f64, double)

int cstone_b200_host_uniform_bins(const uint32_t* counts, int32_t numLeaves, int32_t numBins, int32_t* bins,
                                  uint32_t* binCounts)
{
    if (numBins < 1 || numLeaves < 0 || !counts || !bins || !binCounts) return CSTONE_E_INVALID_ARGUMENT;
    std::vector<int> b;
    std::vector<uint32_t> c;
    csb::uniformBins(counts, numLeaves, numBins, b, c);
    std::copy(b.begin(), b.end(), bins);
    std::copy(c.begin(), c.end(), binCounts);
    return 0;
}
int cstone_b200_host_limit_boundary_shifts_k64(int nranks, const uint64_t* oldB, uint64_t* boundaries,
                                               const uint64_t* leaves, const uint32_t* counts, int32_t numLeaves,
                                               uint32_t* rankCounts)
{
    if (nranks < 1 || numLeaves < 1 || !oldB || !leaves || !counts || !boundaries || !rankCounts)
        return CSTONE_E_INVALID_ARGUMENT;
    csb::limitBoundaryShifts<uint64_t>(nranks, oldB, boundaries, leaves, counts, numLeaves, rankCounts);
    return 0;
}
int cstone_b200_host_limit_boundary_shifts_k32(int nranks, const uint32_t* oldB, uint32_t* boundaries,
                                               const uint32_t* leaves, const uint32_t* counts, int32_t numLeaves,
                                               uint32_t* rankCounts)
{
    if (nranks < 1 || numLeaves < 1 || !oldB || !leaves || !counts || !boundaries || !rankCounts)
        return CSTONE_E_INVALID_ARGUMENT;
    csb::limitBoundaryShifts<uint32_t>(nranks, oldB, boundaries, leaves, counts, numLeaves, rankCounts);
    return 0;
}
void cstone_b200_host_limit_box_shrinking_f64(const double* fit, const double* prev, double* out)
{
    csb::limitBoxShrinking<double>(fit, prev, out);
}
void cstone_b200_host_limit_box_shrinking_f32(const float* fit, const float* prev, float* out)
{
    csb::limitBoxShrinking<float>(fit, prev, out);
}

} // extern "C"
