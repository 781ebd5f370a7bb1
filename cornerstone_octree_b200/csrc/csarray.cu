/*
 * csarray.cu — a6-a9: cornerstone leaf array: particle counts, rebalance decision (fused into its scan),
 * leaf emission, update / build drivers.
 *
 * Replaces computeNodeCountsGpu, computeNodeOpsGpu, rebalanceTreeGpu [ref: tree/csarray_gpu.h:55-84,
 * tree/csarray_gpu.cu:50-226] and updateOctreeGpu [ref: tree/update_gpu.cuh:58-82]; semantics follow the CPU
 * functions in tree/csarray.hpp (cited below).
 */
#include "scan.cuh"

namespace csb
{

// ------------------------------------------------------------------------------------------------
// a6 counts.  counts[i] = min(lb(tree[i+1]) - lb(tree[i]), maxCount)  [ref: tree/csarray.hpp:93-103]
// Each block resolves 257 consecutive leaf boundaries into shared memory with one binary search per boundary
// (the reference does two per leaf), then differences them.  The search runs over the sorted particle keys;
// its upper levels are shared by neighbouring boundaries and stay in L2/L1.
// ------------------------------------------------------------------------------------------------
template<class K>
__device__ inline uint32_t lowerBound(const K* __restrict__ keys, uint32_t n, K value)
{
    uint32_t lo = 0, len = n;
    while (len > 0)
    {
        uint32_t half = len >> 1;
        K probe       = keys[lo + half];
        bool right    = probe < value;
        lo            = right ? lo + half + 1 : lo;
        len           = right ? len - half - 1 : half;
    }
    return lo;
}

constexpr int kCountThreads = 256;

template<class K>
__global__ void __launch_bounds__(kCountThreads) nodeCountsKernel(const K* __restrict__ tree,
                                                                   uint32_t* __restrict__ counts,
                                                                   int numNodes,
                                                                   const K* __restrict__ keys,
                                                                   uint32_t numKeys,
                                                                   uint32_t maxCount)
{
    __shared__ uint32_t sLb[kCountThreads + 1];
    for (int blockBase = blockIdx.x * kCountThreads; blockBase < numNodes; blockBase += gridDim.x * kCountThreads)
    {
        int i = blockBase + threadIdx.x;
        if (i <= numNodes) sLb[threadIdx.x] = lowerBound(keys, numKeys, tree[i]);
        if (threadIdx.x == 0)
        {
            int j = blockBase + kCountThreads;
            if (j <= numNodes) sLb[kCountThreads] = lowerBound(keys, numKeys, tree[j]);
        }
        __syncthreads();
        if (i < numNodes)
        {
            uint32_t c = sLb[threadIdx.x + 1] - sLb[threadIdx.x];
            counts[i]  = min(c, maxCount);
        }
        __syncthreads();
    }
}

template<class K>
int computeNodeCounts(const K* tree, uint32_t* counts, int numNodes, const K* first, const K* last,
                      uint32_t maxCount, cudaStream_t s)
{
    if (numNodes <= 0) return 0;
    if (!tree || !counts || last < first) return CSTONE_E_INVALID_ARGUMENT;
    size_t numKeys = size_t(last - first);
    if (numKeys > 0xFFFFFFFFull) return CSTONE_E_TOO_LARGE;
    unsigned grid = gridFor(size_t(numNodes), kCountThreads, 8);
    nodeCountsKernel<K><<<grid, kCountThreads, 0, s>>>(tree, counts, numNodes, first, uint32_t(numKeys), maxCount);
    CSB_LAUNCHED();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// a7 decision  [ref: tree/csarray.hpp:269-283 siblingAndLevel, :286-309 calculateNodeOp]
// ------------------------------------------------------------------------------------------------
template<class K>
__device__ inline int nodeOp(const K* __restrict__ tree, int i, const uint32_t* __restrict__ counts, uint32_t bucket)
{
    K start        = tree[i];
    unsigned level = treeLevel<K>(K(tree[i + 1] - start));

    int sibling = -1;
    if (level > 0)
    {
        sibling = int(octalDigit<K>(start, level));
        // the parent's range is fully covered by 8 leaves of this level iff the leaf 8 places after octant 0 starts
        // exactly one parent-range later
        if (tree[i - sibling + 8] != K(tree[i - sibling] + nodeRange<K>(level - 1))) sibling = -1;
    }
    if (sibling > 0)
    {
        const uint32_t* g  = counts + (i - sibling);
        unsigned long long parentCount = 0;
#pragma unroll
        for (int o = 0; o < 8; ++o)
            parentCount += g[o];
        if (parentCount <= (unsigned long long)bucket) return 0;
    }
    constexpr unsigned maxLevel = KeyTraits<K>::maxLevel;
    uint32_t c                  = counts[i];
    // 32-bit wrap-around of bucket*512 etc. is the reference's behaviour (unsigned arithmetic)
    if (c > bucket * 512u && level + 3 < maxLevel) return 4096;
    if (c > bucket * 64u && level + 2 < maxLevel) return 512;
    if (c > bucket * 8u && level + 1 < maxLevel) return 64;
    if (c > bucket && level < maxLevel) return 8;
    return 1;
}

template<class K>
struct LoadNodeOp
{
    const K* tree;
    const uint32_t* counts;
    uint32_t bucket;
    int numNodes;
    int* changed; // set to 1 if any op != 1
    __device__ uint32_t operator()(size_t i) const
    {
        if (int(i) >= numNodes) return 0u; // the extra slot that receives the total
        int op = nodeOp<K>(tree, int(i), counts, bucket);
        if (op != 1) *changed = 1; // same-value race, as in the reference [ref: tree/csarray_gpu.cu:150-167]
        return uint32_t(op);
    }
};

template<class K>
__global__ void __launch_bounds__(256)
    decisionKernel(const K* __restrict__ tree, const uint32_t* __restrict__ counts, int numNodes, uint32_t bucket,
                   int* __restrict__ ops)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numNodes; i += gridDim.x * blockDim.x)
        ops[i] = nodeOp<K>(tree, i, counts, bucket);
}

//! device scratch behind the scan temp: [0] total, [1] changed flag
struct OpsScratch
{
    uint32_t* total;
    int* changed;
    void* scanTemp;
};
inline OpsScratch carveOpsScratch(void* temp)
{
    OpsScratch o;
    o.total    = reinterpret_cast<uint32_t*>(temp);
    o.changed  = reinterpret_cast<int*>(temp) + 1;
    o.scanTemp = reinterpret_cast<char*>(temp) + 128;
    return o;
}
inline uint64_t opsTempBytes(uint64_t n) { return 128 + scanTempBytes(n); }

//! enqueue decision+scan; results stay on the device (o.total / o.changed)
template<class K>
int enqueueNodeOps(const K* tree, int numNodes, const uint32_t* counts, uint32_t bucket, int* nodeOps, void* temp,
                   cudaStream_t s)
{
    OpsScratch o = carveOpsScratch(temp);
    CSB_CUDA(cudaMemsetAsync(temp, 0, 128, s));
    LoadNodeOp<K> load{tree, counts, bucket, numNodes, o.changed};
    return launchScan<false>(load, reinterpret_cast<uint32_t*>(nodeOps), size_t(numNodes) + 1, 0u, o.scanTemp,
                             o.total, s);
}

template<class K>
int computeNodeOps(const K* tree, int numNodes, const uint32_t* counts, uint32_t bucket, int* nodeOps,
                   int* newNumNodes, int* converged, void* temp, cudaStream_t s)
{
    if (numNodes <= 0 || !tree || !counts || !nodeOps || !temp || !newNumNodes) return CSTONE_E_INVALID_ARGUMENT;
    int st = enqueueNodeOps<K>(tree, numNodes, counts, bucket, nodeOps, temp, s);
    if (st) return st;
    uint32_t host[2];
    CSB_CUDA(cudaMemcpyAsync(host, temp, 8, cudaMemcpyDeviceToHost, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    *newNumNodes = int(host[0]);
    if (converged) *converged = host[1] ? 0 : 1;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// a8 emission  [ref: tree/csarray.hpp:358-385 processNode, :395-409 rebalanceTree]
// One thread per old leaf writes 1 or 8 keys; multi-level splits (64/512/4096 keys) are written by the whole
// warp so that those stores are coalesced and the work per lane stays bounded.
// ------------------------------------------------------------------------------------------------
template<class K>
__global__ void __launch_bounds__(256) emitKernel(const K* __restrict__ tree,
                                                   int numNodes,
                                                   const int* __restrict__ scan,
                                                   K* __restrict__ newTree)
{
    int nWarpsTotal = (gridDim.x * blockDim.x) >> 5;
    int warpGlobal  = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (int base = warpGlobal * 32; base < numNodes; base += nWarpsTotal * 32)
    {
        int i      = base + int(laneId());
        int op     = 0;
        int dst    = 0;
        K start    = 0;
        K step     = 0;
        if (i < numNodes)
        {
            dst            = scan[i];
            op             = scan[i + 1] - dst;
            start          = tree[i];
            unsigned level = treeLevel<K>(K(tree[i + 1] - start));
            unsigned dl    = (op >= 4096) ? 4u : (op >= 512) ? 3u : (op >= 64) ? 2u : (op >= 8) ? 1u : 0u;
            step           = nodeRange<K>(level + dl);
            if (op == 1) newTree[dst] = start;
            else if (op == 8)
            {
#pragma unroll
                for (int sib = 0; sib < 8; ++sib)
                    newTree[dst + sib] = start + K(sib) * step;
            }
        }
        unsigned big = __ballot_sync(0xffffffffu, op > 8);
        while (big)
        {
            int src = __ffs(int(big)) - 1;
            big &= big - 1;
            int bop   = __shfl_sync(0xffffffffu, op, src);
            int bdst  = __shfl_sync(0xffffffffu, dst, src);
            K bstart  = __shfl_sync(0xffffffffu, start, src);
            K bstep   = __shfl_sync(0xffffffffu, step, src);
            for (int sib = int(laneId()); sib < bop; sib += 32)
                newTree[bdst + sib] = bstart + K(sib) * bstep;
        }
        if (i == numNodes - 1) newTree[scan[numNodes]] = tree[numNodes];
    }
}

template<class K>
int rebalanceTree(const K* tree, int numNodes, int newNumNodes, const int* nodeOps, K* newTree, cudaStream_t s)
{
    if (numNodes <= 0 || newNumNodes <= 0 || !tree || !nodeOps || !newTree) return CSTONE_E_INVALID_ARGUMENT;
    unsigned grid = gridFor(size_t(numNodes), 256, 8);
    emitKernel<K><<<grid, 256, 0, s>>>(tree, numNodes, nodeOps, newTree);
    CSB_LAUNCHED();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// a9 drivers  [ref: tree/csarray.hpp:429-465 updateOctree / computeOctree, tree/update_gpu.cuh:58-82]
// ------------------------------------------------------------------------------------------------
template<class K>
int updateOctree(const K* first, const K* last, uint32_t bucket, uint32_t maxCount, K* treeA, K* treeB,
                 uint32_t* counts, int* nodeOps, int capacity, int* numNodes, int* converged, void* temp,
                 cudaStream_t s)
{
    if (!treeA || !treeB || !counts || !nodeOps || !numNodes || !temp || capacity < 1) return CSTONE_E_INVALID_ARGUMENT;
    int n = *numNodes;
    if (n < 1 || n > capacity) return CSTONE_E_INVALID_ARGUMENT;
    int newN = 0, conv = 0;
    int st = computeNodeOps<K>(treeA, n, counts, bucket, nodeOps, &newN, &conv, temp, s);
    if (st) return st;
    if (newN > capacity)
    {
        *numNodes = newN;
        return CSTONE_E_CAPACITY;
    }
    st = rebalanceTree<K>(treeA, n, newN, nodeOps, treeB, s);
    if (st) return st;
    CSB_CUDA(cudaMemcpyAsync(treeA, treeB, sizeof(K) * (size_t(newN) + 1), cudaMemcpyDeviceToDevice, s));
    st = computeNodeCounts<K>(treeA, counts, newN, first, last, maxCount, s);
    if (st) return st;
    *numNodes = newN;
    if (converged) *converged = conv;
    return 0;
}

template<class K>
int computeOctree(const K* first, const K* last, uint32_t bucket, uint32_t maxCount, K* treeA, K* treeB,
                  uint32_t* counts, int* nodeOps, int capacity, int* numNodes, int* numIterations, void* temp,
                  cudaStream_t s)
{
    if (!treeA || !treeB || !counts || !nodeOps || !numNodes || !temp || capacity < 1) return CSTONE_E_INVALID_ARGUMENT;
    K root[2]       = {K(0), nodeRange<K>(0)};
    uint32_t count0 = uint32_t(last - first); // [ref: tree/csarray.hpp:458] not clamped
    CSB_CUDA(cudaMemcpyAsync(treeA, root, sizeof(root), cudaMemcpyHostToDevice, s));
    CSB_CUDA(cudaMemcpyAsync(counts, &count0, sizeof(count0), cudaMemcpyHostToDevice, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    int n = 1, conv = 0, iters = 0;
    while (!conv)
    {
        int st = updateOctree<K>(first, last, bucket, maxCount, treeA, treeB, counts, nodeOps, capacity, &n, &conv,
                                 temp, s);
        if (st)
        {
            *numNodes = n;
            return st;
        }
        ++iters;
        if (iters > 64) return CSTONE_E_INVALID_ARGUMENT; // cannot happen for sorted keys
    }
    *numNodes = n;
    if (numIterations) *numIterations = iters;
    return 0;
}

} // namespace csb

extern "C" {

#define CSB_TREE_API(KS, K)                                                                                            \
    int cstone_b200_compute_node_counts_##KS(const K* tree, uint32_t* counts, int32_t numNodes, const K* first,        \
                                             const K* last, uint32_t maxCount, int, void* stream)                      \
    {                                                                                                                  \
        return csb::computeNodeCounts<K>(tree, counts, numNodes, first, last, maxCount, csb::asStream(stream));        \
    }                                                                                                                  \
    int cstone_b200_compute_node_ops_##KS(const K* tree, int32_t numNodes, const uint32_t* counts, uint32_t bucket,    \
                                          int32_t* nodeOps, int32_t* newNumNodes, int32_t* converged, void* temp,      \
                                          void* stream)                                                                \
    {                                                                                                                  \
        return csb::computeNodeOps<K>(tree, numNodes, counts, bucket, nodeOps, newNumNodes, converged, temp,           \
                                      csb::asStream(stream));                                                          \
    }                                                                                                                  \
    int cstone_b200_rebalance_decision_##KS(const K* tree, int32_t numNodes, const uint32_t* counts, uint32_t bucket,  \
                                            int32_t* nodeOps, void* stream)                                            \
    {                                                                                                                  \
        if (numNodes <= 0) return 0;                                                                                   \
        csb::decisionKernel<K><<<csb::gridFor(size_t(numNodes), 256), 256, 0, csb::asStream(stream)>>>(                \
            tree, counts, numNodes, bucket, nodeOps);                                                                  \
        CSB_LAUNCHED();                                                                                                \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_rebalance_tree_##KS(const K* tree, int32_t numNodes, int32_t newNumNodes, const int32_t* nodeOps,  \
                                        K* newTree, void* stream)                                                      \
    {                                                                                                                  \
        return csb::rebalanceTree<K>(tree, numNodes, newNumNodes, nodeOps, newTree, csb::asStream(stream));            \
    }                                                                                                                  \
    int cstone_b200_update_octree_##KS(const K* first, const K* last, uint32_t bucket, uint32_t maxCount, K* treeA,    \
                                       K* treeB, uint32_t* counts, int32_t* nodeOps, int32_t capacity,                 \
                                       int32_t* numNodes, int32_t* converged, void* temp, void* stream)                \
    {                                                                                                                  \
        return csb::updateOctree<K>(first, last, bucket, maxCount, treeA, treeB, counts, nodeOps, capacity, numNodes,  \
                                    converged, temp, csb::asStream(stream));                                           \
    }                                                                                                                  \
    int cstone_b200_compute_octree_##KS(const K* first, const K* last, uint32_t bucket, uint32_t maxCount, K* treeA,   \
                                        K* treeB, uint32_t* counts, int32_t* nodeOps, int32_t capacity,                \
                                        int32_t* numNodes, int32_t* numIterations, void* temp, void* stream)           \
    {                                                                                                                  \
        return csb::computeOctree<K>(first, last, bucket, maxCount, treeA, treeB, counts, nodeOps, capacity, numNodes, \
                                     numIterations, temp, csb::asStream(stream));                                      \
    }

CSB_TREE_API(k32, uint32_t)
CSB_TREE_API(k64, uint64_t)

} // extern "C"
