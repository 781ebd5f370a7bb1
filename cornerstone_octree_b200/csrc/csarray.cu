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
// (the reference does two per leaf), then differences them.  The key range of the block is staged in shared memory
// by TMA bulk copies (nodeCountsTmaKernel below).
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

//! fallback (key array not 16-byte aligned): one global binary search per boundary
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

/* TMA-staged variant.  The 257 boundaries of 256 consecutive leaves fall into ONE contiguous range [lo, hi) of the
 * sorted particle keys.  Two warps find lo and hi with a 32-ary cooperative search (5 dependent rounds instead of 25),
 * the range travels to shared memory as bulk-copy tiles (cp.async.bulk + mbarrier, one elected thread), and every
 * thread resolves its boundary with a binary search in shared memory.  Each particle key is read from HBM once
 * (sK per particle) instead of ~25 dependent L2 probes per leaf.  Ranges that are far longer than the leaves they
 * serve (saturated global-tree leaves) keep the direct search, restricted to [lo, hi). */
constexpr int kCountTileKeys = 4096;
constexpr uint32_t kCountMaxChunks = 4;

__device__ inline void mbarInit(uint64_t* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(uint32_t(__cvta_generic_to_shared(bar))), "r"(count));
}
__device__ inline void mbarExpectTx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(uint32_t(__cvta_generic_to_shared(bar))),
                 "r"(bytes)
                 : "memory");
}
__device__ inline void mbarWait(uint64_t* bar, uint32_t phase)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@!p bra WAIT_%=;\n\t}" ::"r"(
                     uint32_t(__cvta_generic_to_shared(bar))),
                 "r"(phase)
                 : "memory");
}
__device__ inline void bulkCopyToShared(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     uint32_t(__cvta_generic_to_shared(dst))),
                 "l"(src), "r"(bytes), "r"(uint32_t(__cvta_generic_to_shared(bar)))
                 : "memory");
}
__device__ inline void fenceProxyAsync() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

//! lower bound of value in keys[0, n) found by one warp: 32 probes per round
template<class K>
__device__ inline uint32_t warpLowerBound(const K* __restrict__ keys, uint32_t n, K value)
{
    const unsigned lane = threadIdx.x & 31u;
    uint32_t lo = 0, len = n;
    while (len > 32u)
    {
        const uint32_t p   = lo + uint32_t((uint64_t(lane + 1u) * len) / 33u);
        const unsigned lt  = __ballot_sync(0xffffffffu, keys[p] < value); // a prefix of the lanes (sorted keys)
        const unsigned cnt = __popc(lt);
        const uint32_t pPrev = __shfl_sync(0xffffffffu, p, (cnt + 31u) & 31u); // probe cnt-1
        const uint32_t pNext = __shfl_sync(0xffffffffu, p, cnt & 31u);         // probe cnt
        const uint32_t newLo = cnt ? pPrev + 1u : lo;
        const uint32_t newHi = cnt < 32u ? pNext : lo + len;                   // answer in [newLo, newHi]
        lo  = newLo;
        len = newHi - newLo;
    }
    const bool lt = lane < len && keys[lo + lane] < value;
    return lo + __popc(__ballot_sync(0xffffffffu, lt));
}

template<class K>
__global__ void __launch_bounds__(kCountThreads) nodeCountsTmaKernel(const K* __restrict__ tree,
                                                                      uint32_t* __restrict__ counts,
                                                                      int numNodes,
                                                                      const K* __restrict__ keys,
                                                                      uint32_t numKeys,
                                                                      uint32_t maxCount)
{
    constexpr uint32_t E = 16u / sizeof(K); // keys per 16 bytes: bulk copies start and end on 16-byte boundaries
    __shared__ alignas(128) K sKeys[kCountTileKeys];
    __shared__ uint32_t sLb[kCountThreads + 1];
    __shared__ uint32_t sRange[2];
    __shared__ alignas(8) uint64_t sBar;

    const unsigned tid = threadIdx.x;
    if (tid == 0)
    {
        mbarInit(&sBar, 1);
        fenceProxyAsync();
    }
    __syncthreads();
    uint32_t phase = 0;
    const uint32_t numKeysAligned = numKeys & ~(E - 1u);

    for (int blockBase = blockIdx.x * kCountThreads; blockBase < numNodes; blockBase += gridDim.x * kCountThreads)
    {
        const int nb = min(kCountThreads, numNodes - blockBase); // leaves of this tile; boundaries blockBase .. +nb
        if (tid < 32u)
        {
            const uint32_t r = warpLowerBound(keys, numKeys, tree[blockBase]);
            if (tid == 0) sRange[0] = r;
        }
        else if (tid < 64u)
        {
            const uint32_t r = warpLowerBound(keys, numKeys, tree[blockBase + nb]);
            if (tid == 32u) sRange[1] = r;
        }
        __syncthreads();
        const uint32_t lo = sRange[0], hi = sRange[1];
        const bool mine   = int(tid) < nb;
        const K v         = mine ? tree[blockBase + int(tid)] : K(0);
        uint32_t ans      = hi;
        if (tid == 0) sLb[nb] = hi;
        if (hi - lo > kCountMaxChunks * uint32_t(kCountTileKeys) - E)
        {
            if (mine) ans = lo + lowerBound(keys + lo, hi - lo, v);
        }
        else
        {
            bool open = mine;
            for (uint32_t start = lo & ~(E - 1u); start < hi; start += uint32_t(kCountTileKeys))
            {
                const uint32_t len  = min(uint32_t(kCountTileKeys), hi - start);
                const uint32_t want = min((len + E - 1u) & ~(E - 1u), numKeysAligned - min(start, numKeysAligned));
                if (tid == 0 && want)
                {
                    fenceProxyAsync();
                    mbarExpectTx(&sBar, want * uint32_t(sizeof(K)));
                    bulkCopyToShared(sKeys, keys + start, want * uint32_t(sizeof(K)), &sBar);
                }
                if (want + tid < len) sKeys[want + tid] = keys[start + want + tid]; // < E keys behind the last 16-byte unit
                if (want)
                {
                    mbarWait(&sBar, phase);
                    phase ^= 1u;
                }
                __syncthreads();
                if (open)
                {
                    const uint32_t off = start < lo ? lo - start : 0u;
                    const uint32_t pos = off + lowerBound(sKeys + off, len - off, v);
                    if (pos < len) ans = start + pos, open = false;
                }
                __syncthreads();
            }
        }
        if (mine) sLb[tid] = ans;
        __syncthreads();
        if (mine) counts[blockBase + int(tid)] = min(sLb[tid + 1] - sLb[tid], maxCount);
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
    unsigned grid = gridFor(size_t(numNodes), kCountThreads, 4);
    if ((reinterpret_cast<uintptr_t>(first) & 15u) == 0)
        nodeCountsTmaKernel<K><<<grid, kCountThreads, 0, s>>>(tree, counts, numNodes, first, uint32_t(numKeys), maxCount);
    else nodeCountsKernel<K><<<grid, kCountThreads, 0, s>>>(tree, counts, numNodes, first, uint32_t(numKeys), maxCount);
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
