/*
 * focus.cu — a19 (first part): the device kernels the locally essential ("focus") tree update is made of.
 *
 * Replaces the GPU entry points of focus/rebalance_gpu.h, traversal/collisions_gpu.h (markMacsGpu) and
 * focus/source_center_gpu.h (geoMacSpheresGpu, setMacGpu, moveCenters); semantics follow the CPU functions
 *   rebalanceDecisionEssential / mergeCountAndMacOp   [ref: focus/rebalance.hpp:49-166]
 *   macRefineOp                                       [ref: focus/rebalance.hpp:91-98, focus/rebalance_gpu.cu:93-118]
 *   protectAncestors / nzAncestorOp                   [ref: focus/rebalance.hpp:113-181]
 *   enforceKeys / enforceKeySingle                    [ref: focus/rebalance.hpp:196-266], containingNode tree/octree.hpp:244-261
 *   rangeCount                                        [ref: focus/rebalance.hpp:277-299]
 *   markMacs / markMacPerBox / evaluateMacPbc         [ref: traversal/macs.hpp:129-141,196-269]
 *   computeMinMacR2 / computeVecMacR2                 [ref: traversal/macs.hpp:58-100]
 * The orchestration around them (FocusedOctree, treelet exchange between peers) is not part of this file.
 *
 * All of them are one thread per node / leaf / key; the tree arrays of a focus tree (10^5..10^7 nodes) are streamed
 * once, the MAC marking walks the linked tree depth first like findHalos.
 */
#include "halo_box.cuh"

namespace csb
{

constexpr int kFocusThreads = 256;

__device__ inline bool overlapKeyRanges(uint64_t a, uint64_t b, uint64_t c, uint64_t d) { return b > c && d > a; }

// ------------------------------------------------------------------------------------------------ rebalance decision
template<class K>
__global__ void __launch_bounds__(kFocusThreads)
    rebalanceEssentialKernel(const K* __restrict__ prefixes, const int* __restrict__ childOffsets,
                             const int* __restrict__ parents, const uint32_t* __restrict__ counts,
                             const char* __restrict__ macs, K focusStart, K focusEnd, uint32_t bucket,
                             int* __restrict__ nodeOps, int numNodes)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numNodes; i += gridDim.x * blockDim.x)
    {
        const K nodeKey      = prefixes[i];
        const unsigned level = decodePrefixLength<K>(nodeKey) / 3u;
        int op               = 1;
        bool merged          = false;
        if (i)
        {
            const int parent      = parents[(i - 1) / 8];
            const bool countMerge = counts[parent] <= bucket;
            const bool macMerge   = macs[parent] == 0;
            const K firstGroupKey = decodePlaceholderBit<K>(prefixes[parent]);
            const K lastGroupKey  = firstGroupKey + K(8) * nodeRange<K>(level);
            // a sibling group that touches the focus is never removed on the MAC alone
            const bool inFringe = overlapKeyRanges(firstGroupKey, lastGroupKey, focusStart, focusEnd);
            merged              = countMerge || (macMerge && !inFringe);
        }
        if (merged) op = 0;
        else
        {
            const K nodeStart  = decodePlaceholderBit<K>(nodeKey);
            const bool isLeaf  = childOffsets[i] == 0;
            const bool inFocus = nodeStart >= focusStart && nodeStart < focusEnd;
            if (isLeaf && level < KeyTraits<K>::maxLevel && counts[i] > bucket && (macs[i] || inFocus)) op = 8;
        }
        nodeOps[i] = op;
    }
}

template<class K>
__global__ void __launch_bounds__(kFocusThreads)
    macRefineKernel(const K* __restrict__ prefixes, const char* __restrict__ macs, const int* __restrict__ l2i,
                    int numLeaves, int focusFirst, int focusLast, int* __restrict__ nodeOps)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numLeaves; i += gridDim.x * blockDim.x)
    {
        int op = 1;
        if (i < focusFirst || i >= focusLast)
        {
            const int n          = l2i[i];
            const unsigned level = decodePrefixLength<K>(prefixes[n]) / 3u;
            if (level < KeyTraits<K>::maxLevel && macs[n]) op = 8;
        }
        nodeOps[i] = op;
    }
}

/* nodeOps[i] = op of the closest ancestor-or-self with a non-zero op if i is its left-most descendant, else 0.
 * In place like the reference: a node whose op is rewritten from 0 to its ancestor's op gives its own descendants the
 * same answer the ancestor would (focus/rebalance_gpu.cu:134-136). */
template<class K>
__global__ void __launch_bounds__(kFocusThreads)
    protectAncestorsKernel(const K* __restrict__ prefixes, const int* __restrict__ parents, int* nodeOps, int numNodes,
                           int* __restrict__ numChanged)
{
    int changed = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numNodes; i += gridDim.x * blockDim.x)
    {
        int op;
        if (i == 0) op = nodeOps[0];
        else
        {
            int a = i;
            int v;
            while ((v = reinterpret_cast<volatile int*>(nodeOps)[a]) == 0)
                a = parents[(a - 1) / 8];
            op = decodePlaceholderBit<K>(prefixes[i]) == decodePlaceholderBit<K>(prefixes[a]) ? v : 0;
        }
        changed += op != 1;
        nodeOps[i] = op;
    }
    changed = __reduce_add_sync(0xffffffffu, changed);
    if ((threadIdx.x & 31) == 0 && changed) atomicAdd(numChanged, changed);
}

// ------------------------------------------------------------------------------------------------ mandatory keys
//! placeholder-bit key of the biggest node that starts at `a`  [ref: sfc/common.hpp:339-356 lastNzPlace / makePrefix]
template<class K>
__device__ inline int lastNzPlaceDev(K key)
{
    if (key == 0) return int(KeyTraits<K>::maxLevel);
    const int tz = (sizeof(K) == 8) ? (__ffsll((long long)key) - 1) : (__ffs(int(key)) - 1);
    return int(KeyTraits<K>::maxLevel) - tz / 3;
}

//! smallest node of the linked tree that contains the node `nodeKey`  [ref: tree/octree.hpp:244-261]
template<class K>
__device__ inline int containingNode(K nodeKey, const K* __restrict__ prefixes, const int* __restrict__ childOffsets)
{
    const int nodeLevel = int(decodePrefixLength<K>(nodeKey)) / 3;
    const K key         = decodePlaceholderBit<K>(nodeKey);
    int ret             = 0;
    for (int i = 1; i <= nodeLevel; ++i)
    {
        if (childOffsets[ret] == 0 || nodeKey == prefixes[ret]) break;
        ret = childOffsets[ret] + int(octalDigit<K>(key, unsigned(i)));
    }
    return ret;
}

/* One thread per mandatory key.  The updates of nodeOps commute (0 -> 1 stores and a maximum), so the result equals
 * the sequential CPU loop whatever the interleaving; status = maximum over the keys. */
template<class K>
__global__ void __launch_bounds__(kFocusThreads)
    enforceKeysKernel(const K* __restrict__ forcedKeys, int numKeys, const K* __restrict__ prefixes,
                      const int* __restrict__ childOffsets, const int* __restrict__ parents, int* nodeOps,
                      int* __restrict__ status)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < numKeys; k += gridDim.x * blockDim.x)
    {
        const K key = forcedKeys[k];
        if (key == 0 || key == nodeRange<K>(0)) continue; // converged
        int st                  = 0;
        const int keyPos        = lastNzPlaceDev<K>(key);
        const K nodeKeyWant     = encodePlaceholderBit<K>(key, 3 * keyPos);
        const int nodeIdx       = containingNode<K>(nodeKeyWant, prefixes, childOffsets);
        const K nodeKeyHave     = prefixes[nodeIdx];
        const int nodeLevelHave = int(decodePrefixLength<K>(nodeKeyHave)) / 3;
        const bool trySplit     = nodeKeyHave != nodeKeyWant && nodeLevelHave < int(KeyTraits<K>::maxLevel);
        const bool undoMerges   = reinterpret_cast<volatile int*>(nodeOps)[nodeIdx] == 0 || trySplit;
        if (undoMerges && nodeIdx > 0)
        {
            st         = 1; // cancelMerge
            int parent = nodeIdx;
            do
            {
                parent          = parents[(parent - 1) / 8];
                const int first = childOffsets[parent];
                for (int i = first; i < first + 8; ++i)
                    atomicCAS(&nodeOps[i], 0, 1);
            } while (parent != 0);
        }
        if (trySplit)
        {
            int levelDiff = keyPos - nodeLevelHave;
            st            = levelDiff > 1 ? 3 : 2; // failed : rebalance (only one level is added per step)
            levelDiff     = min(levelDiff, 1);
            atomicMax(&nodeOps[nodeIdx], 1 << (3 * levelDiff));
        }
        if (st) atomicMax(status, st);
    }
}

// ------------------------------------------------------------------------------------------------ rangeCount
template<class K>
__global__ void __launch_bounds__(kFocusThreads)
    rangeCountKernel(const K* __restrict__ leaves, int numLeafKeys, const uint32_t* __restrict__ counts,
                     const K* __restrict__ leavesFocus, const int* __restrict__ focusIdx, int numIdx,
                     uint32_t* __restrict__ countsFocus)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numIdx; i += gridDim.x * blockDim.x)
    {
        const int leafIdx = focusIdx[i];
        const K startKey = leavesFocus[leafIdx], endKey = leavesFocus[leafIdx + 1];
        // findNodeBelow = upper_bound - 1, findNodeAbove = lower_bound  [ref: tree/csarray.hpp:78-90]
        int lo = 0, len = numLeafKeys;
        while (len > 0)
        {
            const int half = len >> 1;
            const bool le  = !(startKey < leaves[lo + half]);
            lo             = le ? lo + half + 1 : lo;
            len            = le ? len - half - 1 : half;
        }
        const int startIdx = lo - 1;
        lo = 0, len = numLeafKeys;
        while (len > 0)
        {
            const int half = len >> 1;
            const bool lt  = leaves[lo + half] < endKey;
            lo             = lt ? lo + half + 1 : lo;
            len            = lt ? len - half - 1 : half;
        }
        const int endIdx       = lo;
        unsigned long long sum = 0;
        for (int j = startIdx; j < endIdx; ++j)
            sum += counts[j];
        countsFocus[leafIdx] = uint32_t(min(sum, 0xFFFFFFFFull));
    }
}

// ------------------------------------------------------------------------------------------------ MAC spheres
//! geometric centre and half size of the node with the given placeholder-bit key  [ref: sfc/box.hpp:334-351]
template<class K, class T>
__device__ inline void nodeCenterSize(int kind, const uint16_t* dec, K prefix, const DBox<T>& box, T* c, T* s)
{
    constexpr int maxCoord = 1 << KeyTraits<K>::maxLevel;
    const T uL             = T(1.) / T(maxCoord);
    const unsigned level   = decodePrefixLength<K>(prefix) / 3u;
    const IBox b           = sfcIBox<K>(kind, dec, decodePlaceholderBit<K>(prefix), level);
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        const T halfUnit = T(0.5) * uL * box.len[d];
        c[d]             = box.lo[d] + T(b.hi[d] + b.lo[d]) * halfUnit;
        s[d]             = T(b.hi[d] - b.lo[d]) * halfUnit;
    }
}

//! centers[i] = (geometric centre, (2 max(size) invThetaEff)^2)  [ref: traversal/macs.hpp:58-71 computeMinMacR2]
template<class K, class T>
__global__ void __launch_bounds__(kFocusThreads)
    geoMacSpheresKernel(int kind, const uint16_t* __restrict__ dec, const K* __restrict__ prefixes, int numNodes,
                        T* __restrict__ centers, float invTheta, DBox<T> box)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numNodes; i += gridDim.x * blockDim.x)
    {
        T c[3], s[3];
        nodeCenterSize<K, T>(kind, dec, prefixes[i], box, c, s);
        const T l   = T(2) * max(s[0], max(s[1], s[2]));
        const T mac = l * invTheta;
        centers[4 * i] = c[0], centers[4 * i + 1] = c[1], centers[4 * i + 2] = c[2], centers[4 * i + 3] = mac * mac;
    }
}

//! macSpheres[i][3] = (2 max(size) invTheta + |expansion centre - geometric centre|)^2, 0 stays 0
//! [ref: traversal/macs.hpp:84-100 computeVecMacR2, focus/source_center_gpu.cu setMacKernel]
template<class K, class T>
__global__ void __launch_bounds__(kFocusThreads)
    setMacKernel(int kind, const uint16_t* __restrict__ dec, const K* __restrict__ prefixes, int numNodes,
                 T* __restrict__ spheres, float invTheta, DBox<T> box)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numNodes; i += gridDim.x * blockDim.x)
    {
        T c[3], s[3];
        nodeCenterSize<K, T>(kind, dec, prefixes[i], box, c, s);
        const T dx = spheres[4 * i] - c[0], dy = spheres[4 * i + 1] - c[1], dz = spheres[4 * i + 2] - c[2];
        const T dist = sqrt(dx * dx + (dy * dy + dz * dz)); // norm2 is a right fold [ref: util/array.hpp:253-256]
        const T l    = T(2.0) * max(s[0], max(s[1], s[2]));
        const T mac  = l * invTheta + dist;
        spheres[4 * i + 3] = (spheres[4 * i + 3] != T(0)) ? mac * mac : T(0);
    }
}

template<class T>
__global__ void __launch_bounds__(kFocusThreads) moveCentersKernel(const T* __restrict__ src, int numNodes, T* __restrict__ dst)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numNodes; i += gridDim.x * blockDim.x)
    {
        dst[4 * i] = src[3 * i], dst[4 * i + 1] = src[3 * i + 1], dst[4 * i + 2] = src[3 * i + 2], dst[4 * i + 3] = T(1.0);
    }
}

// ------------------------------------------------------------------------------------------------ markMacs
/* One BLOCK (4 warps) per 32 consecutive focus leaves, one lane per leaf.  The reference walks the tree once per focus
 * leaf [traversal/macs.hpp:196-269]; neighbouring leaves open almost the same nodes, so here the UNION of the 32 walks
 * is walked once, with stack entries (first child, lane mask): a lane takes part in the test of a node only if it
 * failed the MAC at every ancestor, exactly as in its own walk, so the set of marked nodes is the reference's.
 * A leaf near the focus surface opens ~2000 nodes, a long serial chain (ncu of the one-warp version: 10 % of the warp
 * slots active, SMs idle 42 % of the time): the warps of the block share the walk through a small pool in shared
 * memory — a busy warp hands surplus stack entries to the pool while it is low, an idle warp takes one, the round ends
 * when the pool is empty and no warp is busy (counter under a shared-memory lock).
 * The eight children of a node are fetched together (uniform addresses, one round trip per opened node: prefixes,
 * centres and child offsets of the eight siblings; the child offset rides on the stack).  Nodes fully inside the focus
 * end the descent; every node that fails the MAC w.r.t. a leaf (evaluateMacPbc, exact expression, no contraction) is
 * marked and opened.  Private stack depth <= 7 per level + 8.
 * Measured (32 M particles, focus = 262144 leaves, profiles/r01h): thread per leaf (the reference's scheme) 1.34 ms,
 * one warp per 32 leaves 0.99 ms (without the batched loads 2.10), leaves with work packed densely per block first
 * 2.02 ms, subtrees of level 3 dealt statically to 8 warps 1.31 ms, this version 0.63 ms with 8 warps per block and
 * 0.58 ms with 4; 0.50 ms with one block per group for the hardware scheduler instead of a grid-stride loop over a
 * capped grid (the groups with long walks are spread unevenly). */
constexpr int kMarkThreads = 128; // the warps of a block share the walk of one group of 32 focus leaves
constexpr int kMarkWarps   = kMarkThreads / 32;
constexpr int kMarkStack   = 160;
constexpr int kMarkPool    = 64;  // shared pool of open nodes, filled by busy warps, drained by idle ones

template<class T>
struct MarkShared
{
    int2 stack[kMarkWarps][kMarkStack];
    int2 pool[kMarkPool];
    T tc[3][32], ts[3][32];
    unsigned maxLevel[32];
    unsigned aliveMask;
    int lock, poolCount, active;
};

__device__ inline void markLock(int* lock)
{
    while (atomicCAS(lock, 0, 1) != 0)
        __nanosleep(64);
    __threadfence_block();
}
__device__ inline void markUnlock(int* lock)
{
    __threadfence_block();
    atomicExch(lock, 0);
}

template<class K, class T>
__global__ void __launch_bounds__(kMarkThreads)
    markMacsKernel(int kind, const uint16_t* __restrict__ enc, const uint16_t* __restrict__ dec,
                   const K* __restrict__ prefixes, const int* __restrict__ childOffsets, const T* __restrict__ centers,
                   DBox<T> box, const K* __restrict__ focusNodes, int numFocusNodes, int limitSource, char* markings)
{
    __shared__ MarkShared<T> sh;
    const int warp      = int(threadIdx.x >> 5);
    const unsigned lane = laneId();
    int2* stack         = sh.stack[warp];

    constexpr int maxCoord = 1 << KeyTraits<K>::maxLevel;
    const T uL             = T(1.) / T(maxCoord);
    const K focusStart = focusNodes[0], focusEnd = focusNodes[numFocusNodes];
    for (int base = int(blockIdx.x) * 32; base < numFocusNodes; base += int(gridDim.x) * 32)
    {
        // ---- warp 0: the 32 targets of this block round (centre, size, deepest source level, has work at all)
        __syncthreads();
        if (warp == 0)
        {
            const int i = base + int(lane);
            bool alive  = i < numFocusNodes;
            T c[3] = {0, 0, 0}, sz[3] = {0, 0, 0};
            unsigned ml = KeyTraits<K>::maxLevel;
            if (alive)
            {
                const K a = focusNodes[i], b = focusNodes[i + 1];
                const unsigned tl = treeLevel<K>(K(b - a));
                const IBox target = sfcIBox<K>(kind, dec, a, tl);
                IBox ext;
#pragma unroll
                for (int d = 0; d < 3; ++d)
                    ext.lo[d] = target.lo[d] - 1, ext.hi[d] = target.hi[d] + 1;
                alive = !containedIn<K>(kind, enc, focusStart, focusEnd, ext);
#pragma unroll
                for (int d = 0; d < 3; ++d)
                {
                    const T halfUnit = T(0.5) * uL * box.len[d];
                    c[d]             = box.lo[d] + T(target.hi[d] + target.lo[d]) * halfUnit;
                    sz[d]            = T(target.hi[d] - target.lo[d]) * halfUnit;
                }
                if (limitSource) ml = unsigned(max(int(tl) - 1, 0));
            }
#pragma unroll
            for (int d = 0; d < 3; ++d)
                sh.tc[d][lane] = c[d], sh.ts[d][lane] = sz[d];
            sh.maxLevel[lane]    = ml;
            const unsigned amask = __ballot_sync(0xffffffffu, alive);
            if (lane == 0) sh.aliveMask = amask, sh.lock = 0, sh.poolCount = 0, sh.active = 1;
        }
        __syncthreads();
        const unsigned aliveMask = sh.aliveMask;
        if (aliveMask == 0) continue; // block-uniform
        const T tc[3] = {sh.tc[0][lane], sh.tc[1][lane], sh.tc[2][lane]};
        const T ts[3] = {sh.ts[0][lane], sh.ts[1][lane], sh.ts[2][lane]};
        const unsigned maxLevel = sh.maxLevel[lane];

        // lanes of `mask` test a node given by its prefix, centre and MAC; returns the lanes that fail the MAC
        auto check = [&](K prefix, T cx, T cy, T cz, T mac, unsigned mask) -> unsigned
        {
            const unsigned level = decodePrefixLength<K>(prefix) / 3u;
            const K nodeStart    = decodePlaceholderBit<K>(prefix);
            const K nodeEnd      = nodeStart + nodeRange<K>(level);
            if (!(nodeStart < focusStart || nodeEnd > focusEnd)) return 0u;
            const T c[3] = {cx, cy, cz};
            T dd[3];
#pragma unroll
            for (int d = 0; d < 3; ++d)
            {
                T dx = tc[d] - c[d];
                if (box.pbc[d]) dx -= box.len[d] * rint(dx * box.ilen[d]); // open axis: the fold term is an exact 0
                dx = fabs(dx);
                dx -= ts[d];
                dx += fabs(dx);
                dx *= T(0.5);
                dd[d] = dx;
            }
            const T r2          = dd[0] * dd[0] + (dd[1] * dd[1] + dd[2] * dd[2]);
            const bool violates = ((mask >> lane) & 1u) && r2 < fabs(mac) && level <= maxLevel;
            return __ballot_sync(0xffffffffu, violates);
        };

        // ---- warp 0 opens the root; the other warps start idle and take work from the pool
        int sp        = 0;
        bool busy     = false;
        if (warp == 0)
        {
            const unsigned rootFail = check(prefixes[0], centers[0], centers[1], centers[2], centers[3], aliveMask);
            const int first         = childOffsets[0];
            if (rootFail && lane == 0 && !markings[0]) markings[0] = 1;
            if (rootFail && first != 0)
            {
                if (lane == 0) stack[0] = make_int2(first, int(rootFail));
                sp   = 1;
                busy = true;
            }
            if (!busy)
            {
                if (lane == 0)
                {
                    markLock(&sh.lock);
                    sh.active = 0;
                    markUnlock(&sh.lock);
                }
            }
        }
        __syncwarp();
        while (true)
        {
            if (sp == 0)
            {
                // out of private work: take an entry from the pool, or leave when nobody can produce any more
                int2 e   = make_int2(0, 0);
                int state = 0; // 0 retry, 1 got work, 2 done
                if (lane == 0)
                {
                    markLock(&sh.lock);
                    if (busy) sh.active -= 1;
                    if (sh.poolCount > 0)
                    {
                        e = sh.pool[--sh.poolCount];
                        sh.active += 1;
                        state = 1;
                    }
                    else if (sh.active == 0) { state = 2; }
                    markUnlock(&sh.lock);
                }
                busy  = false;
                state = __shfl_sync(0xffffffffu, state, 0);
                if (state == 2) break;
                if (state == 0)
                {
                    __nanosleep(200);
                    continue;
                }
                e.x = __shfl_sync(0xffffffffu, e.x, 0), e.y = __shfl_sync(0xffffffffu, e.y, 0);
                if (lane == 0) stack[0] = e;
                sp   = 1;
                busy = true;
                __syncwarp();
            }
            const int2 e        = stack[--sp];
            __syncwarp();
            const int first     = e.x;
            const unsigned mask = unsigned(e.y);
#ifdef CSB_MARK_DEBUG
            if (first <= 0 || first > (1 << 26))
            {
                if (lane == 0) printf("markMacs bad entry: block %d warp %d sp %d first %d mask %08x pool %d active %d\n", blockIdx.x, warp, sp, first, mask, sh.poolCount, sh.active);
                break;
            }
#endif
            // the eight children: all loads first (uniform addresses, independent), then the tests
            K cp[8];
            T cc[8][4];
            int co[8];
#pragma unroll
            for (int o = 0; o < 8; ++o)
            {
                cp[o] = prefixes[first + o];
                co[o] = childOffsets[first + o];
#pragma unroll
                for (int d = 0; d < 4; ++d)
                    cc[o][d] = centers[4 * (first + o) + d];
            }
#pragma unroll
            for (int o = 0; o < 8; ++o)
            {
                const unsigned fail = check(cp[o], cc[o][0], cc[o][1], cc[o][2], cc[o][3], mask);
                if (fail)
                {
                    if (lane == 0 && !markings[first + o]) markings[first + o] = 1;
                    if (co[o] != 0)
                    {
                        if (lane == 0) stack[sp] = make_int2(co[o], int(fail));
                        ++sp;
                    }
                }
            }
            __syncwarp();
            // share: keep one entry, hand the rest of the surplus to idle warps while the pool is low
            // (the pool level is read by lane 0 and broadcast: the lanes must agree on the branch)
            const int poolNow = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile int*>(&sh.poolCount), 0);
            if (sp >= 2 && poolNow < kMarkWarps)
            {
                int give = 0;
                if (lane == 0)
                {
                    markLock(&sh.lock);
                    give = min(sp - 1, kMarkPool - sh.poolCount);
                    give = min(give, 2 * kMarkWarps - sh.poolCount);
                    give = max(give, 0);
                    for (int k = 0; k < give; ++k)
                        sh.pool[sh.poolCount++] = stack[sp - 1 - k]; // from the top: only lane 0 touches the stack
                    markUnlock(&sh.lock);
                }
                give = __shfl_sync(0xffffffffu, give, 0);
                sp -= give;
                __syncwarp();
            }
        }
    }
}

} // namespace csb

extern "C" {

using namespace csb;

#define CSB_FOCUS_API(KS, K)                                                                                           \
    int cstone_b200_rebalance_decision_essential_##KS(const K* prefixes, const int32_t* childOffsets,                  \
                                                      const int32_t* parents, const uint32_t* counts,                  \
                                                      const char* macs, K focusStart, K focusEnd,                      \
                                                      uint32_t bucketSize, int32_t* nodeOps, int32_t numNodes,         \
                                                      void* stream)                                                    \
    {                                                                                                                  \
        if (numNodes <= 0) return 0;                                                                                   \
        if (!prefixes || !childOffsets || !counts || !macs || !nodeOps || (numNodes > 1 && !parents))                  \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        rebalanceEssentialKernel<K><<<gridFor(size_t(numNodes), kFocusThreads, 8), kFocusThreads, 0,                   \
                                      asStream(stream)>>>(prefixes, childOffsets, parents, counts, macs, focusStart,   \
                                                          focusEnd, bucketSize, nodeOps, numNodes);                    \
        CSB_LAUNCHED();                                                                                                \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_mac_refine_decision_##KS(const K* prefixes, const char* macs, const int32_t* l2i,                  \
                                             int32_t numLeafNodes, int32_t focusFirst, int32_t focusLast,              \
                                             int32_t* nodeOps, void* stream)                                           \
    {                                                                                                                  \
        if (numLeafNodes <= 0) return 0;                                                                               \
        if (!prefixes || !macs || !l2i || !nodeOps) return CSTONE_E_INVALID_ARGUMENT;                                  \
        macRefineKernel<K><<<gridFor(size_t(numLeafNodes), kFocusThreads, 8), kFocusThreads, 0, asStream(stream)>>>(   \
            prefixes, macs, l2i, numLeafNodes, focusFirst, focusLast, nodeOps);                                        \
        CSB_LAUNCHED();                                                                                                \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_protect_ancestors_##KS(const K* prefixes, const int32_t* parents, int32_t* nodeOps,                \
                                           int32_t numNodes, int32_t* convergedHost, void* temp, void* stream)         \
    {                                                                                                                  \
        if (!convergedHost) return CSTONE_E_INVALID_ARGUMENT;                                                          \
        *convergedHost = 1;                                                                                            \
        if (numNodes <= 0) return 0;                                                                                   \
        if (!prefixes || !nodeOps || !temp || (numNodes > 1 && !parents)) return CSTONE_E_INVALID_ARGUMENT;            \
        cudaStream_t s = asStream(stream);                                                                             \
        CSB_CUDA(cudaMemsetAsync(temp, 0, 4, s));                                                                      \
        protectAncestorsKernel<K><<<gridFor(size_t(numNodes), kFocusThreads, 8), kFocusThreads, 0, s>>>(               \
            prefixes, parents, nodeOps, numNodes, static_cast<int*>(temp));                                            \
        CSB_LAUNCHED();                                                                                                \
        int changed = 0;                                                                                               \
        CSB_CUDA(cudaMemcpyAsync(&changed, temp, 4, cudaMemcpyDeviceToHost, s));                                       \
        CSB_CUDA(cudaStreamSynchronize(s));                                                                            \
        *convergedHost = changed == 0;                                                                                 \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_enforce_keys_##KS(const K* forcedKeys, int32_t numForcedKeys, const K* nodeKeys,                   \
                                      const int32_t* childOffsets, const int32_t* parents, int32_t* nodeOps,           \
                                      int32_t* statusHost, void* temp, void* stream)                                   \
    {                                                                                                                  \
        if (!statusHost) return CSTONE_E_INVALID_ARGUMENT;                                                             \
        *statusHost = 0;                                                                                               \
        if (numForcedKeys <= 0) return 0;                                                                              \
        if (!forcedKeys || !nodeKeys || !childOffsets || !parents || !nodeOps || !temp)                                \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        cudaStream_t s = asStream(stream);                                                                             \
        CSB_CUDA(cudaMemsetAsync(temp, 0, 4, s));                                                                      \
        enforceKeysKernel<K><<<gridFor(size_t(numForcedKeys), kFocusThreads, 8), kFocusThreads, 0, s>>>(               \
            forcedKeys, numForcedKeys, nodeKeys, childOffsets, parents, nodeOps, static_cast<int*>(temp));             \
        CSB_LAUNCHED();                                                                                                \
        CSB_CUDA(cudaMemcpyAsync(statusHost, temp, 4, cudaMemcpyDeviceToHost, s));                                     \
        CSB_CUDA(cudaStreamSynchronize(s));                                                                            \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_range_count_##KS(const K* leaves, int32_t numLeafKeys, const uint32_t* counts,                     \
                                     const K* leavesFocus, const int32_t* leavesFocusIdx, int32_t numIdx,              \
                                     uint32_t* countsFocus, void* stream)                                              \
    {                                                                                                                  \
        if (numIdx <= 0) return 0;                                                                                     \
        if (!leaves || !counts || !leavesFocus || !leavesFocusIdx || !countsFocus || numLeafKeys < 2)                  \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        rangeCountKernel<K><<<gridFor(size_t(numIdx), kFocusThreads, 8), kFocusThreads, 0, asStream(stream)>>>(        \
            leaves, numLeafKeys, counts, leavesFocus, leavesFocusIdx, numIdx, countsFocus);                            \
        CSB_LAUNCHED();                                                                                                \
        return 0;                                                                                                      \
    }
CSB_FOCUS_API(k32, uint32_t)
CSB_FOCUS_API(k64, uint64_t)

#define CSB_MAC_API(KS, K, TS, T)                                                                                      \
    int cstone_b200_geo_mac_spheres_##KS##_##TS(int kind, const K* prefixes, int32_t numNodes, T* centers4,            \
                                                float invTheta, const cstone_box_##TS* box, void* stream)              \
    {                                                                                                                  \
        if (numNodes <= 0) return 0;                                                                                   \
        if (!prefixes || !centers4 || !box || (kind != CSTONE_SFC_MORTON && kind != CSTONE_SFC_HILBERT))               \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        SfcTables tabs;                                                                                                \
        int st = getSfcTables(&tabs);                                                                                  \
        if (st) return st;                                                                                             \
        geoMacSpheresKernel<K, T><<<gridFor(size_t(numNodes), kFocusThreads, 8), kFocusThreads, 0,                     \
                                    asStream(stream)>>>(kind, tabs.dec, prefixes, numNodes, centers4, invTheta,        \
                                                        makeDBox<T>(box));                                             \
        CSB_LAUNCHED();                                                                                                \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_set_mac_##KS##_##TS(int kind, const K* prefixes, int32_t numNodes, T* macSpheres4, float invTheta, \
                                        const cstone_box_##TS* box, void* stream)                                      \
    {                                                                                                                  \
        if (numNodes <= 0) return 0;                                                                                   \
        if (!prefixes || !macSpheres4 || !box || (kind != CSTONE_SFC_MORTON && kind != CSTONE_SFC_HILBERT))            \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        SfcTables tabs;                                                                                                \
        int st = getSfcTables(&tabs);                                                                                  \
        if (st) return st;                                                                                             \
        setMacKernel<K, T><<<gridFor(size_t(numNodes), kFocusThreads, 8), kFocusThreads, 0, asStream(stream)>>>(       \
            kind, tabs.dec, prefixes, numNodes, macSpheres4, invTheta, makeDBox<T>(box));                              \
        CSB_LAUNCHED();                                                                                                \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_b200_mark_macs_##KS##_##TS(int kind, const K* prefixes, const int32_t* childOffsets,                    \
                                          const T* centers4, const cstone_box_##TS* box, const K* focusNodes,          \
                                          int32_t numFocusNodes, int limitSource, char* markings, void* stream)        \
    {                                                                                                                  \
        if (numFocusNodes <= 0) return 0;                                                                              \
        if (!prefixes || !childOffsets || !centers4 || !box || !focusNodes || !markings ||                             \
            (kind != CSTONE_SFC_MORTON && kind != CSTONE_SFC_HILBERT))                                                 \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        SfcTables tabs;                                                                                                \
        int st = getSfcTables(&tabs);                                                                                  \
        if (st) return st;                                                                                             \
        /* one block per 32 focus leaves, uncapped: groups without work end at once, the hardware scheduler balances */ \
        markMacsKernel<K, T><<<unsigned((numFocusNodes + 31) / 32), kMarkThreads, 0, asStream(stream)>>>(              \
            kind, tabs.enc, tabs.dec, prefixes, childOffsets, centers4, makeDBox<T>(box), focusNodes, numFocusNodes,   \
            limitSource, markings);                                                                                    \
        CSB_LAUNCHED();                                                                                                \
        return 0;                                                                                                      \
    }
CSB_MAC_API(k32, uint32_t, f32, float)
CSB_MAC_API(k32, uint32_t, f64, double)
CSB_MAC_API(k64, uint64_t, f32, float)
CSB_MAC_API(k64, uint64_t, f64, double)

int cstone_b200_move_centers_f32(const float* src3, int32_t numNodes, float* dst4, void* stream)
{
    if (numNodes <= 0) return 0;
    if (!src3 || !dst4) return CSTONE_E_INVALID_ARGUMENT;
    moveCentersKernel<float><<<gridFor(size_t(numNodes), kFocusThreads, 8), kFocusThreads, 0, asStream(stream)>>>(
        src3, numNodes, dst4);
    CSB_LAUNCHED();
    return 0;
}
int cstone_b200_move_centers_f64(const double* src3, int32_t numNodes, double* dst4, void* stream)
{
    if (numNodes <= 0) return 0;
    if (!src3 || !dst4) return CSTONE_E_INVALID_ARGUMENT;
    moveCentersKernel<double><<<gridFor(size_t(numNodes), kFocusThreads, 8), kFocusThreads, 0, asStream(stream)>>>(
        src3, numNodes, dst4);
    CSB_LAUNCHED();
    return 0;
}

} // extern "C"
