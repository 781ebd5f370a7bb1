/*
 * let_tree.cuh -- a19 (third part): the multi-rank focus tree (locally essential tree, LET) of the Domain engine.
 *
 * What the reference's FocusedOctree<KeyType, T, Accelerator> does on P ranks [ref: focus/octree_focus_mpi.hpp:51-734]:
 * every rank keeps ONE octree that is fully resolved (bucketSize) inside its own SFC range, the focus, and only as
 * deep as a minimum-distance MAC w.r.t. the focus demands elsewhere.  Per Domain::sync
 *
 *   updateMinMac   MAC spheres l (1/theta + 1/2) around the node centres, marks w.r.t. the focus leaves   [:457-525]
 *   updateTree     handover of tree parts whose SFC range moved to a neighbour rank (focusTransfer), node ops from
 *                  counts inside / MACs outside the focus, rank boundaries of the peers enforced, rebalance, MAC
 *                  refinement of newly acquired range, then the peers agree on the nodes they hold of each other's
 *                  range (treelets: keys the owner does not have are rejected and removed)             [:108-187]
 *   updateCounts   own particles counted per leaf; leaves of non-peer ranks take sums of the replicated global tree;
 *                  leaves in a peer's range take the owner's node counts (peer exchange); upsweeps      [:205-273]
 *
 * and Halos::discover + computeLayout run on that tree [ref: halos/halos.hpp:127-223].
 *
 * Here: the tree itself, its node ops, rebalance, linking, MAC marks, counts, upsweeps and the peer gather / scatter
 * are this library's kernels on device arrays (class FocusTree, focus_tree.cuh).  The steps that decide WHICH keys
 * travel between ranks work on a host copy of the leaf keys, as in the reference's GPU build, which keeps leaves_ and
 * treeData_ on the host for the same purpose [ref: octree_focus_mpi.hpp:144-168 syncTreeletsGpu + downloadOctree].
 * Messages of a size only the sender knows (the reference probes them with MPI_Probe) are preceded by one all-gather
 * of the sizes (exchangeBytes).
 */
#pragma once

#include <algorithm>
#include <array>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "comm.hpp"
#include "domain_host.hpp"
#include "focus_tree.cuh"

namespace csb
{

//! index of the last leaf that starts at or below key  [ref: tree/csarray.hpp:78-83 findNodeBelow]
template<class K>
inline int findNodeBelow(const K* leaves, int numLeafKeys, K key)
{
    return int(std::upper_bound(leaves, leaves + numLeafKeys, key) - leaves) - 1;
}

//! findNodeAbove for a run of ascending keys: the search starts at the previous hit and gallops forward, so that a
//! treelet of m keys is matched against n leaves in O(m log(n / m)) steps that stay in cache
template<class K>
inline int findNodeAboveFrom(const K* leaves, int numLeafKeys, int from, K key)
{
    int lo = std::min(from, numLeafKeys), step = 1;
    while (lo + step < numLeafKeys && leaves[lo + step] < key)
    {
        lo += step;
        step *= 2;
    }
    const int hi = std::min(lo + step + 1, numLeafKeys);
    return int(std::lower_bound(leaves + lo, leaves + hi, key) - leaves);
}

// ------------------------------------------------------------------------------------------------
// the replicated global tree, linked on the host (a few thousand nodes) for the peer search
// ------------------------------------------------------------------------------------------------
template<class K>
struct HostOctree
{
    struct Node
    {
        K start;
        unsigned level;
        int child; // index of the first of 8 consecutive children, 0 for a leaf
    };
    std::vector<Node> nodes;

    //! nodes in breadth-first order, the 8 children of a node next to each other
    void build(const K* leaves, int numLeaves)
    {
        nodes.clear();
        nodes.push_back({K(0), 0u, 0});
        for (size_t i = 0; i < nodes.size(); ++i)
        {
            const Node n = nodes[i];
            const K end  = n.start + nodeRange<K>(n.level);
            const int lo = findNodeAbove(leaves, numLeaves + 1, n.start);
            if (leaves[lo + 1] == end) continue; // a leaf of the cornerstone array
            nodes[i].child = int(nodes.size());
            for (int o = 0; o < 8; ++o)
                nodes.push_back({K(n.start + K(o) * nodeRange<K>(n.level + 1)), n.level + 1, 0});
        }
    }
    bool isLeaf(int i) const { return nodes[i].child == 0; }
    K codeStart(int i) const { return nodes[i].start; }
    K codeEnd(int i) const { return nodes[i].start + nodeRange<K>(nodes[i].level); }
    //! the node that covers exactly [start, start + nodeRange(level)), -1 if the tree has no such node
    int locate(K start, unsigned level) const
    {
        int i = 0;
        while (nodes[i].level < level)
        {
            if (isLeaf(i)) return -1;
            i = nodes[i].child + int(octalDigit<K>(start, nodes[i].level + 1));
        }
        return nodes[i].start == start ? i : -1;
    }
};

namespace let_detail
{

//! [ref: sfc/box.hpp:334-351 centerAndSize]
template<class K, class T>
inline void centerAndSize(const IBox& ib, const typename BoxT<T>::type& box, T* c, T* s)
{
    constexpr int maxCoord = 1 << KeyTraits<K>::maxLevel;
    constexpr T uL         = T(1.) / maxCoord;
    for (int d = 0; d < 3; ++d)
    {
        const T half = T(0.5) * uL * box.lengths[d];
        c[d]         = box.limits[2 * d] + (ib.hi[d] + ib.lo[d]) * half;
        s[d]         = (ib.hi[d] - ib.lo[d]) * half;
    }
}

//! squared distance of point X to the box (centre bc, half size bs), periodic where the box is
//! [ref: traversal/boxoverlap.hpp:207-217 minDistance, sfc/box.hpp:194-206 applyPbc]
template<class T>
inline T pointBoxDistSq(const T* X, const T* bc, const T* bs, const typename BoxT<T>::type& box)
{
    T d2[3];
    for (int d = 0; d < 3; ++d)
    {
        T dX = bc[d] - X[d];
        dX -= T(box.boundaries[d] == CSTONE_BOUNDARY_PERIODIC) * box.lengths[d] * std::rint(dX * box.inv_lengths[d]);
        dX = std::abs(dX);
        dX -= bs[d];
        dX += std::abs(dX);
        dX *= T(0.5);
        d2[d] = dX * dX;
    }
    return d2[0] + d2[1] + d2[2];
}

//! [ref: traversal/macs.hpp:170-193 minVecMacMutual]
template<class T>
inline bool minVecMacMutual(const T* cA, const T* sA, const T* cB, const T* sB, const typename BoxT<T>::type& box,
                            float invThetaEff)
{
    const T macB = std::max(sB[0], std::max(sB[1], sB[2])) * 2 * invThetaEff;
    const bool passA = pointBoxDistSq(cB, cA, sA, box) > macB * macB;
    const T macA = std::max(sA[0], std::max(sA[1], sA[2])) * 2 * invThetaEff;
    const bool passB = pointBoxDistSq(cA, cB, sB, box) > macA * macA;
    return passA && passB;
}

} // namespace let_detail

/*! ranks that own a global leaf which fails the mutual min-distance MAC against a global leaf of myRank: a dual
 *  traversal of the global tree, pairs (a overlapping the focus, b not inside it) are followed while they fail the MAC
 *  [ref: traversal/peers.hpp:62-117 findPeersMac, traversal/traversal.hpp:135-188 dualTraversal] */
template<class K, class T>
std::vector<int> findPeersMacHost(int kind, int myRank, const std::vector<K>& bnd, const HostOctree<K>& tree,
                                  const typename BoxT<T>::type& box, float invThetaEff)
{
    const int P         = int(bnd.size()) - 1;
    const K domainStart = bnd[myRank], domainEnd = bnd[myRank + 1];
    const uint16_t* dec = hostDecTable();
    const size_t M      = tree.nodes.size();
    std::vector<T> cen(3 * M), siz(3 * M);
    for (size_t i = 0; i < M; ++i)
    {
        IBox ib = sfcIBox<K>(kind, dec, tree.nodes[i].start, tree.nodes[i].level);
        let_detail::centerAndSize<K, T>(ib, box, &cen[3 * i], &siz[3 * i]);
    }
    auto follow = [&](int a, int b)
    {
        const bool aFocusOverlap = domainStart < tree.codeEnd(a) && tree.codeStart(a) < domainEnd; // overlapTwoRanges
        const bool bInFocus      = tree.codeStart(b) >= domainStart && tree.codeEnd(b) <= domainEnd; // containedIn
        if (!aFocusOverlap || bInFocus) return false;
        return !let_detail::minVecMacMutual(&cen[3 * a], &siz[3 * a], &cen[3 * b], &siz[3 * b], box, invThetaEff);
    };
    std::vector<char> isPeer(P, 0);
    auto leafPair = [&](int b)
    {
        // [ref: domain/domaindecomp.hpp SfcAssignment::findRank: upper_bound - 1]
        int r     = int(std::upper_bound(bnd.begin(), bnd.end(), tree.codeStart(b)) - bnd.begin()) - 1;
        isPeer[r] = 1;
    };

    std::vector<std::array<int, 2>> stack;
    // the maximal tree cells that tile [domainStart, domainEnd)  [ref: sfc/common.hpp:392-430 spanSfcRange]
    for (K a = domainStart; a < domainEnd;)
    {
        unsigned level = 0;
        while (level < KeyTraits<K>::maxLevel && ((a % nodeRange<K>(level)) != 0 || nodeRange<K>(level) > domainEnd - a))
            ++level;
        const int start = tree.locate(a, level);
        a += nodeRange<K>(level);
        if (start < 0) continue; // cannot happen: rank boundaries are global leaf boundaries
        if (tree.isLeaf(start) && tree.isLeaf(0))
        {
            if (follow(start, 0)) leafPair(0);
            continue;
        }
        stack.clear();
        stack.push_back({start, 0});
        while (!stack.empty())
        {
            const auto [target, source] = stack.back();
            stack.pop_back();
            auto interact = [&](int ta, int sb)
            {
                if (!follow(ta, sb)) return;
                if (tree.isLeaf(ta) && tree.isLeaf(sb)) leafPair(sb);
                else stack.push_back({ta, sb});
            };
            const bool splitTarget =
                (tree.nodes[target].level < tree.nodes[source].level && !tree.isLeaf(target)) || tree.isLeaf(source);
            if (splitTarget)
            {
                if (!tree.isLeaf(target))
                    for (int o = 0; o < 8; ++o)
                        interact(tree.nodes[target].child + o, source);
            }
            else
            {
                for (int o = 0; o < 8; ++o)
                    interact(target, tree.nodes[source].child + o);
            }
        }
    }
    std::vector<int> peers;
    for (int r = 0; r < P; ++r)
        if (isPeer[r]) peers.push_back(r);
    return peers;
}

// ------------------------------------------------------------------------------------------------
// host-side rebalance of a leaf range before it is handed to another rank
// ------------------------------------------------------------------------------------------------
/*! node op of leaf i of the cornerstone array `tree` (counts alongside)  [ref: tree/csarray.hpp:269-309
 *  siblingAndLevel + calculateNodeOp]; tree/counts point INTO the full arrays, sibling groups may reach outside the
 *  range that is rebalanced, exactly as in the reference's updateTreelet call [ref: focus/exchange_focus.hpp:380-405] */
template<class K>
inline int hostNodeOp(const K* tree, int i, const uint32_t* counts, unsigned bucket)
{
    const K thisNode     = tree[i];
    const unsigned level = treeLevel<K>(tree[i + 1] - thisNode);
    int siblingIdx       = -1;
    if (level > 0)
    {
        siblingIdx = int(octalDigit<K>(thisNode, level));
        if (!(tree[i - siblingIdx + 8] == tree[i - siblingIdx] + nodeRange<K>(level - 1))) siblingIdx = -1;
    }
    if (siblingIdx > 0)
    {
        const uint32_t* g = counts + i - siblingIdx;
        size_t parent     = 0;
        for (int k = 0; k < 8; ++k)
            parent += g[k];
        if (parent <= size_t(bucket)) return 0;
    }
    constexpr unsigned maxL = KeyTraits<K>::maxLevel;
    if (counts[i] > bucket * 512 && level + 3 < maxL) return 4096;
    if (counts[i] > bucket * 64 && level + 2 < maxL) return 512;
    if (counts[i] > bucket * 8 && level + 1 < maxL) return 64;
    if (counts[i] > bucket && level < maxL) return 8;
    return 1;
}

//! leaves [first, last] of `tree` with ops applied: 0 drops the key, 1 keeps it, 8^k splits k levels
//! [ref: tree/csarray.hpp:358-406 processNode + rebalanceTree]
template<class K>
inline std::vector<K> hostApplyOps(const K* tree, int first, int last, const std::vector<int>& ops)
{
    std::vector<K> out;
    for (int i = first; i < last; ++i)
    {
        const int op         = ops[i - first];
        const K thisNode     = tree[i];
        const unsigned level = treeLevel<K>(tree[i + 1] - thisNode);
        if (op == 1) out.push_back(thisNode);
        else if (op >= 8)
        {
            unsigned levelDiff = 0;
            for (int v = op; v > 1; v /= 8)
                ++levelDiff;
            for (int sib = 0; sib < op; ++sib)
                out.push_back(K(thisNode + K(sib) * nodeRange<K>(level + levelDiff)));
        }
    }
    out.push_back(tree[last]);
    return out;
}

/*! out[i] = index, in a linked tree given by its prefixes and level ranges, of the node that covers exactly the key
 *  range [pairs[2i], pairs[2i+1]); -1 if the tree has no such node  [ref: tree/octree.hpp locateNode] */
template<class K>
__global__ void locateNodesKernel(const K* __restrict__ prefixes, const int* __restrict__ levelRange,
                                  const K* __restrict__ pairs, int n, int32_t* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const K start        = pairs[2 * i];
    const unsigned level = treeLevel<K>(K(pairs[2 * i + 1] - start));
    const K prefix       = encodePlaceholderBit<K>(start, 3 * int(level));
    int lo = levelRange[level], hi = levelRange[level + 1];
    const int last = hi;
    while (lo < hi)
    {
        const int mid = (lo + hi) / 2;
        if (prefixes[mid] < prefix) lo = mid + 1;
        else hi = mid;
    }
    out[i] = (lo < last && prefixes[lo] == prefix) ? lo : -1;
}

/*! mass centre of the particles whose keys lie in [ranges[2i], ranges[2i+1]), summed in particle order as the
 *  reference's massCenter does [ref: focus/source_center.hpp:68-80]; one thread per range (a handful of ranges) */
template<class K, class T>
__global__ void rangeMassCenterKernel(const K* __restrict__ keys, uint32_t numKeys, const T* __restrict__ x,
                                      const T* __restrict__ y, const T* __restrict__ z, const T* __restrict__ m,
                                      const K* __restrict__ ranges, int numRanges, T* __restrict__ out4)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numRanges) return;
    auto lowerBound = [&](K v)
    {
        uint32_t lo = 0, hi = numKeys;
        while (lo < hi)
        {
            const uint32_t mid = (lo + hi) / 2;
            if (keys[mid] < v) lo = mid + 1;
            else hi = mid;
        }
        return lo;
    };
    const uint32_t a = lowerBound(ranges[2 * i]), b = lowerBound(ranges[2 * i + 1]);
    T c[4] = {0, 0, 0, 0};
    for (uint32_t j = a; j < b; ++j)
    {
        const T w = fabs(m[j]);
        c[0] += w * x[j], c[1] += w * y[j], c[2] += w * z[j], c[3] += w;
    }
    const T invM = (c[3] != T(0)) ? T(1) / c[3] : T(1);
    out4[4 * i] = c[0] * invM, out4[4 * i + 1] = c[1] * invM, out4[4 * i + 2] = c[2] * invM, out4[4 * i + 3] = c[3];
}

// ------------------------------------------------------------------------------------------------
// the tree
// ------------------------------------------------------------------------------------------------
template<class K, class T>
class LetTree
{
public:
    using CBox = typename BoxT<T>::type;
    static constexpr K kMask = K(1) << (8 * sizeof(K) - 1); // [ref: tree/definitions.h maskKey: top bit]

    LetTree(int rank, int numRanks, unsigned bucket, int kind)
        : t(kind, bucket, 1.0f)
        , rank_(rank)
        , P_(numRanks)
        , bucket_(bucket)
        , kind_(kind)
        , treelets_(numRanks)
        , tlOffset_(numRanks + 1, 0)
        , assign_(numRanks, {0, 0})
    {
        // [ref: octree_focus_mpi.hpp:692 Box<RealType> box_{0, 1}]
        for (int d = 0; d < 3; ++d)
        {
            box_.limits[2 * d] = 0, box_.limits[2 * d + 1] = 1;
            box_.lengths[d] = 1, box_.inv_lengths[d] = 1;
            box_.boundaries[d] = CSTONE_BOUNDARY_OPEN;
        }
        const K root[2]         = {K(0), nodeRange<K>(0)};
        const uint32_t rootCount = bucket + 1;
        leaves_.assign(root, root + 2);
        leafCounts_.assign(&rootCount, &rootCount + 1);
    }

    FocusTree<K> t; // leaves, linked tree, counts, macs: device

    const PinnedVec<K>& leaves() const { return leaves_; }
    const PinnedVec<uint32_t>& leafCounts() const { return leafCounts_; }
    const std::vector<std::array<int, 2>>& assignment() const { return assign_; }
    const std::vector<int>& peers() const { return peers_; }
    int numLeaves() const { return int(leaves_.size()) - 1; }
    /*! true if a step of this rank found the trees of the ranks inconsistent since the last call.  Such a step does not
     *  return early: its peers would wait forever in the next exchange.  The engine reduces the flags of all ranks at
     *  the end of the sync and fails on all of them together. */
    bool takeError()
    {
        const bool e = error_;
        error_       = false;
        return e;
    }

    //! [ref: octree_focus_mpi.hpp:457-525 updateMinMac + updateMacs]  (box: the one of the last updateTree)
    int updateMinMac(const std::vector<K>& bnd, float invThetaEff, cudaStream_t s)
    {
        if (!t.ready_) CSB_TRY(t.init(s));
        return t.template minMac<T>(&box_, invThetaEff, bnd[rank_], bnd[rank_ + 1], s);
    }

    //! [ref: octree_focus_mpi.hpp:108-187 updateTree]
    int updateTree(Comm* comm, const std::vector<int>& peers, const std::vector<K>& bnd, const CBox& box, int* converged,
                   cudaStream_t s)
    {
        if (!t.ready_) CSB_TRY(t.init(s));
        peers_             = peers;
        const K focusStart = bnd[rank_], focusEnd = bnd[rank_ + 1];
        if (prevStart_ == 0 && prevEnd_ == 0) prevStart_ = focusStart, prevEnd_ = focusEnd;

        std::vector<K> enforced;
        CSB_TRY(focusTransfer(comm, focusStart, focusEnd, enforced, s));
        for (int peer : peers_)
        {
            enforced.push_back(bnd[peer]);
            enforced.push_back(bnd[peer + 1]);
        }
        enforced.erase(std::unique(enforced.begin(), enforced.end()), enforced.end());

        static const bool lapOn = std::getenv("CSB_HALO_TIMING") != nullptr;
        auto lapT = std::chrono::steady_clock::now();
        std::string lapLog;
        auto lap = [&](const char* name)
        {
            if (!lapOn) return;
            cudaStreamSynchronize(s);
            auto now = std::chrono::steady_clock::now();
            char buf[64];
            std::snprintf(buf, sizeof(buf), " %s %.2f", name, std::chrono::duration<double, std::milli>(now - lapT).count());
            lapLog += buf;
            lapT = now;
        };
        lap("focusTransfer");
        bool conv = false;
        CSB_TRY(t.rebalanceStep(focusStart, focusEnd, enforced.data(), int(enforced.size()), &conv, s));
        lap("rebalance");
        const float invThetaRefine = float(std::sqrt(3) / 2 + 1e-6);
        for (int c = 0; !c;)
            CSB_TRY(t.template macRefine<T>(&box, prevStart_, prevEnd_, focusStart, focusEnd, invThetaRefine, &c, s));

        lap("macRefine");
        CSB_TRY(downloadLeaves(s));
        translateAssignment(bnd);
        lap("leaves");
        CSB_TRY(syncTreelets(comm, s));
        lap("treelets");
        CSB_TRY(indexTreelets(s));
        translateAssignment(bnd);
        lap("index");
        if (lapOn && rank_ == 0) std::fprintf(stderr, "[csb rank 0]   updateTree ms:%s\n", lapLog.c_str());

        box_       = box;
        globBnd_   = bnd;
        prevStart_ = focusStart, prevEnd_ = focusEnd;
        if (converged) *converged = conv ? 1 : 0;
        return 0;
    }

    //! [ref: octree_focus_mpi.hpp:205-273 updateCounts]
    int updateCounts(Comm* comm, const K* keysDev, size_t numKeys, const std::vector<K>& gLeaves,
                     const std::vector<uint32_t>& gCounts, cudaStream_t s)
    {
        const int L = t.numLeaves_;
        CSB_TRY(t.localCounts(keysDev, numKeys, s));

        // leaves outside of my range and of every peer's range: sums over the global leaves they span
        // [ref: domain/layout.hpp:57-90 invertRanges + enumerateRanges, focus/rebalance.hpp:278-299 rangeCount]
        std::vector<uint32_t> idx = nonPeerLeaves(), val(0);
        val.reserve(idx.size());
        for (uint32_t i : idx)
        {
            const int gs = findNodeBelow(gLeaves.data(), int(gLeaves.size()), leaves_[i]);
            const int ge = findNodeAbove(gLeaves.data(), int(gLeaves.size()), leaves_[i + 1]);
            uint64_t sum = 0;
            for (int g = gs; g < ge; ++g)
                sum += gCounts[g];
            val.push_back(uint32_t(std::min<uint64_t>(0xFFFFFFFFull, sum)));
        }
        if (!idx.empty())
        {
            const size_t n = idx.size();
            CSB_TRY(stage_.reserve(8 * n + 512));
            uint32_t* dIdx = stage_.as<uint32_t>();
            uint32_t* dVal = dIdx + (n + 63) / 64 * 64;
            CSB_CUDA(cudaMemcpyAsync(dIdx, idx.data(), 4 * n, cudaMemcpyHostToDevice, s));
            CSB_CUDA(cudaMemcpyAsync(dVal, val.data(), 4 * n, cudaMemcpyHostToDevice, s));
            CSB_TRY(cstone_b200_scatter(dIdx, n, dVal, t.leafCounts_.p, 4, s));
            CSB_CUDA(cudaStreamSynchronize(s)); // idx / val are host temporaries
        }
        // 1st upsweep with local and global data, counts of the peers' ranges from their owners, 2nd upsweep
        CSB_TRY(t.scatterCounts(s));
        CSB_TRY(t.upsweepCounts(s));
        CSB_TRY(peerExchange(comm, t.counts_.p, 4, s));
        CSB_TRY(t.upsweepCounts(s));
        CSB_TRY(cstone_b200_gather(t.l2i_.template as<uint32_t>() + t.numInternal_, size_t(L), t.counts_.p, t.leafCounts_.p, 4, s));
        leafCounts_.resize(L);
        CSB_CUDA(cudaMemcpyAsync(leafCounts_.data(), t.leafCounts_.p, 4 * size_t(L), cudaMemcpyDeviceToHost, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        return 0;
    }

    //! leaves that lie neither in my range nor in a peer's: their data comes from the replicated global tree
    //! [ref: domain/layout.hpp:57-90 invertRanges + enumerateRanges]
    std::vector<uint32_t> nonPeerLeaves() const
    {
        std::vector<uint32_t> idx;
        const int L = int(leaves_.size()) - 1;
        int current = 0;
        for (const auto& r : assign_)
        {
            if (r[0] == r[1]) continue;
            for (int i = current; i < r[0]; ++i)
                idx.push_back(uint32_t(i));
            current = r[1];
        }
        for (int i = current; i < L; ++i)
            idx.push_back(uint32_t(i));
        return idx;
    }

    //! index of the node [start, end) in a linked tree given by its prefixes and level ranges, -1 if absent
    //! [ref: tree/octree.hpp locateNode]
    static int locateNode(K start, K end, const K* prefixes, const int* levelRange)
    {
        const unsigned level = treeLevel<K>(end - start);
        const K prefix       = encodePlaceholderBit<K>(start, 3 * int(level));
        const K* first       = prefixes + levelRange[level];
        const K* last        = prefixes + levelRange[level + 1];
        const K* it          = std::lower_bound(first, last, prefix);
        return (it == last || *it != prefix) ? -1 : int(it - prefixes);
    }

    /*! expansion (mass) centres of every node [ref: octree_focus_mpi.hpp:368-449 updateCenters]: leaves of my range from
     *  the particles (x, y, z, m: the assigned particles in SFC order), upsweep; cells above the local domains through
     *  the replicated global tree (every rank contributes the centres of its global leaves, upsweep there, the leaves
     *  of non-peer ranks read theirs back: populateGlobal / gatherGlobalLeaves / extractGlobal :288-366); leaves in a
     *  peer's range from the owner; upsweep.  gLeaves: host copy of the global leaf keys. */
    int updateCenters(Comm* comm, const K* keysDev, size_t numKeys, const T* x, const T* y, const T* z, const T* m,
                      const std::vector<K>& gLeaves, cudaStream_t s)
    {
        constexpr int vb   = 4 * int(sizeof(T)); // bytes of one centre
        const int L        = t.numLeaves_, M = t.numNodes_;
        const int firstIdx = assign_[rank_][0], lastIdx = assign_[rank_][1];
        const int levels   = int(KeyTraits<K>::maxLevel) + 1;
        /* particle ranges of my leaves, every other leaf empty.  (The reference leaves layout[i] = 0 behind my range, i.e.
         * start above end there; the kernel wants a monotone layout, so those entries hold the total: the same empty
         * ranges.) */
        CSB_TRY(layoutOwn_.reserve(4 * (size_t(L) + 2)));
        CSB_TRY(t.temp_.reserve(size_t(cstone_b200_scan_temp_bytes(uint64_t(L) + 2)) + 4096));
        CSB_CUDA(cudaMemsetAsync(layoutOwn_.p, 0, 4 * (size_t(L) + 1), s));
        if (lastIdx > firstIdx)
        {
            CSB_TRY(cstone_b200_exclusive_scan_u32(t.leafCounts_.template as<uint32_t>() + firstIdx,
                                                   layoutOwn_.as<uint32_t>() + firstIdx, size_t(lastIdx - firstIdx) + 1, 0u,
                                                   t.temp_.p, s));
            uint32_t total = 0;
            for (int i = firstIdx; i < lastIdx; ++i)
                total += leafCounts_[i];
            if (L > lastIdx) CSB_TRY(cstone_b200_fill(layoutOwn_.as<uint32_t>() + lastIdx + 1, size_t(L - lastIdx), &total, 4, s));
        }
        CSB_TRY(t.centers_.reserve(size_t(M) * vb));
        CSB_TRY(leafCenters(x, y, z, m, t.l2i_.template as<int32_t>() + t.numInternal_, L, layoutOwn_.as<uint32_t>(),
                            t.centers_.template as<T>(), s));
        CSB_TRY(upsweepCenters(levels, t.levelRange_, t.childOffsets_.template as<int32_t>(), t.centers_.template as<T>(), s));

        // ---- cells above the local domains: through the replicated global tree
        CSB_TRY(linkGlobal(gLeaves, s));
        const int Lg = g_.numLeaves_, Mg = g_.numNodes_;
        CSB_TRY(gLeafCenters_.reserve(size_t(Lg) * vb + 256));
        CSB_TRY(g_.centers_.reserve(size_t(Mg) * vb + 256));
        /* A global leaf of my range that is no node of my tree (the focus tree follows the particles one rebalance step
         * per sync and can be coarser than the global tree for a step; the reference then reads an unrelated node) takes
         * its centre from the particles directly. */
        std::vector<uint32_t> missing;
        CSB_TRY(populateGlobal(t.centers_.p, vb, gLeaves, gLeafCenters_.p, &missing, s));
        if (!missing.empty())
        {
            const size_t nm = missing.size();
            std::vector<K> missingRanges;
            for (uint32_t i : missing)
                missingRanges.push_back(gLeaves[i]), missingRanges.push_back(gLeaves[i + 1]);
            const size_t oOut = (nm * 2 * sizeof(K) + 255) / 256 * 256, oIdx = oOut + (nm * vb + 255) / 256 * 256;
            CSB_TRY(xchg_.reserve(oIdx + 4 * nm + 256));
            K* dRanges     = xchg_.as<K>();
            T* dOut        = reinterpret_cast<T*>(xchg_.as<char>() + oOut);
            uint32_t* dIdx = reinterpret_cast<uint32_t*>(xchg_.as<char>() + oIdx);
            CSB_CUDA(cudaMemcpyAsync(dRanges, missingRanges.data(), nm * 2 * sizeof(K), cudaMemcpyHostToDevice, s));
            CSB_CUDA(cudaMemcpyAsync(dIdx, missing.data(), 4 * nm, cudaMemcpyHostToDevice, s));
            rangeMassCenterKernel<K, T><<<unsigned((nm + 63) / 64), 64, 0, s>>>(keysDev, uint32_t(numKeys), x, y, z, m, dRanges,
                                                                               int(nm), dOut);
            CSB_LAUNCHED();
            CSB_TRY(cstone_b200_scatter(dIdx, nm, dOut, gLeafCenters_.p, vb, s));
            CSB_CUDA(cudaStreamSynchronize(s));
        }
        CSB_TRY(gatherGlobalLeaves(comm, gLeafCenters_.p, vb, gLeaves, s));
        CSB_TRY(cstone_b200_scatter(g_.l2i_.template as<uint32_t>() + g_.numInternal_, size_t(Lg), gLeafCenters_.p, g_.centers_.p, vb, s));
        CSB_TRY(upsweepCenters(levels, g_.levelRange_, g_.childOffsets_.template as<int32_t>(), g_.centers_.template as<T>(), s));
        CSB_TRY(extractGlobal(g_.centers_.p, vb, t.centers_.p, s));
        CSB_TRY(peerExchange(comm, t.centers_.p, vb, s));
        CSB_TRY(upsweepCenters(levels, t.levelRange_, t.childOffsets_.template as<int32_t>(), t.centers_.template as<T>(), s));
        return 0;
    }

    // ---- the three steps of globalFocusExchange [ref: octree_focus_mpi.hpp:762-784] for node quantities of any element
    //      size (4 - 32 bytes); between gatherGlobalLeaves and extractGlobal the caller scatters the global leaf
    //      quantities to the nodes of the global tree (globalTree()) and runs ITS upsweep there

    //! the replicated global tree, linked on the device
    int linkGlobal(const std::vector<K>& gLeaves, cudaStream_t s)
    {
        CSB_TRY(g_.setLeavesHost(gLeaves.data(), int(gLeaves.size()) - 1, s));
        gPrefixes_.resize(size_t(g_.numNodes_));
        CSB_CUDA(cudaMemcpyAsync(gPrefixes_.data(), g_.prefixes_.p, size_t(g_.numNodes_) * sizeof(K), cudaMemcpyDeviceToHost, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        return 0;
    }
    const FocusTree<K>& globalTree() const { return g_; }

    /*! [ref: octree_focus_mpi.hpp:288-316 populateGlobal] gLeafQ (device, one element per global leaf) is zeroed, then
     *  the global leaves of my range take the quantities of the nodes of my tree that cover them.  Global leaves without
     *  such a node are listed in *missing if given, otherwise they count as an inconsistency. */
    int populateGlobal(const void* nodeQ, int elemBytes, const std::vector<K>& gLeaves, void* gLeafQ,
                       std::vector<uint32_t>* missing, cudaStream_t s)
    {
        const int nk = int(gLeaves.size()), Lg = nk - 1;
        CSB_CUDA(cudaMemsetAsync(gLeafQ, 0, size_t(Lg) * elemBytes, s));
        const int firstGlb = findNodeAbove(gLeaves.data(), nk, prevStart_), lastGlb = findNodeAbove(gLeaves.data(), nk, prevEnd_);
        std::vector<K> myGlobal;
        for (int i = firstGlb; i < lastGlb; ++i)
            myGlobal.push_back(gLeaves[i]), myGlobal.push_back(gLeaves[i + 1]);
        std::vector<int32_t> located;
        CSB_TRY(locateNodes(t, myGlobal, located, s));
        std::vector<uint32_t> gmap;
        for (int i = firstGlb; i < lastGlb; ++i)
        {
            const int node = located[size_t(i - firstGlb)];
            if (node < 0)
            {
                if (missing) missing->push_back(uint32_t(i));
                else fail(__LINE__);
            }
            gmap.push_back(uint32_t(std::max(node, 0)));
        }
        if (!gmap.empty())
        {
            CSB_TRY(stage_.reserve(4 * gmap.size() + 1024));
            CSB_CUDA(cudaMemcpyAsync(stage_.p, gmap.data(), 4 * gmap.size(), cudaMemcpyHostToDevice, s));
            CSB_TRY(cstone_b200_gather(stage_.as<uint32_t>(), gmap.size(), nodeQ, static_cast<char*>(gLeafQ) + size_t(firstGlb) * elemBytes,
                                       elemBytes, s));
            CSB_CUDA(cudaStreamSynchronize(s)); // gmap is a host temporary
        }
        return 0;
    }

    //! [ref: octree_focus_mpi.hpp:350-366 gatherGlobalLeaves] every rank's part of gLeafQ goes to every rank (the
    //! in-place all-gather with ragged counts of the reference, as one exchange step)
    int gatherGlobalLeaves(Comm* comm, void* gLeafQ, int elemBytes, const std::vector<K>& gLeaves, cudaStream_t s)
    {
        const int Lg = int(gLeaves.size()) - 1;
        std::vector<int> displ(size_t(P_) + 1);
        for (int r = 0; r < P_; ++r)
            displ[r] = findNodeAbove(gLeaves.data(), Lg, globBnd_[r]);
        displ[P_] = Lg;
        std::vector<CommMsg> sends, recvs;
        char* q = static_cast<char*>(gLeafQ);
        for (int r = 0; r < P_; ++r)
        {
            if (r == rank_) continue;
            const size_t mineBytes = size_t(displ[rank_ + 1] - displ[rank_]) * elemBytes;
            const size_t hisBytes  = size_t(displ[r + 1] - displ[r]) * elemBytes;
            if (mineBytes) sends.push_back({r, q + size_t(displ[rank_]) * elemBytes, mineBytes});
            if (hisBytes) recvs.push_back({r, q + size_t(displ[r]) * elemBytes, hisBytes});
        }
        return comm->exchange(sends, recvs, s);
    }

    //! [ref: octree_focus_mpi.hpp:325-347 extractGlobal] the leaves of my tree outside of my range and of every peer's
    //! range take the quantity of the global node with their key range; gNodeQ: one element per node of globalTree()
    int extractGlobal(const void* gNodeQ, int elemBytes, void* nodeQ, cudaStream_t s)
    {
        std::vector<uint32_t> leafIdx, gIdx;
        for (uint32_t i : nonPeerLeaves())
        {
            const int gi = locateNode(leaves_[i], leaves_[i + 1], gPrefixes_.data(), g_.levelRange_);
            if (gi < 0)
            {
                // finer than the global tree outside of every peer: no source for this cell.  The other ranks are not
                // left alone in the exchanges that follow; the error surfaces on all ranks at the end of the sync
                fail(__LINE__);
                continue;
            }
            leafIdx.push_back(i);
            gIdx.push_back(uint32_t(gi));
        }
        const size_t n = leafIdx.size();
        if (n == 0) return 0;
        const size_t stride = (n + 63) / 64 * 64;
        CSB_TRY(stage_.reserve(12 * stride + 1024));
        CSB_TRY(xchg_.reserve(n * elemBytes + 512));
        uint32_t* dG = stage_.as<uint32_t>();
        uint32_t* dL = dG + stride;
        uint32_t* dN = dL + stride;
        CSB_CUDA(cudaMemcpyAsync(dG, gIdx.data(), 4 * n, cudaMemcpyHostToDevice, s));
        CSB_CUDA(cudaMemcpyAsync(dL, leafIdx.data(), 4 * n, cudaMemcpyHostToDevice, s));
        // node index of those leaves, then quantity: global node -> buffer -> node of my tree
        CSB_TRY(cstone_b200_gather(dL, n, t.l2i_.template as<uint32_t>() + t.numInternal_, dN, 4, s));
        CSB_TRY(cstone_b200_gather(dG, n, gNodeQ, xchg_.p, elemBytes, s));
        CSB_TRY(cstone_b200_scatter(dN, n, xchg_.p, nodeQ, elemBytes, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        return 0;
    }

    //! MAC radius around the expansion centres, marks w.r.t. my leaves [ref: octree_focus_mpi.hpp:481-525
    //! setMacRadius + updateMacs]
    int updateMacs(const std::vector<K>& bnd, float invTheta, cudaStream_t s)
    {
        using A = FocusApi<K>;
        CSB_TRY(t.temp_.reserve(size_t(cstone_b200_reduce_temp_bytes()) + 4096));
        CSB_TRY(setMac(kind_, t.prefixes_.template as<K>(), t.numNodes_, t.centers_.template as<T>(), invTheta, &box_, s));
        CSB_TRY(t.macs_.reserve(size_t(t.numNodes_)));
        CSB_CUDA(cudaMemsetAsync(t.macs_.p, 0, size_t(t.numNodes_), s));
        const int nk     = int(leaves_.size()) - 1;
        const int fStart = findNodeAbove(leaves_.data(), nk, bnd[rank_]), fEnd = findNodeAbove(leaves_.data(), nk, bnd[rank_ + 1]);
        if (fEnd > fStart)
            CSB_TRY(A::mark(kind_, t.prefixes_.template as<K>(), t.childOffsets_.template as<int>(), t.centers_.template as<T>(),
                            &box_, t.leaves_.template as<K>() + fStart, fEnd - fStart, 0, t.macs_.template as<char>(), s));
        return 0;
    }

    /*! node quantities (elemBytes each, indexed like the linked tree) of the peers' ranges are replaced by the owners'
     *  values  [ref: focus/exchange_focus.hpp:265-330 exchangeTreeletGeneral] */
    int peerExchange(Comm* comm, void* quantities, int elemBytes, cudaStream_t s)
    {
        const size_t nSend = size_t(tlOffset_[P_]);
        size_t nRecv       = 0;
        for (int peer : peers_)
            nRecv += size_t(assign_[peer][1] - assign_[peer][0]);
        const size_t sendBytes = (nSend * elemBytes + 255) / 256 * 256;
        CSB_TRY(xchg_.reserve(sendBytes + nRecv * elemBytes + 512));
        char* sendBuf = xchg_.as<char>();
        char* recvBuf = sendBuf + sendBytes;
        if (nSend) CSB_TRY(cstone_b200_gather(tlIdxDev_.as<uint32_t>(), nSend, quantities, sendBuf, elemBytes, s));
        std::vector<CommMsg> sends, recvs;
        size_t ro = 0;
        for (int peer : peers_)
        {
            const size_t ns = size_t(tlOffset_[peer + 1] - tlOffset_[peer]);
            const size_t nr = size_t(assign_[peer][1] - assign_[peer][0]);
            if (ns) sends.push_back({peer, sendBuf + size_t(tlOffset_[peer]) * elemBytes, ns * elemBytes});
            if (nr) recvs.push_back({peer, recvBuf + ro * elemBytes, nr * elemBytes});
            ro += nr;
        }
        CSB_TRY(comm->exchange(sends, recvs, s));
        ro = 0;
        for (int peer : peers_)
        {
            const size_t nr = size_t(assign_[peer][1] - assign_[peer][0]);
            if (nr)
                CSB_TRY(cstone_b200_scatter(t.l2i_.template as<uint32_t>() + t.numInternal_ + assign_[peer][0], nr,
                                            recvBuf + ro * elemBytes, quantities, elemBytes, s));
            ro += nr;
        }
        return 0;
    }

    /*! variable-size messages between ranks: to[q] goes to rank q, from[q] is what rank q sent here.  Every rank calls
     *  it (one all-gather of the sizes, one exchange step).  Stands in for the reference's MPI_Probe'd messages. */
    int exchangeBytes(Comm* comm, const std::vector<std::vector<char>>& to, std::vector<std::vector<char>>& from,
                      cudaStream_t s)
    {
        const int P = P_;
        CSB_TRY(sizes_.reserve(size_t(P) * 4 + size_t(P) * P * 4 + 512));
        uint32_t* dMine = sizes_.as<uint32_t>();
        uint32_t* dAll  = dMine + (P + 63) / 64 * 64;
        std::vector<uint32_t> mine(P, 0), all(size_t(P) * P);
        size_t total = 0;
        for (int q = 0; q < P; ++q)
            mine[q] = uint32_t(to[q].size()), total += (to[q].size() + 15) / 16 * 16;
        CSB_CUDA(cudaMemcpyAsync(dMine, mine.data(), 4 * size_t(P), cudaMemcpyHostToDevice, s));
        CSB_TRY(comm->allGather(dMine, dAll, 4 * size_t(P), s));
        CSB_CUDA(cudaMemcpyAsync(all.data(), dAll, 4 * size_t(P) * P, cudaMemcpyDeviceToHost, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        size_t totalRecv = 0;
        for (int q = 0; q < P; ++q)
            totalRecv += (size_t(all[size_t(q) * P + rank_]) + 15) / 16 * 16;
        const size_t sendBytes = (total + 255) / 256 * 256;
        CSB_TRY(xchg_.reserve(sendBytes + totalRecv + 512));
        char* sendBuf = xchg_.as<char>();
        char* recvBuf = sendBuf + sendBytes;
        std::vector<CommMsg> sends, recvs;
        size_t so = 0, ro = 0;
        for (int q = 0; q < P; ++q)
        {
            if (!to[q].empty())
            {
                CSB_CUDA(cudaMemcpyAsync(sendBuf + so, to[q].data(), to[q].size(), cudaMemcpyHostToDevice, s));
                sends.push_back({q, sendBuf + so, to[q].size()});
                so += (to[q].size() + 15) / 16 * 16;
            }
            const size_t nr = all[size_t(q) * P + rank_];
            if (nr)
            {
                recvs.push_back({q, recvBuf + ro, nr});
                ro += (nr + 15) / 16 * 16;
            }
        }
        CSB_TRY(comm->exchange(sends, recvs, s));
        from.assign(P, {});
        ro = 0;
        for (int q = 0; q < P; ++q)
        {
            const size_t nr = all[size_t(q) * P + rank_];
            if (!nr) continue;
            from[q].resize(nr);
            CSB_CUDA(cudaMemcpyAsync(from[q].data(), recvBuf + ro, nr, cudaMemcpyDeviceToHost, s));
            ro += (nr + 15) / 16 * 16;
        }
        CSB_CUDA(cudaStreamSynchronize(s));
        return 0;
    }

    //! a replaced leaf array (host) goes to the device, the tree is relinked and the host copies follow
    int uploadLeaves(cudaStream_t s)
    {
        CSB_TRY(t.setLeavesHost(leaves_.data(), int(leaves_.size()) - 1, s));
        return 0;
    }

private:
    void fail(int line)
    {
        error_ = true;
        traceStatus(CSTONE_E_LET, __FILE__, line);
    }
    bool error_ = false;
    static int leafCenters(const float* x, const float* y, const float* z, const float* m, const int32_t* l2i, int L,
                           const uint32_t* layout, float* c, void* s)
    {
        return cstone_b200_compute_leaf_source_center_f32_f32_f32(x, y, z, m, l2i, L, layout, c, s);
    }
    static int leafCenters(const double* x, const double* y, const double* z, const double* m, const int32_t* l2i, int L,
                           const uint32_t* layout, double* c, void* s)
    {
        return cstone_b200_compute_leaf_source_center_f64_f64_f64(x, y, z, m, l2i, L, layout, c, s);
    }
    static int upsweepCenters(int levels, const int* lr, const int32_t* co, float* c, void* s)
    {
        return cstone_b200_upsweep_centers_f32(levels, lr, co, c, s);
    }
    static int upsweepCenters(int levels, const int* lr, const int32_t* co, double* c, void* s)
    {
        return cstone_b200_upsweep_centers_f64(levels, lr, co, c, s);
    }
    static int setMac(int kind, const uint32_t* p, int n, float* c, float it, const cstone_box_f32* b, void* s)
    {
        return cstone_b200_set_mac_k32_f32(kind, p, n, c, it, b, s);
    }
    static int setMac(int kind, const uint32_t* p, int n, double* c, float it, const cstone_box_f64* b, void* s)
    {
        return cstone_b200_set_mac_k32_f64(kind, p, n, c, it, b, s);
    }
    static int setMac(int kind, const uint64_t* p, int n, float* c, float it, const cstone_box_f32* b, void* s)
    {
        return cstone_b200_set_mac_k64_f32(kind, p, n, c, it, b, s);
    }
    static int setMac(int kind, const uint64_t* p, int n, double* c, float it, const cstone_box_f64* b, void* s)
    {
        return cstone_b200_set_mac_k64_f64(kind, p, n, c, it, b, s);
    }
    static std::vector<char> asBytes(const std::vector<K>& v)
    {
        std::vector<char> b(v.size() * sizeof(K));
        if (!v.empty()) std::memcpy(b.data(), v.data(), b.size());
        return b;
    }
    static std::vector<K> asKeys(const std::vector<char>& b)
    {
        std::vector<K> v(b.size() / sizeof(K));
        if (!v.empty()) std::memcpy(v.data(), b.data(), v.size() * sizeof(K));
        return v;
    }

    int downloadLeaves(cudaStream_t s)
    {
        leaves_.resize(size_t(t.numLeaves_) + 1);
        CSB_CUDA(cudaMemcpyAsync(leaves_.data(), t.leaves_.p, leaves_.size() * sizeof(K), cudaMemcpyDeviceToHost, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        return 0;
    }

    /*! parts of the tree whose SFC range moved to the neighbour rank are handed over, rebalanced once with their counts,
     *  and become mandatory keys at the receiver  [ref: focus/exchange_focus.hpp:364-433 focusTransfer] */
    int focusTransfer(Comm* comm, K newStart, K newEnd, std::vector<K>& buffer, cudaStream_t s)
    {
        const int nk = int(leaves_.size());
        std::vector<std::vector<char>> to(P_), from;
        auto treeletOf = [&](K a, K b)
        {
            const int first = findNodeAbove(leaves_.data(), nk, a);
            const int last  = findNodeAbove(leaves_.data(), nk, b);
            std::vector<int> ops(size_t(std::max(last - first, 0)));
            for (int i = first; i < last; ++i)
                ops[i - first] = hostNodeOp<K>(leaves_.data(), i, leafCounts_.data(), bucket_);
            std::vector<K> tl = hostApplyOps<K>(leaves_.data(), first, last, ops);
            tl.pop_back(); // the upper boundary does not travel
            return tl;
        };
        if (prevStart_ < newStart && rank_ > 0) to[rank_ - 1] = asBytes(treeletOf(prevStart_, newStart));
        if (newEnd < prevEnd_ && rank_ + 1 < P_) to[rank_ + 1] = asBytes(treeletOf(newEnd, prevEnd_));
        CSB_TRY(exchangeBytes(comm, to, from, s));
        if (newStart < prevStart_ && rank_ > 0)
        {
            std::vector<K> k = asKeys(from[rank_ - 1]);
            buffer.insert(buffer.end(), k.begin(), k.end());
        }
        if (prevEnd_ < newEnd && rank_ + 1 < P_)
        {
            std::vector<K> k = asKeys(from[rank_ + 1]);
            buffer.insert(buffer.end(), k.begin(), k.end());
        }
        return 0;
    }

    //! leaf index ranges of the peers' (and my) SFC ranges in my tree; a range without exact boundaries is narrowed
    //! [ref: domain/domaindecomp.hpp:183-207 translateAssignment]
    void translateAssignment(const std::vector<K>& bnd)
    {
        const int nk = int(leaves_.size());
        std::fill(assign_.begin(), assign_.end(), std::array<int, 2>{0, 0});
        for (int peer : peers_)
        {
            int a = findNodeAbove(leaves_.data(), nk, bnd[peer]);
            int b = findNodeBelow(leaves_.data(), nk, bnd[peer + 1]);
            if (b < a) b = a;
            assign_[peer] = {a, b};
        }
        assign_[rank_] = {findNodeAbove(leaves_.data(), nk, bnd[rank_]), findNodeBelow(leaves_.data(), nk, bnd[rank_ + 1])};
    }

    /*! every peer learns which nodes I hold of ITS range; keys it does not have are rejected and leave my tree
     *  [ref: focus/exchange_focus.hpp:61-217 exchangeTreelets, checkTreelets, exchangeRejectedKeys, pruneTreelets,
     *  syncTreeletsGpu] */
    int syncTreelets(Comm* comm, cudaStream_t s)
    {
        const int nk = int(leaves_.size());
        std::vector<std::vector<char>> to(P_), from;
        for (int peer : peers_)
        {
            const K* first = leaves_.data() + assign_[peer][0];
            to[peer].resize(size_t(assign_[peer][1] - assign_[peer][0] + 1) * sizeof(K));
            std::memcpy(to[peer].data(), first, to[peer].size());
        }
        CSB_TRY(exchangeBytes(comm, to, from, s));
        for (auto& v : treelets_)
            v.clear();
        std::vector<std::vector<char>> rejected(P_), rejectedHere;
        for (int peer : peers_)
        {
            std::vector<K> tl = asKeys(from[peer]);
            std::vector<K> rej, kept;
            const int numNodes = int(tl.size()) - 1;
            kept.reserve(tl.size());
            int pos = 0;
            for (int i = 0; i < numNodes; ++i)
            {
                const K k = tl[i];
                pos       = findNodeAboveFrom(leaves_.data(), nk - 1, pos, k); // (treelet keys ascend)
                if (k != leaves_[pos]) rej.push_back(k);
                else kept.push_back(k);
            }
            if (!tl.empty()) kept.push_back(tl.back());
            treelets_[peer] = std::move(kept);
            rejected[peer]  = asBytes(rej);
        }
        CSB_TRY(exchangeBytes(comm, rejected, rejectedHere, s));
        // the rejected keys leave my tree: on the device, then the host copy of the leaves is refreshed
        std::vector<uint32_t> drop;
        for (int peer : peers_)
            for (K k : asKeys(rejectedHere[peer]))
                drop.push_back(uint32_t(findNodeAbove(leaves_.data(), nk, k)));
        if (!drop.empty())
        {
            std::sort(drop.begin(), drop.end());
            drop.erase(std::unique(drop.begin(), drop.end()), drop.end());
            CSB_TRY(stage_.reserve(4 * drop.size() + 256));
            CSB_CUDA(cudaMemcpyAsync(stage_.p, drop.data(), 4 * drop.size(), cudaMemcpyHostToDevice, s));
            CSB_TRY(t.dropLeaves(stage_.as<uint32_t>(), int(drop.size()), s));
            CSB_TRY(downloadLeaves(s));
        }
        return 0;
    }

    //! node index in MY linked tree of every node the peers hold of my range
    //! [ref: focus/exchange_focus.hpp:237-262 indexTreelets, tree/octree.hpp locateNode]
    int indexTreelets(cudaStream_t s)
    {
        std::fill(tlOffset_.begin(), tlOffset_.end(), 0);
        for (int r = 0; r < P_; ++r)
            tlOffset_[r + 1] = tlOffset_[r] + (treelets_[r].empty() ? 0 : int(treelets_[r].size()) - 1);
        std::vector<K> pairs(2 * size_t(tlOffset_[P_]));
        for (int peer : peers_)
        {
            const std::vector<K>& tl = treelets_[peer];
            for (int i = 0; i + 1 < int(tl.size()); ++i)
            {
                pairs[2 * (size_t(tlOffset_[peer]) + i)]     = tl[i];
                pairs[2 * (size_t(tlOffset_[peer]) + i) + 1] = tl[i + 1];
            }
        }
        // (ranks that are no peers any more keep a count of zero nodes: treelets_ was cleared in syncTreelets)
        CSB_TRY(locateNodes(t, pairs, tlIdx_, s));
        bool missing = false;
        for (int32_t& v : tlIdx_)
            if (v < 0)
            {
                fail(__LINE__);
                v = 0, missing = true;
            }
        CSB_TRY(tlIdxDev_.reserve(4 * tlIdx_.size() + 256));
        if (!tlIdx_.empty())
        {
            if (missing) CSB_CUDA(cudaMemcpyAsync(tlIdxDev_.p, tlIdx_.data(), 4 * tlIdx_.size(), cudaMemcpyHostToDevice, s));
            else CSB_CUDA(cudaMemcpyAsync(tlIdxDev_.p, locOut_.p, 4 * tlIdx_.size(), cudaMemcpyDeviceToDevice, s));
        }
        CSB_CUDA(cudaStreamSynchronize(s));
        return 0;
    }

    //! out[i] = node of `tree` covering [pairs[2i], pairs[2i+1]) or -1, located on the device (the prefixes stay in HBM);
    //! the result is left in locOut_ as well
    int locateNodes(const FocusTree<K>& tree, const std::vector<K>& pairs, std::vector<int32_t>& out, cudaStream_t s)
    {
        const size_t n = pairs.size() / 2;
        out.assign(n, 0);
        if (n == 0) return 0;
        CSB_TRY(locIn_.reserve(2 * n * sizeof(K) + 256));
        CSB_TRY(locOut_.reserve(4 * n + 256));
        CSB_CUDA(cudaMemcpyAsync(locIn_.p, pairs.data(), 2 * n * sizeof(K), cudaMemcpyHostToDevice, s));
        locateNodesKernel<K><<<unsigned((n + 255) / 256), 256, 0, s>>>(tree.prefixes_.template as<K>(), tree.levelRangeDev_.template as<int>(),
                                                                      locIn_.template as<K>(), int(n), locOut_.template as<int32_t>());
        CSB_LAUNCHED();
        CSB_CUDA(cudaMemcpyAsync(out.data(), locOut_.p, 4 * n, cudaMemcpyDeviceToHost, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        return 0;
    }

    int rank_, P_;
    unsigned bucket_;
    int kind_;
    CBox box_;
    K prevStart_ = 0, prevEnd_ = 0;
    PinnedVec<K> leaves_;               // host copy of the leaf keys
    PinnedVec<uint32_t> leafCounts_;    // host copy of the leaf counts
    FBuf locIn_, locOut_;               // key ranges to locate in a linked tree / their node indices
    std::vector<int> peers_;
    std::vector<std::vector<K>> treelets_;
    std::vector<int> tlOffset_;
    std::vector<int32_t> tlIdx_;
    std::vector<std::array<int, 2>> assign_;
    FBuf stage_, xchg_, sizes_, tlIdxDev_;
    // gravity side: the global tree linked on the device, centres of its leaves / nodes, helpers
    FocusTree<K> g_{CSTONE_SFC_HILBERT, 1u, 1.0f};
    std::vector<K> gPrefixes_, globBnd_;
    FBuf layoutOwn_, gLeafCenters_;
};

} // namespace csb
