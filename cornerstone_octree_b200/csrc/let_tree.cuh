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
#include <cstring>
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
        leaves_ = {K(0), nodeRange<K>(0)};
        leafCounts_ = {bucket + 1};
    }

    FocusTree<K> t; // leaves, linked tree, counts, macs: device

    const std::vector<K>& leaves() const { return leaves_; }
    const std::vector<uint32_t>& leafCounts() const { return leafCounts_; }
    const std::vector<std::array<int, 2>>& assignment() const { return assign_; }
    const std::vector<int>& peers() const { return peers_; }
    int numLeaves() const { return int(leaves_.size()) - 1; }

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

        bool conv = false;
        CSB_TRY(t.rebalanceStep(focusStart, focusEnd, enforced.data(), int(enforced.size()), &conv, s));
        const float invThetaRefine = float(std::sqrt(3) / 2 + 1e-6);
        for (int c = 0; !c;)
            CSB_TRY(t.template macRefine<T>(&box, prevStart_, prevEnd_, focusStart, focusEnd, invThetaRefine, &c, s));

        CSB_TRY(downloadLeaves(s));
        translateAssignment(bnd);
        CSB_TRY(syncTreelets(comm, s));
        CSB_TRY(indexTreelets(s));
        translateAssignment(bnd);

        box_       = box;
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
        std::vector<uint32_t> idx, val;
        {
            int current = 0;
            auto fill = [&](int a, int b)
            {
                for (int i = a; i < b; ++i)
                {
                    const int gs = findNodeBelow(gLeaves.data(), int(gLeaves.size()), leaves_[i]);
                    const int ge = findNodeAbove(gLeaves.data(), int(gLeaves.size()), leaves_[i + 1]);
                    uint64_t sum = 0;
                    for (int g = gs; g < ge; ++g)
                        sum += gCounts[g];
                    idx.push_back(uint32_t(i));
                    val.push_back(uint32_t(std::min<uint64_t>(0xFFFFFFFFull, sum)));
                }
            };
            for (const auto& r : assign_)
            {
                if (r[0] == r[1]) continue;
                if (current < r[0]) fill(current, r[0]);
                current = r[1];
            }
            if (current < L) fill(current, L);
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
            for (int i = 0; i < numNodes; ++i)
            {
                const K k = tl[i];
                if (k != leaves_[findNodeAbove(leaves_.data(), nk - 1, k)]) rej.push_back(k);
                else kept.push_back(k);
            }
            if (!tl.empty()) kept.push_back(tl.back());
            treelets_[peer] = std::move(kept);
            rejected[peer]  = asBytes(rej);
        }
        CSB_TRY(exchangeBytes(comm, rejected, rejectedHere, s));
        std::vector<char> keep(size_t(nk), 1);
        bool any = false;
        for (int peer : peers_)
            for (K k : asKeys(rejectedHere[peer]))
            {
                keep[findNodeAbove(leaves_.data(), nk, k)] = 0;
                any = true;
            }
        if (any)
        {
            std::vector<K> nl;
            nl.reserve(size_t(nk));
            for (int i = 0; i + 1 < nk; ++i)
                if (keep[i]) nl.push_back(leaves_[i]);
            nl.push_back(leaves_.back());
            leaves_.swap(nl);
            CSB_TRY(uploadLeaves(s));
        }
        return 0;
    }

    //! node index in MY linked tree of every node the peers hold of my range
    //! [ref: focus/exchange_focus.hpp:237-262 indexTreelets, tree/octree.hpp locateNode]
    int indexTreelets(cudaStream_t s)
    {
        const int M = t.numNodes_;
        prefixes_.resize(size_t(M));
        CSB_CUDA(cudaMemcpyAsync(prefixes_.data(), t.prefixes_.p, size_t(M) * sizeof(K), cudaMemcpyDeviceToHost, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        std::fill(tlOffset_.begin(), tlOffset_.end(), 0);
        for (int r = 0; r < P_; ++r)
            tlOffset_[r + 1] = tlOffset_[r] + (treelets_[r].empty() ? 0 : int(treelets_[r].size()) - 1);
        tlIdx_.assign(size_t(tlOffset_[P_]), 0);
        for (int peer : peers_)
        {
            const std::vector<K>& tl = treelets_[peer];
            for (int i = 0; i + 1 < int(tl.size()); ++i)
            {
                const unsigned level = treeLevel<K>(tl[i + 1] - tl[i]);
                const K prefix       = encodePlaceholderBit<K>(tl[i], 3 * int(level));
                const K* first       = prefixes_.data() + t.levelRange_[level];
                const K* last        = prefixes_.data() + t.levelRange_[level + 1];
                const K* it          = std::lower_bound(first, last, prefix);
                if (it == last || *it != prefix) return CSTONE_E_LET;
                tlIdx_[size_t(tlOffset_[peer]) + i] = uint32_t(it - prefixes_.data());
            }
        }
        CSB_TRY(tlIdxDev_.reserve(4 * tlIdx_.size() + 256));
        if (!tlIdx_.empty())
            CSB_CUDA(cudaMemcpyAsync(tlIdxDev_.p, tlIdx_.data(), 4 * tlIdx_.size(), cudaMemcpyHostToDevice, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        return 0;
    }

    int rank_, P_;
    unsigned bucket_;
    int kind_;
    CBox box_;
    K prevStart_ = 0, prevEnd_ = 0;
    std::vector<K> leaves_, prefixes_;
    std::vector<uint32_t> leafCounts_;
    std::vector<int> peers_;
    std::vector<std::vector<K>> treelets_;
    std::vector<int> tlOffset_;
    std::vector<uint32_t> tlIdx_;
    std::vector<std::array<int, 2>> assign_;
    FBuf stage_, xchg_, sizes_, tlIdxDev_;
};

} // namespace csb
