/*
 * cstone_oracle.cpp — TEST INFRASTRUCTURE ONLY ("port" oracle).
 *
 * A plain CPU restatement of the reference algorithms on the hot path
 * (SFC keys -> sort -> cornerstone leaf array -> internal tree -> halos -> neighbours).
 * Each function cites the reference file:line it follows (paths relative to
 * /root/reference/include/cstone/).  It is written from the algorithm description,
 * not copied, and is PINNED against
 *   (1) the golden vectors of the reference's own unit tests (tests/test_oracle_kat.py) and
 *   (2) the reference headers themselves compiled in oracle/_ref (tests/test_oracle_vs_ref.py).
 * It exists so the checker travels to the GPU box without /root/reference, and as the
 * "port" cpu_baseline.  The product library never links or loads it.
 *
 * Build: see oracle/Makefile (-O3 -fopenmp -ffp-contract=off, no -march=native, no fast-math).
 */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_PREFIX cstone_oracle_
#include "oracle_api.h"

namespace
{

// ---------------------------------------------------------------------------------------------
// key traits  [ref: tree/definitions.h:45-91]
// ---------------------------------------------------------------------------------------------
template<class K>
struct KT;
template<>
struct KT<uint32_t>
{
    static constexpr unsigned maxLevel = 10; // 3*10 = 30 key bits, 2 unused
    static constexpr unsigned unused   = 2;
    static int clz(uint32_t v) { return v ? __builtin_clz(v) : 32; }
    static int ctz(uint32_t v) { return v ? __builtin_ctz(v) : 32; }
    typedef int32_t Signed;
};
template<>
struct KT<uint64_t>
{
    static constexpr unsigned maxLevel = 21; // 63 key bits, 1 unused
    static constexpr unsigned unused   = 1;
    static int clz(uint64_t v) { return v ? __builtin_clzll(v) : 64; }
    static int ctz(uint64_t v) { return v ? __builtin_ctzll(v) : 64; }
    typedef int64_t Signed;
};

//! key span of one node at `level`  [ref: sfc/common.hpp:130-137 nodeRange]
template<class K>
K nodeRange(unsigned level)
{
    return K(1) << (3 * (KT<K>::maxLevel - level));
}

//! subdivision level of a power-of-8 key range  [ref: sfc/common.hpp:178-183 treeLevel]
template<class K>
unsigned treeLevel(K range)
{
    return (KT<K>::clz(K(range - 1)) - KT<K>::unused) / 3;
}

//! number of leading key bits shared by two keys  [ref: sfc/common.hpp:164-168 commonPrefix]
template<class K>
int commonPrefix(K a, K b)
{
    return KT<K>::clz(a ^ b) - int(KT<K>::unused);
}

//! octal digit at tree level `pos` (1 = most significant)  [ref: sfc/common.hpp:275-279 octalDigit]
template<class K>
unsigned octalDigit(K key, unsigned pos)
{
    return unsigned(key >> (3 * (KT<K>::maxLevel - pos))) & 7u;
}

//! Warren-Salmon placeholder-bit format  [ref: sfc/common.hpp:196-235]
template<class K>
K encodePlaceholderBit(K key, int prefixLength)
{
    int shift = 3 * int(KT<K>::maxLevel) - prefixLength;
    return (K(1) << prefixLength) | (key >> shift);
}
template<class K>
unsigned decodePrefixLength(K prefix)
{
    return 8 * sizeof(K) - 1 - KT<K>::clz(prefix);
}
template<class K>
K decodePlaceholderBit(K prefix)
{
    int len = decodePrefixLength(prefix);
    return K(prefix ^ (K(1) << len)) << (3 * int(KT<K>::maxLevel) - len);
}

// ---------------------------------------------------------------------------------------------
// Morton  [ref: sfc/morton.hpp:52-125]  bit spreading restated as a loop over bits (same result as
// the reference's magic-number version: bit b of v moves to bit 3b)
// ---------------------------------------------------------------------------------------------
template<class K>
K spreadBits3(unsigned v)
{
    K r = 0;
    for (unsigned b = 0; b < KT<K>::maxLevel; ++b)
        r |= K((v >> b) & 1u) << (3 * b);
    return r;
}
template<class K>
unsigned compactBits3(K v)
{
    unsigned r = 0;
    for (unsigned b = 0; b < KT<K>::maxLevel; ++b)
        r |= unsigned((v >> (3 * b)) & 1u) << b;
    return r;
}
template<class K>
K iMorton(unsigned ix, unsigned iy, unsigned iz)
{
    return spreadBits3<K>(ix) * 4 + spreadBits3<K>(iy) * 2 + spreadBits3<K>(iz);
}
template<class K>
void decodeMorton(K key, unsigned* xyz)
{
    xyz[0] = compactBits3<K>(key >> 2);
    xyz[1] = compactBits3<K>(key >> 1);
    xyz[2] = compactBits3<K>(key);
}

// ---------------------------------------------------------------------------------------------
// Hilbert  [ref: sfc/hilbert.hpp:58-109 iHilbert, :145-188 decodeHilbert]
// ---------------------------------------------------------------------------------------------
template<class K>
K iHilbert(unsigned px, unsigned py, unsigned pz)
{
    static const unsigned mortonToHilbert[8] = {0, 1, 3, 2, 7, 6, 4, 5};
    K key                                    = 0;
    for (int level = int(KT<K>::maxLevel) - 1; level >= 0; --level)
    {
        unsigned xi = (px >> level) & 1u, yi = (py >> level) & 1u, zi = (pz >> level) & 1u;
        key = (key << 3) + mortonToHilbert[(xi << 2) | (yi << 1) | zi];

        // conditional complement of ALL remaining bits (the masks are 0 or ~0)
        px ^= -(xi & ((!yi) | zi));
        py ^= -((xi & (yi | zi)) | (yi & (!zi)));
        pz ^= -((xi & (!yi) & (!zi)) | (yi & (!zi)));

        if (zi) // cyclic rotation x <- y <- z <- x
        {
            unsigned t = px;
            px = py, py = pz, pz = t;
        }
        else if (!yi) // exchange x and z
        {
            std::swap(px, pz);
        }
    }
    return key;
}

template<class K>
void decodeHilbert(K key, unsigned* xyz)
{
    unsigned px = 0, py = 0, pz = 0;
    for (unsigned level = 0; level < KT<K>::maxLevel; ++level)
    {
        unsigned octant = unsigned(key >> (3 * level)) & 7u;
        unsigned xi = octant >> 2, yi = (octant >> 1) & 1u, zi = octant & 1u;

        if (yi ^ zi) // inverse cyclic rotation
        {
            unsigned t = px;
            px = pz, pz = py, py = t;
        }
        else if ((!xi & !yi & !zi) || (xi & yi & zi)) { std::swap(px, pz); }

        unsigned mask = (1u << level) - 1;
        px ^= mask & (-(xi & (yi | zi)));
        py ^= mask & (-((xi & ((!yi) | (!zi))) | ((!xi) & yi & zi)));
        pz ^= mask & (-((xi & (!yi) & (!zi)) | (yi & zi)));

        px |= xi << level;
        py |= (xi ^ yi) << level;
        pz |= (yi ^ zi) << level;
    }
    xyz[0] = px, xyz[1] = py, xyz[2] = pz;
}

template<class K>
K encodeSfc(int kind, unsigned ix, unsigned iy, unsigned iz)
{
    return kind == CSTONE_SFC_MORTON ? iMorton<K>(ix, iy, iz) : iHilbert<K>(ix, iy, iz);
}
template<class K>
void decodeSfc(int kind, K key, unsigned* xyz)
{
    if (kind == CSTONE_SFC_MORTON) decodeMorton<K>(key, xyz);
    else decodeHilbert<K>(key, xyz);
}

//! integer box {xmin,xmax,ymin,ymax,zmin,zmax} of the node that starts at `key` on `level`
//! [ref: sfc/hilbert.hpp:274-290 hilbertIBox, sfc/morton.hpp:175-182 mortonIBox]
struct IBox
{
    int v[6];
};
template<class K>
IBox sfcIBox(int kind, K key, unsigned level)
{
    unsigned cube = (1u << KT<K>::maxLevel) >> level;
    unsigned c[3];
    decodeSfc<K>(kind, key, c);
    unsigned mask = ~(cube - 1);
    if (kind != CSTONE_SFC_MORTON) { c[0] &= mask, c[1] &= mask, c[2] &= mask; }
    return IBox{{int(c[0]), int(c[0] + cube), int(c[1]), int(c[1] + cube), int(c[2]), int(c[2] + cube)}};
}

// ---------------------------------------------------------------------------------------------
// box access
// ---------------------------------------------------------------------------------------------
template<class T>
struct BoxOf;
template<>
struct BoxOf<float>
{
    typedef cstone_box_f32 type;
};
template<>
struct BoxOf<double>
{
    typedef cstone_box_f64 type;
};
template<class B>
bool periodic(const B& b, int axis)
{
    return b.boundaries[axis] == CSTONE_BOUNDARY_PERIODIC;
}

// ---------------------------------------------------------------------------------------------
// keys from coordinates  [ref: sfc/sfc.hpp:157-194 sfc3D, :283-291 computeSfcKeys]
//   ix = int( floor(x*mx) - xmin*mx ), mx = 2^maxLevel * (1/lx); evaluated in T without contraction,
//   truncated, clamped to 2^maxLevel-1; keys equal to removeKey = 2^(3*maxLevel) are left alone.
// ---------------------------------------------------------------------------------------------
template<class K, class T>
void computeSfcKeys(int kind, const T* x, const T* y, const T* z, K* keys, size_t n,
                    const typename BoxOf<T>::type* box)
{
    const unsigned cube = 1u << KT<K>::maxLevel;
    const int mcoord    = int(cube) - 1;
    const T mx = cube * box->inv_lengths[0], my = cube * box->inv_lengths[1], mz = cube * box->inv_lengths[2];
    const T xmin = box->limits[0], ymin = box->limits[2], zmin = box->limits[4];
    const K removeKey = nodeRange<K>(0);

#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i)
    {
        if (keys[i] == removeKey) continue;
        int ix  = int(std::floor(x[i] * mx) - xmin * mx);
        int iy  = int(std::floor(y[i] * my) - ymin * my);
        int iz  = int(std::floor(z[i] * mz) - zmin * mz);
        ix      = std::min(ix, mcoord);
        iy      = std::min(iy, mcoord);
        iz      = std::min(iz, mcoord);
        keys[i] = encodeSfc<K>(kind, ix, iy, iz);
    }
}

} // namespace

namespace
{

// ---------------------------------------------------------------------------------------------
// stable sort of (key, value) pairs by key  [ref: primitives/gather.hpp:58-90 sort_by_key]
// ---------------------------------------------------------------------------------------------
template<class K>
void sortByKey(K* keys, uint32_t* values, size_t n)
{
    struct Pair
    {
        K k;
        uint32_t v;
    };
    std::vector<Pair> pairs(n);
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i)
        pairs[i] = Pair{keys[i], values[i]};
    std::stable_sort(pairs.begin(), pairs.end(), [](const Pair& a, const Pair& b) { return a.k < b.k; });
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i)
    {
        keys[i]   = pairs[i].k;
        values[i] = pairs[i].v;
    }
}

//! dst[i] = src[map[i]]  [ref: primitives/gather.hpp:106-119 gather]
void gatherBytes(const uint32_t* map, size_t n, const char* src, char* dst, int elemBytes)
{
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i)
        std::memcpy(dst + i * size_t(elemBytes), src + size_t(map[i]) * size_t(elemBytes), size_t(elemBytes));
}

//! serial in-place exclusive scan; returns the total  [ref: primitives/scan.hpp:106-118]
template<class V>
V exclusiveScanInplace(V* a, size_t n)
{
    V run = 0;
    for (size_t i = 0; i < n; ++i)
    {
        V t  = a[i];
        a[i] = run;
        run += t;
    }
    return run;
}

// ---------------------------------------------------------------------------------------------
// leaf particle counts  [ref: tree/csarray.hpp:93-103 calculateNodeCount, :199-254 computeNodeCounts]
// The reference optionally narrows its binary search with guesses taken from the previous counts
// (findSearchBounds :116-150, updateNodeCount :165-185); the guess never changes the result, which is
// the exact number of keys in [tree[i], tree[i+1]) clamped to maxCount and 0 outside the populated
// leaf range.  The port computes exactly that.
// ---------------------------------------------------------------------------------------------
template<class K>
void computeNodeCounts(const K* tree, uint32_t* counts, int32_t numNodes, const K* first, const K* last,
                       uint32_t maxCount, int /*useCountsAsGuess*/)
{
    int32_t firstNode = numNodes, lastNode = numNodes;
    if (first != last)
    {
        firstNode = int32_t(std::upper_bound(tree, tree + numNodes, *first) - tree) - 1;
        lastNode  = int32_t(std::upper_bound(tree, tree + numNodes, *(last - 1)) - tree);
    }
#pragma omp parallel for schedule(static)
    for (int32_t i = 0; i < numNodes; ++i)
    {
        if (i < firstNode || i >= lastNode)
        {
            counts[i] = 0;
            continue;
        }
        size_t c  = size_t(std::lower_bound(first, last, tree[i + 1]) - std::lower_bound(first, last, tree[i]));
        counts[i] = uint32_t(std::min(c, size_t(maxCount)));
    }
}

// ---------------------------------------------------------------------------------------------
// rebalance decision  [ref: tree/csarray.hpp:269-283 siblingAndLevel, :286-309 calculateNodeOp,
//                      :328-348 rebalanceDecision]
// ---------------------------------------------------------------------------------------------
template<class K>
int nodeOp(const K* tree, int32_t i, const uint32_t* counts, uint32_t bucket)
{
    K start        = tree[i];
    unsigned level = treeLevel<K>(K(tree[i + 1] - start));

    int sibling = -1;
    if (level > 0)
    {
        sibling = int(octalDigit<K>(start, level));
        // all eight children of the parent present at this level?
        if (tree[i - sibling + 8] != K(tree[i - sibling] + nodeRange<K>(level - 1))) sibling = -1;
    }
    if (sibling > 0) // octant 0 stays (it turns into the parent), octants 1-7 vanish
    {
        const uint32_t* g = counts + (i - sibling);
        uint64_t parentCount = 0;
        for (int s = 0; s < 8; ++s)
            parentCount += g[s];
        if (parentCount <= uint64_t(bucket)) return 0;
    }
    const unsigned maxLevel = KT<K>::maxLevel;
    uint32_t c              = counts[i];
    // note: bucket*512 etc. wrap in 32-bit unsigned arithmetic exactly like the reference
    if (c > bucket * 512u && level + 3 < maxLevel) return 4096;
    if (c > bucket * 64u && level + 2 < maxLevel) return 512;
    if (c > bucket * 8u && level + 1 < maxLevel) return 64;
    if (c > bucket && level < maxLevel) return 8;
    return 1;
}

template<class K>
int rebalanceDecision(const K* tree, const uint32_t* counts, int32_t numNodes, uint32_t bucket, int32_t* ops)
{
    int changed = 0;
#pragma omp parallel for schedule(static) reduction(| : changed)
    for (int32_t i = 0; i < numNodes; ++i)
    {
        ops[i] = nodeOp<K>(tree, i, counts, bucket);
        changed |= (ops[i] != 1);
    }
    return !changed;
}

// ---------------------------------------------------------------------------------------------
// emit new leaves  [ref: tree/csarray.hpp:358-385 processNode, :395-409 rebalanceTree]
// ops: numNodes+1 entries, un-scanned on input (last entry ignored), scanned on output
// ---------------------------------------------------------------------------------------------
template<class K>
int32_t rebalanceTree(const K* tree, int32_t numNodes, int32_t* ops, K* newTree, int32_t capacity)
{
    ops[numNodes]   = 0;
    int32_t newSize = exclusiveScanInplace(ops, size_t(numNodes) + 1);
    if (newSize > capacity) return -newSize;

#pragma omp parallel for schedule(static)
    for (int32_t i = 0; i < numNodes; ++i)
    {
        int32_t op = ops[i + 1] - ops[i];
        if (op == 0) continue;
        K start        = tree[i];
        unsigned level = treeLevel<K>(K(tree[i + 1] - start));
        unsigned dl    = 0; // number of levels to descend: op = 8^dl
        for (int32_t t = op; t > 1; t >>= 3)
            ++dl;
        K step = nodeRange<K>(level + dl);
        for (int32_t s = 0; s < op; ++s)
            newTree[ops[i] + s] = start + K(s) * step;
    }
    newTree[newSize] = tree[numNodes];
    return newSize;
}

//! one rebalance + recount step  [ref: tree/csarray.hpp:429-448 updateOctree]
template<class K>
int updateOctree(const K* first, const K* last, uint32_t bucket, uint32_t maxCount, std::vector<K>& tree,
                 std::vector<uint32_t>& counts)
{
    int32_t numNodes = int32_t(tree.size()) - 1;
    std::vector<int32_t> ops(numNodes + 1);
    int converged = rebalanceDecision<K>(tree.data(), counts.data(), numNodes, bucket, ops.data());

    int64_t total = 0;
    for (int32_t i = 0; i < numNodes; ++i)
        total += ops[i];
    std::vector<K> newTree(total + 1);
    rebalanceTree<K>(tree.data(), numNodes, ops.data(), newTree.data(), int32_t(total));
    tree.swap(newTree);

    counts.resize(total);
    computeNodeCounts<K>(tree.data(), counts.data(), int32_t(total), first, last, maxCount, 1);
    return converged;
}

//! build from the root until converged  [ref: tree/csarray.hpp:451-465 computeOctree]
template<class K>
void computeOctree(const K* first, const K* last, uint32_t bucket, uint32_t maxCount, std::vector<K>& tree,
                   std::vector<uint32_t>& counts)
{
    tree   = {K(0), nodeRange<K>(0)};
    counts = {uint32_t(last - first)};
    while (!updateOctree<K>(first, last, bucket, maxCount, tree, counts)) {}
}

} // namespace

namespace
{

// ---------------------------------------------------------------------------------------------
// internal octree  [ref: tree/octree.hpp:72-82 binaryKeyWeight, :95-118 createUnsortedLayoutCpu,
//                   :132-166 linkTreeCpu, :169-178 getLevelRangeCpu, :185-214 buildOctreeCpu]
// ---------------------------------------------------------------------------------------------
//! [ref: sfc/common.hpp:288-292 digitWeight] digits 0..3 weigh -digit, digits 4..7 weigh 7-digit
inline int digitWeight(int digit) { return digit >= 4 ? 7 - digit : -digit; }

template<class K>
void buildOctree(const K* leaves, int32_t L, K* prefixes, int32_t* childOffsets, int32_t* parents,
                 int32_t* levelRange, int32_t* internalToLeaf, int32_t* leafToInternal)
{
    const int32_t I = (L - 1) / 7, M = L + I;
    const unsigned maxLevel = KT<K>::maxLevel;

    // 1. unsorted layout: leaf prefixes behind the internal part; an internal node is born at every leaf whose
    //    common prefix with its successor ends on an octal digit boundary
    std::vector<uint32_t> order(M);
#pragma omp parallel for schedule(static)
    for (int32_t i = 0; i < L; ++i)
    {
        K key          = leaves[i];
        unsigned level = treeLevel<K>(K(leaves[i + 1] - key));
        prefixes[i + I] = encodePlaceholderBit<K>(key, 3 * level);
        order[i + I]    = uint32_t(i + I);

        int cpl = commonPrefix<K>(key, leaves[i + 1]);
        if (cpl % 3 == 0 && i < L - 1)
        {
            int weight = 0;
            for (unsigned l = 1; l <= unsigned(cpl / 3) + 1; ++l)
                weight += digitWeight(int(octalDigit<K>(key, l)));
            int32_t slot   = (i + weight) / 7;
            prefixes[slot] = encodePlaceholderBit<K>(key, cpl);
            order[slot]    = uint32_t(slot);
        }
    }

    // 2. sort by placeholder-bit key: level-major, SFC order inside each level
    sortByKey<K>(prefixes, order.data(), size_t(M));

    // 3. both direction maps
#pragma omp parallel for schedule(static)
    for (int32_t i = 0; i < M; ++i)
    {
        leafToInternal[order[i]] = i;
        internalToLeaf[i]        = int32_t(order[i]) - I;
    }

    // 4. first node of each level
    for (unsigned l = 0; l <= maxLevel; ++l)
        levelRange[l] = int32_t(std::lower_bound(prefixes, prefixes + M, K(K(1) << (3 * l))) - prefixes);
    levelRange[maxLevel + 1] = M;

    // 5. child / parent links of the internal nodes
    std::fill(childOffsets, childOffsets + M, 0);
#pragma omp parallel for schedule(static)
    for (int32_t i = 0; i < I; ++i)
    {
        int32_t node    = leafToInternal[i];
        K prefix        = prefixes[node];
        unsigned len    = decodePrefixLength<K>(prefix);
        unsigned level  = len / 3;
        K childPrefix   = encodePlaceholderBit<K>(decodePlaceholderBit<K>(prefix), int(len) + 3);
        const K* lo     = prefixes + levelRange[level + 1];
        const K* hi     = prefixes + levelRange[level + 2];
        const K* it     = std::lower_bound(lo, hi, childPrefix);
        if (it != hi && *it == childPrefix)
        {
            int32_t child            = int32_t(it - prefixes);
            childOffsets[node]       = child;
            parents[(child - 1) / 8] = node;
        }
    }
}

//! saturating sum of the 8 children, deepest level first  [ref: tree/octree.hpp:583-626 upsweep + NodeCount]
void upsweepSum(int32_t numLevels, const int32_t* levelRange, const int32_t* childOffsets, uint32_t* counts)
{
    for (int32_t level = numLevels - 1; level >= 0; --level)
    {
#pragma omp parallel for schedule(static)
        for (int32_t i = levelRange[level]; i < levelRange[level + 1]; ++i)
        {
            int32_t c = childOffsets[i];
            if (!c) continue;
            uint64_t sum = 0;
            for (int o = 0; o < 8; ++o)
                sum += counts[c + o];
            counts[i] = uint32_t(std::min<uint64_t>(sum, 0xFFFFFFFFull));
        }
    }
}

// ---------------------------------------------------------------------------------------------
// geometric centres  [ref: focus/source_center.hpp:145-157 nodeFpCenters, sfc/box.hpp:334-351 centerAndSize]
// ---------------------------------------------------------------------------------------------
template<class K, class T>
void geoCenters(int kind, const K* prefixes, int32_t n, T* centers, T* sizes, const typename BoxOf<T>::type* box)
{
    const T uL = T(1.) / T(1u << KT<K>::maxLevel); // the reference divides by an int constant: exact power of two
#pragma omp parallel for schedule(static)
    for (int32_t i = 0; i < n; ++i)
    {
        K prefix       = prefixes[i];
        unsigned level = decodePrefixLength<K>(prefix) / 3;
        IBox b         = sfcIBox<K>(kind, decodePlaceholderBit<K>(prefix), level);
        for (int d = 0; d < 3; ++d)
        {
            T halfUnit         = T(0.5) * uL * box->lengths[d];
            centers[3 * i + d] = box->limits[2 * d] + (b.v[2 * d + 1] + b.v[2 * d]) * halfUnit;
            sizes[3 * i + d]   = (b.v[2 * d + 1] - b.v[2 * d]) * halfUnit;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// halo radii  [ref: halos/halos.hpp:148-160 (segmentMax + scaleGpu), primitives_gpu.cu:240-259]
// ---------------------------------------------------------------------------------------------
template<class Th>
void haloRadii(const Th* h, const uint32_t* seg, size_t numSegments, float scale, float* radii)
{
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < numSegments; ++i)
    {
        Th m = 0;
        for (uint32_t j = seg[i]; j < seg[i + 1]; ++j)
            m = std::max(m, h[j]);
        radii[i] = float(m) * scale;
    }
}

// ---------------------------------------------------------------------------------------------
// halo discovery  [ref: traversal/collisions.hpp:39-105, traversal/traversal.hpp:68-110,
//                  traversal/boxoverlap.hpp:44-83 overlap, :93-116 containedIn, :145-182 makeHaloBox,
//                  sfc/common.hpp:80-90 toNBitIntCeil, :312-319 smallestCommonBox]
// ---------------------------------------------------------------------------------------------
inline bool overlapTwoRanges(int a, int b, int c, int d) { return b > c && d > a; }
//! periodic interval test with the three images the reference checks
inline bool overlapRange(int R, int a, int b, int c, int d)
{
    return overlapTwoRanges(a, b, c, d) || overlapTwoRanges(a + R, b + R, c, d) || overlapTwoRanges(a, b, c + R, d + R);
}
template<class K>
bool overlap(const IBox& a, const IBox& b)
{
    const int R = 1 << KT<K>::maxLevel;
    return overlapRange(R, a.v[0], a.v[1], b.v[0], b.v[1]) && overlapRange(R, a.v[2], a.v[3], b.v[2], b.v[3]) &&
           overlapRange(R, a.v[4], a.v[5], b.v[4], b.v[5]);
}

template<class K, class T>
IBox makeHaloBox(int kind, K start, K end, float radius, const typename BoxOf<T>::type* box)
{
    const int maxCoord = 1 << KT<K>::maxLevel;
    IBox node          = sfcIBox<K>(kind, start, treeLevel<K>(K(end - start)));
    IBox halo;
    for (int d = 0; d < 3; ++d)
    {
        // float radius times T inverse length, evaluated in T; ceil; clamp to maxCoord-1
        T scaled        = radius * box->inv_lengths[d];
        unsigned cells  = unsigned(std::ceil(scaled * T(1u << KT<K>::maxLevel)));
        int delta       = int(std::min(cells, unsigned(maxCoord - 1)));
        int lo          = node.v[2 * d] - delta;
        int hi          = node.v[2 * d + 1] + delta;
        if (!periodic(*box, d))
        {
            lo = std::min(std::max(0, lo), maxCoord);
            hi = std::min(std::max(0, hi), maxCoord);
        }
        halo.v[2 * d]     = lo;
        halo.v[2 * d + 1] = hi;
    }
    return halo;
}

//! is the integer box inside the key range [codeStart, codeEnd) ?  (curve is ALWAYS Hilbert in the reference)
template<class K>
bool containedIn(int kind, K codeStart, K codeEnd, const IBox& b)
{
    const int R = 1 << KT<K>::maxLevel;
    if (std::min(std::min(b.v[0], b.v[2]), b.v[4]) < 0 || std::max(std::max(b.v[1], b.v[3]), b.v[5]) > R)
        return codeStart == 0 && codeEnd == nodeRange<K>(0);

    K lowCode  = encodeSfc<K>(kind, b.v[0], b.v[2], b.v[4]);
    K highCode = encodeSfc<K>(kind, b.v[1] - 1, b.v[3] - 1, b.v[5] - 1);
    unsigned commonLevel = unsigned(commonPrefix<K>(lowCode, highCode)) / 3;
    K envStart = lowCode & ~K(nodeRange<K>(commonLevel) - 1);
    K envEnd   = envStart + nodeRange<K>(commonLevel);
    return envStart >= codeStart && envEnd <= codeEnd;
}

//! depth-first walk in the reference's order: the 8 children of the current node are tested in octant order,
//! leaf children are handled at once, internal children are pushed and later popped last-in-first-out
template<class Descend, class Leaf>
void singleTraversal(const int32_t* childOffsets, Descend&& descend, Leaf&& leafAction)
{
    if (!descend(0)) return;
    if (childOffsets[0] == 0)
    {
        leafAction(0);
        return;
    }
    int32_t stack[128];
    int sp      = 0;
    stack[sp++] = 0; // sentinel: popping the root ends the walk
    int32_t node = 0;
    do
    {
        for (int o = 0; o < 8; ++o)
        {
            int32_t child = childOffsets[node] + o;
            if (!descend(child)) continue;
            if (childOffsets[child] == 0) leafAction(child);
            else stack[sp++] = child;
        }
        node = stack[--sp];
    } while (node != 0);
}

template<class K, class T>
void findHalos(int kind, const K* prefixes, const int32_t* childOffsets, const int32_t* internalToLeaf,
               const K* leaves, const float* radii, const typename BoxOf<T>::type* box, int32_t firstNode,
               int32_t lastNode, int32_t* flags)
{
    const K lowest = leaves[firstNode], highest = leaves[lastNode];
#pragma omp parallel for schedule(dynamic, 64)
    for (int32_t leaf = firstNode; leaf < lastNode; ++leaf)
    {
        IBox halo = makeHaloBox<K, T>(kind, leaves[leaf], leaves[leaf + 1], radii[leaf], box);
        if (containedIn<K>(kind, lowest, highest, halo)) continue;

        auto descend = [&](int32_t idx)
        {
            K prefix       = prefixes[idx];
            unsigned level = decodePrefixLength<K>(prefix) / 3;
            K nodeStart    = decodePlaceholderBit<K>(prefix);
            K nodeEnd      = nodeStart + nodeRange<K>(level);
            bool inOwn     = !(nodeStart < lowest || nodeEnd > highest);
            return !inOwn && overlap<K>(sfcIBox<K>(kind, nodeStart, level), halo);
        };
        auto mark = [&](int32_t idx) { flags[internalToLeaf[idx]] = 1; };
        singleTraversal(childOffsets, descend, mark);
    }
}

//! [ref: domain/layout.hpp:150-164 computeNodeLayout]
void computeNodeLayout(const uint32_t* counts, const int32_t* flags, int32_t first, int32_t last, int32_t L,
                       uint32_t* layout)
{
    for (int32_t i = 0; i < L; ++i)
        layout[i] = ((first <= i && i < last) || flags[i]) ? counts[i] : 0;
    layout[L] = 0;
    exclusiveScanInplace(layout, size_t(L) + 1);
}

} // namespace

namespace
{

// ---------------------------------------------------------------------------------------------
// neighbour search  [ref: findneighbors.hpp:51-78 distanceSq, :95-165 findNeighbors (one particle),
//                    :167-188 driver; traversal/boxoverlap.hpp:185-217 insideBox / minDistance;
//                    sfc/box.hpp:194-206 applyPbc; util/array.hpp:253-256 (norm2 is a RIGHT fold:
//                    x*x + (y*y + z*z))]
// ---------------------------------------------------------------------------------------------
template<class T>
void findNeighbors(const T* x, const T* y, const T* z, const T* h, uint32_t firstId, uint32_t lastId,
                   const typename BoxOf<T>::type* box, const int32_t* childOffsets, const int32_t* internalToLeaf,
                   const uint32_t* layout, const T* centers, const T* sizes, float ext, uint32_t ngmax,
                   uint32_t* neighbors, uint32_t* neighborsCount)
{
    const bool pbc[3] = {periodic(*box, 0), periodic(*box, 1), periodic(*box, 2)};
    const bool anyPbc = pbc[0] || pbc[1] || pbc[2];
    const T* L        = box->lengths;
    const T* iL       = box->inv_lengths;

#pragma omp parallel for schedule(dynamic, 256)
    for (uint32_t w = 0; w < lastId - firstId; ++w)
    {
        const uint32_t i = w + firstId;
        const T p[3]     = {x[i], y[i], z[i]};
        const T hi       = h[i];
        const T radiusSq = T(4.0) * hi * hi;
        const T cellRadiusSq = radiusSq * ext * ext;
        uint32_t* out    = neighbors + size_t(w) * ngmax;
        uint32_t found   = 0;

        // PBC path only if the 2h-cube around the particle pokes out of the box
        bool usePbc = false;
        if (anyPbc)
        {
            T r2h       = T(2) * hi;
            bool inside = true;
            for (int d = 0; d < 3; ++d)
                inside = inside && (p[d] - r2h >= box->limits[2 * d]) && (p[d] + r2h <= box->limits[2 * d + 1]);
            usePbc = !inside;
        }

        auto descend = [&](int32_t idx)
        {
            T d[3];
            for (int k = 0; k < 3; ++k)
            {
                T dx = centers[3 * idx + k] - p[k];
                if (usePbc) dx -= T(pbc[k]) * L[k] * std::rint(dx * iL[k]);
                dx = std::abs(dx) - sizes[3 * idx + k];
                dx += std::abs(dx);
                dx *= T(0.5);
                d[k] = dx;
            }
            return d[0] * d[0] + (d[1] * d[1] + d[2] * d[2]) < cellRadiusSq;
        };
        auto leafAction = [&](int32_t idx)
        {
            int32_t leaf = internalToLeaf[idx];
            for (uint32_t j = layout[leaf]; j < layout[leaf + 1]; ++j)
            {
                if (j == i) continue;
                T dx = x[j] - p[0], dy = y[j] - p[1], dz = z[j] - p[2];
                if (usePbc)
                {
                    dx -= T(pbc[0]) * L[0] * std::rint(dx * iL[0]);
                    dy -= T(pbc[1]) * L[1] * std::rint(dy * iL[1]);
                    dz -= T(pbc[2]) * L[2] * std::rint(dz * iL[2]);
                }
                if (dx * dx + dy * dy + dz * dz < radiusSq)
                {
                    if (found < ngmax) out[found] = j;
                    ++found;
                }
            }
        };
        singleTraversal(childOffsets, descend, leafAction);
        neighborsCount[w] = found;
    }
}

template<class K>
int updateOctreeC(const K* f, const K* l, uint32_t bucket, uint32_t maxCount, K* tree, uint32_t* counts,
                  int32_t capacity, int32_t* numNodes)
{
    std::vector<K> t(tree, tree + *numNodes + 1);
    std::vector<uint32_t> c(counts, counts + *numNodes);
    int converged = updateOctree<K>(f, l, bucket, maxCount, t, c);
    int32_t newN  = int32_t(t.size()) - 1;
    *numNodes     = newN;
    if (newN > capacity) return -1;
    std::copy(t.begin(), t.end(), tree);
    std::copy(c.begin(), c.end(), counts);
    return converged;
}

template<class K>
int32_t computeOctreeC(const K* f, const K* l, uint32_t bucket, uint32_t maxCount, K* tree, uint32_t* counts,
                       int32_t capacity)
{
    std::vector<K> t;
    std::vector<uint32_t> c;
    computeOctree<K>(f, l, bucket, maxCount, t, c);
    int32_t newN = int32_t(t.size()) - 1;
    if (newN > capacity) return -newN;
    std::copy(t.begin(), t.end(), tree);
    std::copy(c.begin(), c.end(), counts);
    return newN;
}

template<class K>
void iboxC(int kind, K key, unsigned level, int* out)
{
    IBox b = sfcIBox<K>(kind, key, level);
    std::copy(b.v, b.v + 6, out);
}

} // namespace

// ---------------------------------------------------------------------------------------------
// C exports (oracle_api.h)
// ---------------------------------------------------------------------------------------------
extern "C" {

int cstone_oracle_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void cstone_oracle_set_num_threads(int n)
{
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

uint32_t cstone_oracle_encode_k32(int kind, unsigned ix, unsigned iy, unsigned iz) { return encodeSfc<uint32_t>(kind, ix, iy, iz); }
uint64_t cstone_oracle_encode_k64(int kind, unsigned ix, unsigned iy, unsigned iz) { return encodeSfc<uint64_t>(kind, ix, iy, iz); }
void cstone_oracle_decode_k32(int kind, uint32_t key, unsigned* xyz) { decodeSfc<uint32_t>(kind, key, xyz); }
void cstone_oracle_decode_k64(int kind, uint64_t key, unsigned* xyz) { decodeSfc<uint64_t>(kind, key, xyz); }

#define SFCKEYS(KS, K, TS, T)                                                                                          \
    void cstone_oracle_compute_sfc_keys_##KS##_##TS(int kind, const T* x, const T* y, const T* z, K* keys, size_t n,   \
                                                    const cstone_box_##TS* box)                                        \
    {                                                                                                                  \
        computeSfcKeys<K, T>(kind, x, y, z, keys, n, box);                                                             \
    }
SFCKEYS(k32, uint32_t, f32, float)
SFCKEYS(k32, uint32_t, f64, double)
SFCKEYS(k64, uint64_t, f32, float)
SFCKEYS(k64, uint64_t, f64, double)

void cstone_oracle_sfc_ibox_k32(int kind, uint32_t key, unsigned level, int* b) { iboxC<uint32_t>(kind, key, level, b); }
void cstone_oracle_sfc_ibox_k64(int kind, uint64_t key, unsigned level, int* b) { iboxC<uint64_t>(kind, key, level, b); }

void cstone_oracle_sort_by_key_k32(uint32_t* keys, uint32_t* values, size_t n) { sortByKey(keys, values, n); }
void cstone_oracle_sort_by_key_k64(uint64_t* keys, uint32_t* values, size_t n) { sortByKey(keys, values, n); }
void cstone_oracle_gather(const uint32_t* map, size_t n, const void* src, void* dst, int elemBytes)
{
    gatherBytes(map, n, (const char*)src, (char*)dst, elemBytes);
}

#define TREEFUNCS(KS, K)                                                                                               \
    void cstone_oracle_compute_node_counts_##KS(const K* tree, uint32_t* counts, int32_t n, const K* f, const K* l,    \
                                                uint32_t maxCount, int guess)                                          \
    {                                                                                                                  \
        computeNodeCounts<K>(tree, counts, n, f, l, maxCount, guess);                                                  \
    }                                                                                                                  \
    int cstone_oracle_rebalance_decision_##KS(const K* tree, const uint32_t* counts, int32_t n, uint32_t bucket,       \
                                              int32_t* ops)                                                            \
    {                                                                                                                  \
        return rebalanceDecision<K>(tree, counts, n, bucket, ops);                                                     \
    }                                                                                                                  \
    int32_t cstone_oracle_rebalance_tree_##KS(const K* tree, int32_t n, int32_t* ops, K* newTree, int32_t cap)         \
    {                                                                                                                  \
        return rebalanceTree<K>(tree, n, ops, newTree, cap);                                                           \
    }                                                                                                                  \
    int cstone_oracle_update_octree_##KS(const K* f, const K* l, uint32_t bucket, uint32_t maxCount, K* tree,          \
                                         uint32_t* counts, int32_t cap, int32_t* n)                                    \
    {                                                                                                                  \
        return updateOctreeC<K>(f, l, bucket, maxCount, tree, counts, cap, n);                                         \
    }                                                                                                                  \
    int32_t cstone_oracle_compute_octree_##KS(const K* f, const K* l, uint32_t bucket, uint32_t maxCount, K* tree,     \
                                              uint32_t* counts, int32_t cap)                                           \
    {                                                                                                                  \
        return computeOctreeC<K>(f, l, bucket, maxCount, tree, counts, cap);                                           \
    }                                                                                                                  \
    void cstone_oracle_build_octree_##KS(const K* leaves, int32_t L, K* p, int32_t* co, int32_t* pa, int32_t* lr,      \
                                         int32_t* i2l, int32_t* l2i)                                                   \
    {                                                                                                                  \
        buildOctree<K>(leaves, L, p, co, pa, lr, i2l, l2i);                                                            \
    }
TREEFUNCS(k32, uint32_t)
TREEFUNCS(k64, uint64_t)

/* [ref: sfc/common.hpp:392-430 spanSfcRange, :339-346 lastNzPlace]: walk up from a in growing powers of 8 until the
 * first octal place where a and b differ, then down towards b in shrinking powers of 8 */
extern "C++" {
template<class K>
static int spanSfcRangeT(K a, K b, K* out)
{
    auto lastNz = [](K x) { return x ? int(KT<K>::maxLevel) - KT<K>::ctz(x) / 3 : int(KT<K>::maxLevel); };
    int n             = 0;
    const int diffPos = (KT<K>::clz(K(a ^ b)) + 3 - int(KT<K>::unused)) / 3;
    const int aLast = lastNz(a), bLast = lastNz(b);
    for (int pos = aLast; pos > diffPos; --pos)
        for (int d = (8 - int(octalDigit<K>(a, unsigned(pos)))) % 8; d > 0; --d, ++n)
        {
            if (out) out[n] = a;
            a += nodeRange<K>(unsigned(pos));
        }
    for (int pos = diffPos; pos <= bLast; ++pos)
        for (int d = int(octalDigit<K>(b, unsigned(pos))) - int(octalDigit<K>(a, unsigned(pos))); d > 0; --d, ++n)
        {
            if (out) out[n] = a;
            a += nodeRange<K>(unsigned(pos));
        }
    return n;
}
}
int cstone_oracle_span_sfc_range_k32(uint32_t a, uint32_t b, uint32_t* out) { return spanSfcRangeT<uint32_t>(a, b, out); }
int cstone_oracle_span_sfc_range_k64(uint64_t a, uint64_t b, uint64_t* out) { return spanSfcRangeT<uint64_t>(a, b, out); }

void cstone_oracle_upsweep_sum(int32_t numLevels, const int32_t* lr, const int32_t* co, uint32_t* counts)
{
    upsweepSum(numLevels, lr, co, counts);
}

#define GEOFUNCS(KS, K, TS, T)                                                                                         \
    void cstone_oracle_geo_centers_##KS##_##TS(int kind, const K* p, int32_t n, T* c, T* s,                            \
                                               const cstone_box_##TS* box)                                             \
    {                                                                                                                  \
        geoCenters<K, T>(kind, p, n, c, s, box);                                                                       \
    }                                                                                                                  \
    void cstone_oracle_find_halos_##KS##_##TS(int kind, const K* p, const int32_t* co, const int32_t* i2l,             \
                                              const K* leaves, const float* radii, const cstone_box_##TS* box,         \
                                              int32_t first, int32_t last, int32_t* flags)                             \
    {                                                                                                                  \
        findHalos<K, T>(kind, p, co, i2l, leaves, radii, box, first, last, flags);                                     \
    }
GEOFUNCS(k32, uint32_t, f32, float)
GEOFUNCS(k32, uint32_t, f64, double)
GEOFUNCS(k64, uint64_t, f32, float)
GEOFUNCS(k64, uint64_t, f64, double)

void cstone_oracle_halo_radii_f32(const float* h, const uint32_t* seg, size_t n, float scale, float* r) { haloRadii(h, seg, n, scale, r); }
void cstone_oracle_halo_radii_f64(const double* h, const uint32_t* seg, size_t n, float scale, float* r) { haloRadii(h, seg, n, scale, r); }

void cstone_oracle_compute_node_layout(const uint32_t* counts, const int32_t* flags, int32_t first, int32_t last,
                                       int32_t L, uint32_t* layout)
{
    computeNodeLayout(counts, flags, first, last, L, layout);
}

#define NBFUNCS(TS, T)                                                                                                 \
    void cstone_oracle_find_neighbors_##TS(const T* x, const T* y, const T* z, const T* h, uint32_t f, uint32_t l,     \
                                           const cstone_box_##TS* box, const int32_t* co, const int32_t* i2l,          \
                                           const uint32_t* layout, const T* c, const T* s, float ext, uint32_t ngmax,  \
                                           uint32_t* nb, uint32_t* nc)                                                 \
    {                                                                                                                  \
        findNeighbors<T>(x, y, z, h, f, l, box, co, i2l, layout, c, s, ext, ngmax, nb, nc);                            \
    }
NBFUNCS(f32, float)
NBFUNCS(f64, double)

} // extern "C"
