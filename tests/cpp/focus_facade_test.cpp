// Target groups and the single-rank focus tree through include/cstone_b200/focus.hpp, on the known answers of the
// reference's own tests [ref: test/unit_cuda/traversal/groups.cu (t0, groupVolumes),
// test/unit/focus/octree_focus.cpp:339-421 (computeEssentialTree), :457-470 (MacRefinement.fullSurface)].
// Plain C++17, no CUDA headers.  Usage: focus_facade_test run
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <numeric>
#include <vector>

#include "cstone_b200/focus.hpp"

using namespace cstone;

static int failures = 0;
#define EXPECT(cond)                                                                                                   \
    do                                                                                                                 \
    {                                                                                                                  \
        if (!(cond))                                                                                                   \
        {                                                                                                              \
            std::printf("FAILED %s:%d  %s\n", __FILE__, __LINE__, #cond);                                              \
            ++failures;                                                                                                \
        }                                                                                                              \
    } while (0)

template<class K>
constexpr int maxTreeLevel()
{
    return sizeof(K) == 4 ? 10 : 21;
}
template<class K>
constexpr K nodeRange(int level)
{
    return K(1) << (3 * (maxTreeLevel<K>() - level));
}
template<class K>
constexpr K pad(K prefix, int length)
{
    return prefix << (3 * maxTreeLevel<K>() - length);
}

static void fixedGroups()
{
    GroupData<GpuTag> groups;
    computeFixedGroups(4, 34, 8, groups);
    std::vector<LocalIndex> h(groups.data.size());
    cstone_b200_device_memcpy(h.data(), groups.data.data(), h.size() * sizeof(LocalIndex), 1);
    EXPECT((h == std::vector<LocalIndex>{4, 12, 20, 28, 34}));
    EXPECT(groups.view().numGroups == 4 && groups.view().groupEnd == groups.view().groupStart + 1);
}

static void groupVolumes()
{
    using T = double;
    using K = uint64_t;
    LocalIndex first = 4, last = 128;
    Box<T> box(0, last);
    double distCrit = std::cbrt(box.lx() * box.ly() * box.lz() / 64);

    // OctreeMaker{}.divide().divide(2): 15 leaves
    std::vector<K> leaves{0, nodeRange<K>(1), 2 * nodeRange<K>(1)};
    for (int i = 1; i < 8; ++i)
        leaves.push_back(2 * nodeRange<K>(1) + i * nodeRange<K>(2));
    for (int i = 3; i <= 8; ++i)
        leaves.push_back(i * nodeRange<K>(1));
    std::vector<unsigned> counts{4, 1, 8, 8, 8, 8, 31, 8, 8, 8, 16, 16, 16, 0, 0};
    std::vector<LocalIndex> layout(counts.size() + 1, 0);
    std::inclusive_scan(counts.begin(), counts.end(), layout.begin() + 1);

    std::vector<T> x(last);
    std::iota(x.begin(), x.end(), T(0));
    x[5] -= 0.01;
    DeviceVector<T> dx(x), dh(std::vector<T>(last, 0.0));
    DeviceVector<K> dLeaves(leaves);
    DeviceVector<LocalIndex> dLayout(layout), temp, groups;

    float tolFactor = std::sqrt(3.0) / distCrit * 1.01;
    computeGroupSplits(first, last, dx.data(), dx.data(), dx.data(), dh.data(), dLeaves.data(),
                       TreeNodeIndex(leaves.size() - 1), dLayout.data(), box, 64, tolFactor, temp, groups);
    std::vector<LocalIndex> h(groups.size());
    cstone_b200_device_memcpy(h.data(), groups.data(), h.size() * sizeof(LocalIndex), 1);
    EXPECT((h == std::vector<LocalIndex>{4, 6, 68, 128}));

    bool threw = false;
    try
    {
        computeGroupSplits(first, last, dx.data(), dx.data(), dx.data(), dh.data(), dLeaves.data(),
                           TreeNodeIndex(leaves.size() - 1), dLayout.data(), box, 128, tolFactor, temp, groups);
    }
    catch (std::runtime_error&)
    {
        threw = true;
    }
    EXPECT(threw);
}

template<class K>
static std::vector<int> octantNodeCount(const std::vector<K>& tree)
{
    std::vector<int> counts;
    for (int octant = 0; octant < 8; ++octant)
    {
        K range  = nodeRange<K>(1);
        auto ita = std::lower_bound(tree.begin(), tree.end(), K(octant * range));
        auto itb = std::lower_bound(tree.begin(), tree.end(), K(octant * range + range));
        EXPECT(*ita == K(octant * range) && *itb == K(octant * range + range));
        counts.push_back(int(itb - ita));
    }
    std::sort(counts.begin(), counts.end());
    counts.pop_back();
    return counts;
}

template<class K>
static void computeEssentialTree()
{
    Box<double> box{-1, 1};
    int nParticles = 200000;
    std::vector<K> codes(nParticles);
    std::iota(codes.begin(), codes.end(), 0);
    std::for_each(codes.begin(), codes.end(), [n = nParticles](auto& k) { k *= double(nodeRange<K>(0)) / n; });
    DeviceVector<K> dCodes(codes);

    FocusedOctreeSingleNode<K> tree(16, 0.9f);
    std::vector<int> refCounts{29, 302, 302, 302, 1184, 1184, 1184};

    K focusStart = 1;
    K focusEnd   = pad(K(1), 3);
    tree.update(box, dCodes, focusStart, focusEnd);
    EXPECT(tree.treeLeaves().toHost()[1] == focusStart);
    int steps = 0;
    while (!tree.update(box, dCodes, focusStart, focusEnd) && ++steps < 50) {}
    EXPECT(steps < 50);
    EXPECT(octantNodeCount(tree.treeLeaves().toHost()) == refCounts);

    focusStart = pad(K(7), 3);
    focusEnd   = nodeRange<K>(0) - 1;
    while (!tree.update(box, dCodes, focusStart, focusEnd) && ++steps < 100) {}
    EXPECT(octantNodeCount(tree.treeLeaves().toHost()) == refCounts);

    focusStart = 0;
    focusEnd   = pad(K(1), 3);
    while (!tree.update(box, dCodes, focusStart, focusEnd) && ++steps < 150) {}
    EXPECT(steps < 150);
    EXPECT(octantNodeCount(tree.treeLeaves().toHost()) == refCounts);

    auto counts = tree.leafCounts().toHost();
    EXPECT(std::accumulate(counts.begin(), counts.end(), 0ul) == std::size_t(nParticles));
}

static void macRefinementFullSurface()
{
    using K = uint64_t;
    // L3 in the first octant (L4 in its first cell), L1 otherwise
    std::vector<K> leaves;
    for (int i = 0; i <= 8; ++i)
        leaves.push_back(i * nodeRange<K>(1));
    for (int i = 0; i < 8; ++i)
        leaves.push_back(i * nodeRange<K>(2));
    for (int i = 0; i < 64; ++i)
        leaves.push_back(i * nodeRange<K>(3));
    for (int i = 0; i < 8; ++i)
        leaves.push_back(i * nodeRange<K>(4));
    std::sort(leaves.begin(), leaves.end());
    leaves.erase(std::unique(leaves.begin(), leaves.end()), leaves.end());
    EXPECT(leaves.size() - 1 == 78);
    DeviceVector<K> dLeaves(leaves);

    FocusedOctreeSingleNode<K> tree(16, 0.5f);
    tree.setLeaves(dLeaves.data(), TreeNodeIndex(leaves.size() - 1));
    Box<double> box(0, 1);
    float invTheta = std::sqrt(3) / 2 + 1e-6;
    K focusStart = 0, focusEnd = nodeRange<K>(1);
    int steps = 0;
    while (!tree.macRefine(box, focusEnd, focusEnd, focusStart, focusEnd, invTheta) && ++steps < 20) {}
    int numNodesVertex = 7 + 8, numNodesEdge = 6 + 2 * 8, numNodesFace = 4 + 4 * 8;
    EXPECT(int(tree.treeLeaves().size()) - 1 == 64 + 7 + 3 * numNodesFace + 3 * numNodesEdge + numNodesVertex);
}

// never called: instantiates every type combination of the facade templates so that the host-compiler check covers them
[[maybe_unused]] static void instantiateAll()
{
    DeviceVector<LocalIndex> temp, groups, layout;
    DeviceVector<float> xf, hf;
    DeviceVector<double> xd, hd;
    DeviceVector<uint32_t> l32;
    DeviceVector<uint64_t> l64;
    Box<float> bf(0, 1);
    Box<double> bd(0, 1);
    computeGroupSplits(0u, 0u, xf.data(), xf.data(), xf.data(), hf.data(), l32.data(), 0, layout.data(), bf, 32, 1.0f, temp, groups);
    computeGroupSplits(0u, 0u, xf.data(), xf.data(), xf.data(), hf.data(), l64.data(), 0, layout.data(), bf, 32, 1.0f, temp, groups);
    computeGroupSplits(0u, 0u, xd.data(), xd.data(), xd.data(), hf.data(), l32.data(), 0, layout.data(), bd, 64, 1.0f, temp, groups);
    computeGroupSplits(0u, 0u, xd.data(), xd.data(), xd.data(), hf.data(), l64.data(), 0, layout.data(), bd, 64, 1.0f, temp, groups);
    computeGroupSplits(0u, 0u, xd.data(), xd.data(), xd.data(), hd.data(), l32.data(), 0, layout.data(), bd, 64, 1.0f, temp, groups);
    FocusedOctreeSingleNode<unsigned> t32(8, 0.5f);
    FocusedOctreeSingleNode<uint64_t> t64(8, 0.5f);
    t32.update(bf, l32, 0u, 1u);
    t32.update(bd, l32, 0u, 1u);
    t64.update(bf, l64, uint64_t(0), uint64_t(1));
    t32.macRefine(bf, 0u, 1u, 0u, 2u, 1.0f);
    t64.macRefine(bd, uint64_t(0), uint64_t(1), uint64_t(0), uint64_t(2), 1.0f);
}

int main(int argc, char** argv)
{
    if (argc < 2 || std::strcmp(argv[1], "run") != 0)
    {
        std::printf("usage: %s run   (needs a CUDA device)\n", argv[0]);
        return 2;
    }
    fixedGroups();
    groupVolumes();
    computeEssentialTree<unsigned>();
    computeEssentialTree<uint64_t>();
    macRefinementFullSurface();
    if (failures == 0) std::printf("FOCUS_FACADE_OK\n");
    return failures ? 1 : 0;
}
