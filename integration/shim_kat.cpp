/*
 * shim_kat.cpp — TEST PROGRAM (not shipped): a client translation unit of the REFERENCE.  It includes the reference's
 * own headers only (the *_gpu.h declarations of its GPU library and its CPU algorithms), calls the pipeline through
 * the reference's GPU names — which cstone_b200_shim.cpp defines on top of libcstone_b200.so — and compares every
 * stage bit for bit with the reference's CPU functions on the same input.
 *
 *   keys (computeSfcKeysGpu, Hilbert + Morton)  ->  sort (sortByKeyGpu, both forms; sequenceGpu)  ->  gather (gatherGpu)
 *   ->  leaf array (computeNodeCountsGpu / computeNodeOpsGpu / rebalanceTreeGpu to convergence)
 *   ->  internal tree (buildOctreeGpu)  ->  halo radii (segmentMax + scaleGpu)  ->  halo flags (findHalosGpu)
 *   +   the small primitives (fill, increment, scans, count, lowerBound, reduce, MinMax, maxNormSquare, gatherRanges)
 *
 * Built by integration/Makefile where /root/reference is present; the binary travels to the GPU box.
 * Device memory: the C ABI's allocator (the reference's DeviceVector is part of the library this shim replaces).
 */
#define USE_CUDA
#include <algorithm>
#include <cstdio>
#include <numeric>
#include <random>
#include <vector>

#include "cstone/sfc/sfc_gpu.h"
#include "cstone/primitives/primitives_gpu.h"
#include "cstone/tree/csarray_gpu.h"
#include "cstone/tree/octree_gpu.h"
#include "cstone/traversal/collisions_gpu.h"
#include "cstone/halos/gather_halos_gpu.h"
// the reference's CPU side of the same path (the checker)
#include "cstone/sfc/sfc.hpp"
#include "cstone/tree/csarray.hpp"
#include "cstone/tree/octree.hpp"
#include "cstone/traversal/collisions.hpp"

#include "cstone_b200.h" // device malloc / memcpy only

using namespace cstone;

static int failures = 0;
#define EXPECT(c)                                                                                                      \
    do                                                                                                                 \
    {                                                                                                                  \
        if (!(c))                                                                                                      \
        {                                                                                                              \
            std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c);                                                    \
            ++failures;                                                                                                \
        }                                                                                                              \
    } while (0)

template<class T>
struct Dev
{
    T* p          = nullptr;
    std::size_t n = 0;
    explicit Dev(std::size_t n_) : n(n_)
    {
        void* q = nullptr;
        if (cstone_b200_device_malloc(&q, std::max<std::size_t>(n, 1) * sizeof(T))) std::abort();
        p = static_cast<T*>(q);
    }
    explicit Dev(const std::vector<T>& h) : Dev(h.size()) { up(h); }
    Dev(const Dev&)            = delete;
    Dev& operator=(const Dev&) = delete;
    ~Dev() { cstone_b200_device_free(p); }
    void up(const std::vector<T>& h) { cstone_b200_device_memcpy(p, h.data(), h.size() * sizeof(T), 0); }
    std::vector<T> down(std::size_t m) const
    {
        std::vector<T> h(m);
        cstone_b200_device_memcpy(h.data(), p, m * sizeof(T), 1);
        return h;
    }
    std::vector<T> down() const { return down(n); }
};

template<class KeyType, class T>
static void pipeline(std::size_t n, unsigned bucket)
{
    using HK = HilbertKey<KeyType>;
    Box<T> box(-1, 1, -2, 2, 0, 3, BoundaryType::open, BoundaryType::periodic, BoundaryType::open);
    std::mt19937 gen(42);
    std::normal_distribution<T> gx(0.0, 0.3), gz(1.5, 0.4);
    std::uniform_real_distribution<T> uy(-2, 2);
    std::vector<T> x(n), y(n), z(n), h(n);
    for (std::size_t i = 0; i < n; ++i)
    {
        x[i] = std::min(std::max(gx(gen), T(-1)), T(1));
        y[i] = uy(gen);
        z[i] = std::min(std::max(gz(gen), T(0)), T(3));
        h[i] = T(0.01) + T(0.02) * T(i % 7);
    }

    // ---- keys: both curves through the reference's strong key types
    Dev<T> dx(x), dy(y), dz(z), dh(h);
    Dev<KeyType> dkeys(n);
    std::vector<KeyType> keysCpu(n);
    computeSfcKeys(x.data(), y.data(), z.data(), reinterpret_cast<MortonKey<KeyType>*>(keysCpu.data()), n, box);
    computeSfcKeysGpu(dx.p, dy.p, dz.p, reinterpret_cast<MortonKey<KeyType>*>(dkeys.p), n, box);
    EXPECT(dkeys.down() == keysCpu);
    computeSfcKeys(x.data(), y.data(), z.data(), reinterpret_cast<HK*>(keysCpu.data()), n, box);
    computeSfcKeysGpu(dx.p, dy.p, dz.p, reinterpret_cast<HK*>(dkeys.p), n, box);
    EXPECT(dkeys.down() == keysCpu);

    // ---- sort with permutation: double-buffer form with caller scratch, then the allocating form on a copy
    std::vector<LocalIndex> orderCpu(n);
    std::iota(orderCpu.begin(), orderCpu.end(), 0u);
    std::stable_sort(orderCpu.begin(), orderCpu.end(), [&](LocalIndex a, LocalIndex b) { return keysCpu[a] < keysCpu[b]; });
    std::vector<KeyType> sortedCpu(n);
    for (std::size_t i = 0; i < n; ++i)
        sortedCpu[i] = keysCpu[orderCpu[i]];

    Dev<KeyType> dkeys2(keysCpu), dkeyBuf(n);
    Dev<LocalIndex> dorder(n), dvalBuf(n), dorder2(n);
    sequenceGpu(dorder.p, n, LocalIndex(0));
    sequenceGpu(dorder2.p, n, LocalIndex(0));
    {
        uint64_t tb = sortByKeyTempStorage<KeyType, LocalIndex>(n);
        Dev<char> temp(tb);
        sortByKeyGpu(dkeys.p, dkeys.p + n, dorder.p, dkeyBuf.p, dvalBuf.p, temp.p, tb);
        bool threw = false;
        try
        {
            if (tb > 16) sortByKeyGpu(dkeys2.p, dkeys2.p + n, dorder2.p, dkeyBuf.p, dvalBuf.p, temp.p, uint64_t(16));
            else threw = true;
        }
        catch (std::runtime_error&)
        {
            threw = true; // too small a temp buffer throws [ref: primitives/primitives_gpu.cu:353]
        }
        EXPECT(threw);
    }
    EXPECT(dkeys.down() == sortedCpu);
    EXPECT(dorder.down() == orderCpu); // equal keys keep their input order
    sortByKeyGpu(dkeys2.p, dkeys2.p + n, dorder2.p);
    EXPECT(dkeys2.down() == sortedCpu);
    EXPECT(dorder2.down() == orderCpu);

    // ---- gather x, y, z, h into SFC order
    Dev<T> sx(n), sy(n), sz(n), sh(n);
    gatherGpu(dorder.p, n, dx.p, sx.p);
    gatherGpu(dorder.p, n, dy.p, sy.p);
    gatherGpu(dorder.p, n, dz.p, sz.p);
    gatherGpu(dorder.p, n, dh.p, sh.p);
    std::vector<T> hs(n);
    for (std::size_t i = 0; i < n; ++i)
        hs[i] = h[orderCpu[i]];
    EXPECT(sh.down() == hs);

    // ---- leaf array: the update loop of tree/update_gpu.cuh:58-82 written with the three library calls
    auto [treeCpu, countsCpu] = computeOctree(sortedCpu.data(), sortedCpu.data() + n, bucket);
    std::size_t cap = treeCpu.size() * 2 + 64;
    Dev<KeyType> treeA(cap), treeB(cap);
    Dev<unsigned> counts(cap);
    Dev<TreeNodeIndex> nodeOps(cap);
    treeA.up(std::vector<KeyType>{0, nodeRange<KeyType>(0)});
    TreeNodeIndex numNodes = 1;
    computeNodeCountsGpu(treeA.p, counts.p, numNodes, dkeys.p, dkeys.p + n, std::numeric_limits<unsigned>::max(), false);
    bool converged = false;
    int steps      = 0;
    while (!converged && steps++ < 40)
    {
        TreeNodeIndex newNumNodes = computeNodeOpsGpu(treeA.p, numNodes, counts.p, bucket, nodeOps.p);
        EXPECT(std::size_t(newNumNodes) + 1 <= cap);
        converged = rebalanceTreeGpu(treeA.p, numNodes, newNumNodes, nodeOps.p, treeB.p);
        std::swap(treeA.p, treeB.p);
        numNodes = newNumNodes;
        computeNodeCountsGpu(treeA.p, counts.p, numNodes, dkeys.p, dkeys.p + n, std::numeric_limits<unsigned>::max(), true);
    }
    EXPECT(converged);
    EXPECT(std::size_t(numNodes) + 1 == treeCpu.size());
    EXPECT(treeA.down(numNodes + 1) == treeCpu);
    EXPECT(counts.down(numNodes) == countsCpu);
    EXPECT(reduceGpu(counts.p, std::size_t(numNodes), std::size_t(0)) == n);

    // ---- internal tree
    OctreeData<KeyType, CpuTag> oc;
    oc.resize(numNodes);
    updateInternalTree<KeyType>(treeCpu, oc.data());
    TreeNodeIndex M = oc.numNodes;
    Dev<KeyType> prefixes(M);
    Dev<TreeNodeIndex> childOffsets(M + 1), parents((M - 1) / 8 + 1), levelRange(maxTreeLevel<KeyType>{} + 2), i2l(M), l2i(M);
    OctreeView<KeyType> view{numNodes,       oc.numInternalNodes, M,     prefixes.p, childOffsets.p,
                             parents.p,      levelRange.p,        i2l.p, l2i.p};
    buildOctreeGpu(treeA.p, view);
    EXPECT(prefixes.down() == std::vector<KeyType>(oc.prefixes.begin(), oc.prefixes.end()));
    EXPECT(childOffsets.down(M) == std::vector<TreeNodeIndex>(oc.childOffsets.begin(), oc.childOffsets.begin() + M));
    EXPECT(parents.down((M - 1) / 8) == std::vector<TreeNodeIndex>(oc.parents.begin(), oc.parents.begin() + (M - 1) / 8));
    EXPECT(levelRange.down() == std::vector<TreeNodeIndex>(oc.levelRange.begin(), oc.levelRange.end()));
    EXPECT(i2l.down() == std::vector<TreeNodeIndex>(oc.internalToLeaf.begin(), oc.internalToLeaf.end()));
    EXPECT(l2i.down() == std::vector<TreeNodeIndex>(oc.leafToInternal.begin(), oc.leafToInternal.end()));

    // ---- layout (exclusive scan of the counts), halo radii, halo flags  [ref: halos/halos.hpp:148-169]
    Dev<LocalIndex> layout(numNodes + 1);
    exclusiveScanGpu(counts.p, counts.p + numNodes + 1, layout.p);
    std::vector<LocalIndex> layoutCpu(numNodes + 1, 0);
    std::partial_sum(countsCpu.begin(), countsCpu.end(), layoutCpu.begin() + 1);
    EXPECT(layout.down() == layoutCpu);

    Dev<float> radii(numNodes);
    segmentMax(sh.p, layout.p, std::size_t(numNodes), radii.p);
    scaleGpu(radii.p, radii.p + numNodes, 2.0f);
    std::vector<float> radiiCpu(numNodes);
    for (TreeNodeIndex i = 0; i < numNodes; ++i)
    {
        T m = 0;
        for (LocalIndex j = layoutCpu[i]; j < layoutCpu[i + 1]; ++j)
            m = std::max(m, hs[j]);
        radiiCpu[i] = float(m) * 2.0f;
    }
    EXPECT(radii.down() == radiiCpu);

    TreeNodeIndex first = numNodes / 4, last = numNodes / 2;
    Dev<int> flags(numNodes);
    fillGpu(flags.p, flags.p + numNodes, 0);
    findHalosGpu(prefixes.p, childOffsets.p, i2l.p, treeA.p, radii.p, box, first, last, flags.p);
    std::vector<int> flagsCpu(numNodes, 0);
    findHalos(oc.prefixes.data(), oc.childOffsets.data(), oc.internalToLeaf.data(), treeCpu.data(), radiiCpu.data(), box,
              first, last, flagsCpu.data());
    EXPECT(flags.down() == flagsCpu);
    EXPECT(countGpu(flags.p, flags.p + numNodes, 1) == std::size_t(std::count(flagsCpu.begin(), flagsCpu.end(), 1)));

    // ---- small primitives between the stages
    {
        auto [mn, mx] = MinMaxGpu<T>{}(dx.p, dx.p + n);
        EXPECT(mn == *std::min_element(x.begin(), x.end()) && mx == *std::max_element(x.begin(), x.end()));
        T nrm = 0;
        for (std::size_t i = 0; i < n; ++i)
            nrm = std::max(nrm, x[i] * x[i] + y[i] * y[i] + z[i] * z[i]);
        EXPECT(maxNormSquareGpu(dx.p, dy.p, dz.p, n) == nrm);

        KeyType probe = sortedCpu[n / 3];
        EXPECT(lowerBoundGpu(dkeys.p, dkeys.p + n, probe) ==
               std::size_t(std::lower_bound(sortedCpu.begin(), sortedCpu.end(), probe) - sortedCpu.begin()));
        std::vector<KeyType> probes{0, sortedCpu[n / 5], sortedCpu[n / 2] + 1, nodeRange<KeyType>(0)};
        Dev<KeyType> dprobes(probes);
        Dev<LocalIndex> dres(probes.size());
        lowerBoundGpu(dkeys.p, dkeys.p + n, dprobes.p, dprobes.p + probes.size(), dres.p);
        auto res = dres.down();
        for (std::size_t k = 0; k < probes.size(); ++k)
            EXPECT(res[k] == LocalIndex(std::lower_bound(sortedCpu.begin(), sortedCpu.end(), probes[k]) - sortedCpu.begin()));

        Dev<LocalIndex> inc(numNodes + 1), incl(numNodes + 1);
        incrementGpu(layout.p, layout.p + numNodes + 1, inc.p, LocalIndex(7));
        inclusiveScanGpu(counts.p, counts.p + numNodes, incl.p);
        auto hinc = inc.down(), hincl = incl.down(numNodes);
        bool ok = true;
        for (TreeNodeIndex i = 0; i <= numNodes; ++i)
            ok = ok && hinc[i] == layoutCpu[i] + 7;
        for (TreeNodeIndex i = 0; i < numNodes; ++i)
            ok = ok && hincl[i] == layoutCpu[i + 1];
        EXPECT(ok);

        // halo send-buffer packing [ref: halos/exchange_halos_gpu.cuh:70-73]
        std::vector<LocalIndex> offs{LocalIndex(n / 10), LocalIndex(n / 2), LocalIndex(n - 50)}, lens{30, 17, 50}, scan{0, 30, 47};
        Dev<LocalIndex> doffs(offs), dscan(scan);
        Dev<int> src(n), buf(97);
        std::vector<int> hsrc(n);
        std::iota(hsrc.begin(), hsrc.end(), 1000);
        src.up(hsrc);
        gatherRanges(dscan.p, doffs.p, 3, src.p, buf.p, 97);
        auto hbuf = buf.down();
        ok        = true;
        for (int r = 0, t = 0; r < 3; ++r)
            for (LocalIndex k = 0; k < lens[r]; ++k, ++t)
                ok = ok && hbuf[t] == hsrc[offs[r] + k];
        EXPECT(ok);
    }
    std::printf("%s / %s n=%zu bucket=%u: %d leaves, %d nodes, %d update steps, %d failure(s) so far\n",
                sizeof(KeyType) == 8 ? "u64" : "u32", sizeof(T) == 8 ? "f64" : "f32", n, bucket, int(numNodes), int(M), steps, failures);
}

int main()
{
    pipeline<uint64_t, double>(200000, 64);
    pipeline<uint64_t, float>(50000, 16);
    pipeline<unsigned, double>(50000, 64);
    pipeline<unsigned, float>(30000, 8);
    std::printf(failures ? "SHIM KAT FAILED (%d)\n" : "SHIM KAT OK (%d failures)\n", failures);
    return failures != 0;
}
