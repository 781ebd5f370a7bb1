/*
 * focus_tree.cuh -- the device-resident focus tree (class FocusTree) shared by focus_tree.cu (single-rank C entry
 * points) and let_tree.cuh (the multi-rank locally essential tree of the Domain engine).  See focus_tree.cu for the
 * description of the update steps and the reference lines they follow.
 */
#pragma once
#include <cstdlib>
#include <cstring>
#include <initializer_list>
#include <vector>

#include "common.cuh"

namespace csb
{


//! grow-only device array (5 % head room, as the reference's reallocate policy, util/reallocate.hpp:37-66)
struct FBuf
{
    void* p      = nullptr;
    size_t bytes = 0;
    int reserve(size_t need)
    {
        if (need <= bytes) return 0;
        size_t cap = size_t(double(need) * 1.05) + 256;
        if (p) cudaFree(p);
        p = nullptr, bytes = 0;
        CSB_CUDA(cudaMalloc(&p, cap));
        bytes = cap;
        return 0;
    }
    template<class U>
    U* as() const
    {
        return reinterpret_cast<U*>(p);
    }
    void swap(FBuf& o)
    {
        std::swap(p, o.p);
        std::swap(bytes, o.bytes);
    }
    ~FBuf()
    {
        if (p) cudaFree(p);
    }
    FBuf()                       = default;
    FBuf(const FBuf&)            = delete;
    FBuf& operator=(const FBuf&) = delete;
};

/*! host array in page-locked memory (grow-only, 30 % head room): the host copies of leaf keys, counts, halo flags and
 *  layout that the multi-rank focus tree reads and writes every sync cross PCIe at full speed and asynchronously.
 *  Falls back to pageable memory if the pinned allocation fails.  Only what the callers need of std::vector. */
template<class U>
class PinnedVec
{
public:
    PinnedVec() = default;
    PinnedVec(std::initializer_list<U> init) { assign(init.begin(), init.end()); }
    PinnedVec(const PinnedVec&)            = delete;
    PinnedVec& operator=(const PinnedVec&) = delete;
    ~PinnedVec() { release(p_, pinned_); }

    void resize(size_t n)
    {
        if (n > cap_)
        {
            const size_t cap = size_t(double(n) * 1.3) + 1024;
            U* q             = nullptr;
            bool pinned      = cudaMallocHost(reinterpret_cast<void**>(&q), cap * sizeof(U)) == cudaSuccess;
            if (!pinned)
            {
                cudaGetLastError();
                q = static_cast<U*>(std::malloc(cap * sizeof(U)));
            }
            if (n_) std::memcpy(q, p_, n_ * sizeof(U));
            release(p_, pinned_);
            p_ = q, cap_ = cap, pinned_ = pinned;
        }
        n_ = n;
    }
    void assign(const U* first, const U* last)
    {
        n_ = 0;
        resize(size_t(last - first));
        if (n_) std::memcpy(p_, first, n_ * sizeof(U));
    }
    U* data() { return p_; }
    const U* data() const { return p_; }
    size_t size() const { return n_; }
    bool empty() const { return n_ == 0; }
    U& operator[](size_t i) { return p_[i]; }
    const U& operator[](size_t i) const { return p_[i]; }
    const U& back() const { return p_[n_ - 1]; }
    const U* begin() const { return p_; }
    const U* end() const { return p_ + n_; }

private:
    static void release(U* p, bool pinned)
    {
        if (!p) return;
        if (pinned) cudaFreeHost(p);
        else std::free(p);
    }
    U* p_        = nullptr;
    size_t n_    = 0, cap_ = 0;
    bool pinned_ = false;
};

template<class K>
struct FocusApi;
#define CSB_FOCUS_TREE_API(KS, K)                                                                                      \
    template<>                                                                                                         \
    struct FocusApi<K>                                                                                                 \
    {                                                                                                                  \
        static constexpr auto decision = cstone_b200_rebalance_decision_essential_##KS;                                \
        static constexpr auto enforce  = cstone_b200_enforce_keys_##KS;                                                \
        static constexpr auto protect  = cstone_b200_protect_ancestors_##KS;                                           \
        static constexpr auto refine   = cstone_b200_mac_refine_decision_##KS;                                         \
        static constexpr auto rebal    = cstone_b200_rebalance_tree_##KS;                                              \
        static constexpr auto sortKeys = cstone_b200_sort_keys_##KS;                                                   \
        static constexpr auto sortTemp = cstone_b200_sort_keys_temp_bytes_##KS;                                        \
        static constexpr auto gapCount = cstone_b200_count_sfc_gaps_##KS;                                              \
        static constexpr auto gapFill  = cstone_b200_fill_sfc_gaps_##KS;                                               \
        static constexpr auto build    = cstone_b200_build_octree_##KS;                                                \
        static constexpr auto buildTmp = cstone_b200_build_octree_temp_bytes_##KS;                                     \
        static constexpr auto counts   = cstone_b200_compute_node_counts_##KS;                                         \
        static int geoMac(int kind, const K* p, int n, float* c, float it, const cstone_box_f32* b, void* s)           \
        {                                                                                                              \
            return cstone_b200_geo_mac_spheres_##KS##_f32(kind, p, n, c, it, b, s);                                    \
        }                                                                                                              \
        static int geoMac(int kind, const K* p, int n, double* c, float it, const cstone_box_f64* b, void* s)          \
        {                                                                                                              \
            return cstone_b200_geo_mac_spheres_##KS##_f64(kind, p, n, c, it, b, s);                                    \
        }                                                                                                              \
        static int mark(int kind, const K* p, const int* co, const float* c, const cstone_box_f32* b, const K* f,      \
                        int nf, int lim, char* m, void* s)                                                             \
        {                                                                                                              \
            return cstone_b200_mark_macs_##KS##_f32(kind, p, co, c, b, f, nf, lim, m, s);                              \
        }                                                                                                              \
        static int mark(int kind, const K* p, const int* co, const double* c, const cstone_box_f64* b, const K* f,     \
                        int nf, int lim, char* m, void* s)                                                             \
        {                                                                                                              \
            return cstone_b200_mark_macs_##KS##_f64(kind, p, co, c, b, f, nf, lim, m, s);                              \
        }                                                                                                              \
    };
CSB_FOCUS_TREE_API(k32, uint32_t)
CSB_FOCUS_TREE_API(k64, uint64_t)

inline int lowerBoundKey(const uint32_t* first, size_t n, uint32_t v, uint64_t* out, void* temp, void* s)
{
    return cstone_b200_lower_bound_u32(first, n, v, out, temp, s);
}
inline int lowerBoundKey(const uint64_t* first, size_t n, uint64_t v, uint64_t* out, void* temp, void* s)
{
    return cstone_b200_lower_bound_u64(first, n, v, out, temp, s);
}

//! ops[idx[i]] = 0
template<class I>
__global__ void zeroAtKernel(const I* __restrict__ idx, int n, int* __restrict__ ops)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) ops[idx[i]] = 0;
}

template<class K>
class FocusTree
{
public:
    FocusTree(int kind, unsigned bucket, float theta)
        : kind_(kind)
        , bucket_(bucket)
        , theta_(theta)
    {
    }

    //! leaves {0, 2^(3 maxLevel)}, counts {bucket + 1}, macs {1}  [ref: focus/octree_focus.hpp:354-363]
    int init(cudaStream_t s)
    {
        std::vector<K> root{K(0), nodeRange<K>(0)};
        CSB_TRY(leaves_.reserve(2 * sizeof(K)));
        CSB_CUDA(cudaMemcpyAsync(leaves_.p, root.data(), 2 * sizeof(K), cudaMemcpyHostToDevice, s));
        numLeaves_ = 1;
        CSB_TRY(link(s));
        unsigned c = bucket_ + 1;
        char m     = 1;
        CSB_TRY(counts_.reserve(4));
        CSB_TRY(macs_.reserve(1));
        CSB_CUDA(cudaMemcpyAsync(counts_.p, &c, 4, cudaMemcpyHostToDevice, s));
        CSB_CUDA(cudaMemcpyAsync(macs_.p, &m, 1, cudaMemcpyHostToDevice, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        ready_ = true;
        return 0;
    }

    template<class T>
    int update(const typename BoxT<T>::type* box, const K* keys, size_t numKeys, K focusStart, K focusEnd,
               const K* mandatoryHost, int numMandatory, int* convergedHost, cudaStream_t s);

    //! replace the leaf array (e.g. a tree received from elsewhere) and relink; counts and macs are zeroed
    int setLeaves(const K* leavesDev, int numLeaves, cudaStream_t s)
    {
        CSB_TRY(leaves_.reserve((size_t(numLeaves) + 1) * sizeof(K)));
        CSB_CUDA(cudaMemcpyAsync(leaves_.p, leavesDev, (size_t(numLeaves) + 1) * sizeof(K), cudaMemcpyDeviceToDevice, s));
        numLeaves_ = numLeaves;
        CSB_TRY(link(s));
        CSB_TRY(counts_.reserve(size_t(numNodes_) * 4));
        CSB_TRY(macs_.reserve(size_t(numNodes_)));
        CSB_CUDA(cudaMemsetAsync(counts_.p, 0, size_t(numNodes_) * 4, s));
        CSB_CUDA(cudaMemsetAsync(macs_.p, 0, size_t(numNodes_), s));
        CSB_TRY(leafCounts_.reserve(size_t(numLeaves_) * 4));
        CSB_CUDA(cudaMemsetAsync(leafCounts_.p, 0, size_t(numLeaves_) * 4, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        ready_ = true;
        return 0;
    }

    template<class T>
    int macRefine(const typename BoxT<T>::type* box, K oldFocusStart, K oldFocusEnd, K focusStart, K focusEnd,
                  float invTheta, int* convergedHost, cudaStream_t s);

    // ---- the pieces of update(), also driven one by one by the multi-rank tree (let_tree.cuh)
    //! steps 1-6: node ops from counts_ / macs_, mandatory keys, rebalance (+ injection), relink
    int rebalanceStep(K focusStart, K focusEnd, const K* mandatoryHost, int numMandatory, bool* converged, cudaStream_t s);
    //! step 7: MAC spheres l * invThetaEff around the geometric centres, marks w.r.t. the leaves in [focusStart, focusEnd)
    template<class T>
    int minMac(const typename BoxT<T>::type* box, float invThetaEff, K focusStart, K focusEnd, cudaStream_t s);
    //! step 8a: leafCounts_ = particles per leaf
    int localCounts(const K* keys, size_t numKeys, cudaStream_t s);
    //! step 8b: counts_ = leafCounts_ scattered to the leaves of the linked tree, internal nodes zeroed
    int scatterCounts(cudaStream_t s);
    //! step 8c: internal node counts = sums of their children
    int upsweepCounts(cudaStream_t s);
    /*! remove the leaves with the given indices (device array, n > 0): their key ranges join the leaf before them;
     *  rebalance with ops 0 / 1 and relink  [ref: focus/exchange_focus.hpp:188-217 syncTreeletsGpu] */
    int dropLeaves(const uint32_t* idxDev, int n, cudaStream_t s)
    {
        using A     = FocusApi<K>;
        const int L = numLeaves_;
        CSB_TRY(leafOps_.reserve((size_t(L) + 1) * 4));
        CSB_TRY(nodeOps_.reserve((size_t(numNodes_) + 1) * 4));
        CSB_TRY(temp_.reserve(size_t(cstone_b200_scan_temp_bytes(uint64_t(L) + 1)) + 4096));
        const int one = 1;
        CSB_TRY(cstone_b200_fill(leafOps_.p, size_t(L), &one, 4, s));
        CSB_CUDA(cudaMemsetAsync(leafOps_.as<int>() + L, 0, 4, s));
        zeroAtKernel<uint32_t><<<unsigned((n + 255) / 256), 256, 0, s>>>(idxDev, n, leafOps_.as<int>());
        CSB_LAUNCHED();
        CSB_TRY(cstone_b200_exclusive_scan_u32(leafOps_.as<uint32_t>(), nodeOps_.as<uint32_t>(), size_t(L) + 1, 0u, temp_.p, s));
        int newL = 0;
        CSB_CUDA(cudaMemcpyAsync(&newL, nodeOps_.as<int>() + L, 4, cudaMemcpyDeviceToHost, s));
        CSB_CUDA(cudaStreamSynchronize(s));
        CSB_TRY(newLeaves_.reserve((size_t(newL) + 1) * sizeof(K)));
        CSB_TRY(A::rebal(leaves_.as<K>(), L, newL, nodeOps_.as<int>(), newLeaves_.as<K>(), s));
        leaves_.swap(newLeaves_);
        numLeaves_ = newL;
        return link(s);
    }
    //! replace the leaves by a HOST array and relink (counts / macs keep their buffers, contents are stale)
    int setLeavesHost(const K* leavesHost, int numLeaves, cudaStream_t s)
    {
        CSB_TRY(leaves_.reserve((size_t(numLeaves) + 1) * sizeof(K)));
        CSB_CUDA(cudaMemcpyAsync(leaves_.p, leavesHost, (size_t(numLeaves) + 1) * sizeof(K), cudaMemcpyHostToDevice, s));
        numLeaves_ = numLeaves;
        return link(s);
    }

    int kind_;
    unsigned bucket_;
    float theta_;
    bool ready_ = false;

    int numLeaves_ = 0, numInternal_ = 0, numNodes_ = 0;
    int levelRange_[KeyTraits<K>::maxLevel + 2];
    FBuf leaves_, newLeaves_, prefixes_, childOffsets_, parents_, levelRangeDev_, i2l_, l2i_;
    FBuf counts_, leafCounts_, macs_, centers_, nodeOps_, leafOps_, forced_, temp_;

private:
    int link(cudaStream_t s);
    int inject(const K* forcedDev, int numForced, cudaStream_t s);
};

//! rebuild the linked tree from leaves_  [ref: tree/octree.hpp:377-383 updateInternalTree, OctreeData::resize :327-342]
template<class K>
int FocusTree<K>::link(cudaStream_t s)
{
    using A      = FocusApi<K>;
    numInternal_ = (numLeaves_ - 1) / 7;
    numNodes_    = numLeaves_ + numInternal_;
    size_t M     = size_t(numNodes_);
    CSB_TRY(prefixes_.reserve(M * sizeof(K)));
    CSB_TRY(childOffsets_.reserve((M + 1) * 4));
    CSB_TRY(parents_.reserve(std::max<size_t>(1, (M - 1) / 8) * 4 + 4));
    CSB_TRY(levelRangeDev_.reserve(sizeof(levelRange_)));
    CSB_TRY(i2l_.reserve(M * 4));
    CSB_TRY(l2i_.reserve(M * 4));
    uint64_t tb = A::buildTmp(numLeaves_);
    CSB_TRY(temp_.reserve(size_t(tb) + 4096));
    CSB_TRY(A::build(leaves_.as<K>(), numLeaves_, prefixes_.as<K>(), childOffsets_.as<int>(), parents_.as<int>(),
                     levelRangeDev_.as<int>(), i2l_.as<int>(), l2i_.as<int>(), temp_.p, temp_.bytes, s));
    CSB_CUDA(cudaMemcpyAsync(levelRange_, levelRangeDev_.p, sizeof(levelRange_), cudaMemcpyDeviceToHost, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    return 0;
}

/*! keys that the rebalance step could not introduce: append them to the leaf array, sort, and fill every gap between
 *  consecutive keys with the cornerstone nodes spanning it  [ref: focus/inject.hpp:99-128 injectKeysGpu; the CPU
 *  twin :52-74 unites the leaves with computeSpanningTree(keys) — the same key set] */
template<class K>
int FocusTree<K>::inject(const K* forcedDev, int numForced, cudaStream_t s)
{
    using A     = FocusApi<K>;
    size_t nOld = size_t(numLeaves_) + 1, nAll = nOld + size_t(numForced);
    // leaves_ must hold nAll keys: grow through newLeaves_ (contents kept)
    CSB_TRY(newLeaves_.reserve(nAll * sizeof(K)));
    CSB_CUDA(cudaMemcpyAsync(newLeaves_.p, leaves_.p, nOld * sizeof(K), cudaMemcpyDeviceToDevice, s));
    CSB_CUDA(cudaMemcpyAsync(newLeaves_.as<K>() + nOld, forcedDev, size_t(numForced) * sizeof(K),
                             cudaMemcpyDeviceToDevice, s));
    leaves_.swap(newLeaves_);
    CSB_TRY(newLeaves_.reserve(nAll * sizeof(K)));
    uint64_t tb = A::sortTemp(nAll);
    CSB_TRY(temp_.reserve(size_t(tb) + 4096));
    CSB_TRY(A::sortKeys(leaves_.as<K>(), newLeaves_.as<K>(), nAll, temp_.p, temp_.bytes, s));

    // gaps: nAll - 1 intervals, nAll scan entries
    CSB_TRY(leafOps_.reserve(nAll * 4));
    CSB_TRY(nodeOps_.reserve(nAll * 4));
    CSB_CUDA(cudaMemsetAsync(leafOps_.p, 0, nAll * 4, s));
    CSB_TRY(A::gapCount(leaves_.as<K>(), int(nAll - 1), leafOps_.as<int>(), s));
    CSB_TRY(temp_.reserve(size_t(cstone_b200_scan_temp_bytes(nAll)) + 4096));
    CSB_TRY(cstone_b200_exclusive_scan_u32(leafOps_.as<uint32_t>(), nodeOps_.as<uint32_t>(), nAll, 0u, temp_.p, s));
    int numNodesGap = 0;
    CSB_CUDA(cudaMemcpyAsync(&numNodesGap, nodeOps_.as<int>() + nAll - 1, 4, cudaMemcpyDeviceToHost, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    CSB_TRY(newLeaves_.reserve((size_t(numNodesGap) + 1) * sizeof(K)));
    CSB_TRY(A::gapFill(leaves_.as<K>(), int(nAll - 1), nodeOps_.as<int>(), newLeaves_.as<K>(), s));
    leaves_.swap(newLeaves_);
    numLeaves_ = numNodesGap;
    return 0;
}

template<class K>
int FocusTree<K>::rebalanceStep(K focusStart, K focusEnd, const K* mandatoryHost, int numMandatory, bool* convergedOut,
                                cudaStream_t s)
{
    using A = FocusApi<K>;
    if (!ready_) CSB_TRY(init(s));
    const int L = numLeaves_, M = numNodes_;

    // ---- 1-3: op per node  [ref: focus/octree_focus.hpp:98-110]
    CSB_TRY(nodeOps_.reserve((size_t(M) + 1) * 4));
    CSB_TRY(temp_.reserve(size_t(cstone_b200_scan_temp_bytes(uint64_t(M) + 1)) + cstone_b200_reduce_temp_bytes() + 4096));
    CSB_TRY(A::decision(prefixes_.as<K>(), childOffsets_.as<int>(), parents_.as<int>(), counts_.as<uint32_t>(),
                        macs_.as<char>(), focusStart, focusEnd, bucket_, nodeOps_.as<int>(), M, s));
    std::vector<K> forced{focusStart, focusEnd};
    forced.insert(forced.end(), mandatoryHost, mandatoryHost + numMandatory);
    CSB_TRY(forced_.reserve(forced.size() * sizeof(K)));
    CSB_CUDA(cudaMemcpyAsync(forced_.p, forced.data(), forced.size() * sizeof(K), cudaMemcpyHostToDevice, s));
    int status = 0, protectedOk = 0;
    CSB_TRY(A::enforce(forced_.as<K>(), int(forced.size()), prefixes_.as<K>(), childOffsets_.as<int>(),
                       parents_.as<int>(), nodeOps_.as<int>(), &status, temp_.p, s));
    CSB_TRY(A::protect(prefixes_.as<K>(), parents_.as<int>(), nodeOps_.as<int>(), M, &protectedOk, temp_.p, s));
    bool converged = protectedOk != 0;

    // ---- 4: leaf ops, scan, new leaf array  [ref: :112-125]
    CSB_TRY(leafOps_.reserve((size_t(L) + 1) * 4));
    CSB_CUDA(cudaMemsetAsync(leafOps_.as<int>() + L, 0, 4, s));
    CSB_TRY(cstone_b200_gather(l2i_.as<uint32_t>() + numInternal_, size_t(L), nodeOps_.p, leafOps_.p, 4, s));
    if (status == 1) // cancelMerge: converged iff every leaf op is 1
    {
        uint64_t ones = 0;
        CSB_TRY(cstone_b200_count_u32(leafOps_.as<uint32_t>(), size_t(L), 1u, &ones, temp_.p, s));
        converged = ones == uint64_t(L);
    }
    else if (status == 2) { converged = false; }
    // scanned ops go to nodeOps_ (M + 1 >= L + 1 entries)
    CSB_TRY(cstone_b200_exclusive_scan_u32(leafOps_.as<uint32_t>(), nodeOps_.as<uint32_t>(), size_t(L) + 1, 0u,
                                           temp_.p, s));
    int newL = 0;
    CSB_CUDA(cudaMemcpyAsync(&newL, nodeOps_.as<int>() + L, 4, cudaMemcpyDeviceToHost, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    CSB_TRY(newLeaves_.reserve((size_t(newL) + 1) * sizeof(K)));
    CSB_TRY(A::rebal(leaves_.as<K>(), L, newL, nodeOps_.as<int>(), newLeaves_.as<K>(), s));
    leaves_.swap(newLeaves_);
    numLeaves_ = newL;

    // ---- 5: forced injection  [ref: :127-132]
    if (status == 3)
    {
        converged = false;
        CSB_TRY(inject(forced_.as<K>(), int(forced.size()), s));
    }
    // ---- 6
    CSB_TRY(link(s));
    *convergedOut = converged;
    return 0;
}

template<class K>
template<class T>
int FocusTree<K>::minMac(const typename BoxT<T>::type* box, float invThetaEff, K focusStart, K focusEnd, cudaStream_t s)
{
    using A = FocusApi<K>;
    // ---- 7: MAC marks of the tree w.r.t. the focus  [ref: focus/octree_focus.hpp:375-391, octree_focus_mpi.hpp:457-525]
    CSB_TRY(temp_.reserve(size_t(cstone_b200_reduce_temp_bytes()) + 4096));
    CSB_TRY(centers_.reserve(size_t(numNodes_) * 4 * sizeof(T)));
    CSB_TRY(A::geoMac(kind_, prefixes_.as<K>(), numNodes_, centers_.as<T>(), invThetaEff, box, s));
    CSB_TRY(macs_.reserve(size_t(numNodes_)));
    CSB_CUDA(cudaMemsetAsync(macs_.p, 0, size_t(numNodes_), s));
    uint64_t fStart = 0, fEnd = 0;
    CSB_TRY(lowerBoundKey(leaves_.as<K>(), size_t(numLeaves_), focusStart, &fStart, temp_.p, s));
    CSB_TRY(lowerBoundKey(leaves_.as<K>(), size_t(numLeaves_), focusEnd, &fEnd, temp_.p, s));
    if (fEnd > fStart)
        CSB_TRY(A::mark(kind_, prefixes_.as<K>(), childOffsets_.as<int>(), centers_.as<T>(), box,
                        leaves_.as<K>() + fStart, int(fEnd - fStart), 0, macs_.as<char>(), s));
    return 0;
}

template<class K>
int FocusTree<K>::localCounts(const K* keys, size_t numKeys, cudaStream_t s)
{
    using A = FocusApi<K>;
    CSB_TRY(leafCounts_.reserve(size_t(numLeaves_) * 4));
    return A::counts(leaves_.as<K>(), leafCounts_.as<uint32_t>(), numLeaves_, keys, keys + numKeys, 0xFFFFFFFFu, 1, s);
}

template<class K>
int FocusTree<K>::scatterCounts(cudaStream_t s)
{
    CSB_TRY(counts_.reserve(size_t(numNodes_) * 4));
    CSB_CUDA(cudaMemsetAsync(counts_.p, 0, size_t(numNodes_) * 4, s));
    return cstone_b200_scatter(l2i_.as<uint32_t>() + numInternal_, size_t(numLeaves_), leafCounts_.p, counts_.p, 4, s);
}

template<class K>
int FocusTree<K>::upsweepCounts(cudaStream_t s)
{
    return cstone_b200_upsweep_sum(int(KeyTraits<K>::maxLevel) + 1, levelRange_, childOffsets_.as<int>(),
                                   counts_.as<uint32_t>(), s);
}

template<class K>
template<class T>
int FocusTree<K>::update(const typename BoxT<T>::type* box, const K* keys, size_t numKeys, K focusStart, K focusEnd,
                         const K* mandatoryHost, int numMandatory, int* convergedHost, cudaStream_t s)
{
    bool converged = false;
    CSB_TRY(rebalanceStep(focusStart, focusEnd, mandatoryHost, numMandatory, &converged, s));
    CSB_TRY(minMac<T>(box, float(double(1.0f / theta_) + 0.5), focusStart, focusEnd, s));
    CSB_TRY(localCounts(keys, numKeys, s));
    CSB_TRY(scatterCounts(s));
    CSB_TRY(upsweepCounts(s));
    CSB_CUDA(cudaStreamSynchronize(s));
    if (convergedHost) *convergedHost = converged ? 1 : 0;
    return 0;
}

/*! one step of macRefineGpu / updateMacRefineGpu [ref: focus/octree_focus.hpp:297-342]: after the focus moved from
 *  [oldFocusStart, oldFocusEnd) to [focusStart, focusEnd), leaves OUTSIDE the focus that fail the minimum-distance MAC
 *  w.r.t. the newly covered part of the focus are split by one level.  converged = no leaf was split. */
template<class K>
template<class T>
int FocusTree<K>::macRefine(const typename BoxT<T>::type* box, K oldFocusStart, K oldFocusEnd, K focusStart,
                            K focusEnd, float invTheta, int* convergedHost, cudaStream_t s)
{
    using A = FocusApi<K>;
    if (!ready_) CSB_TRY(init(s));
    if (oldFocusStart == focusStart && oldFocusEnd == focusEnd)
    {
        if (convergedHost) *convergedHost = 1;
        return 0;
    }
    const int L = numLeaves_, M = numNodes_;
    CSB_TRY(temp_.reserve(size_t(cstone_b200_scan_temp_bytes(uint64_t(M) + 1)) + cstone_b200_reduce_temp_bytes() + 4096));
    CSB_TRY(centers_.reserve(size_t(M) * 4 * sizeof(T)));
    CSB_TRY(A::geoMac(kind_, prefixes_.as<K>(), M, centers_.as<T>(), invTheta, box, s));
    CSB_TRY(macs_.reserve(size_t(M)));
    CSB_CUDA(cudaMemsetAsync(macs_.p, 0, size_t(M), s));

    K growthLower = focusStart < oldFocusStart ? oldFocusStart : focusStart;
    K growthUpper = oldFocusEnd < focusEnd ? oldFocusEnd : focusEnd;
    uint64_t fGrowL = 0, fGrowU = 0, fStart = 0, fEnd = 0;
    CSB_TRY(lowerBoundKey(leaves_.as<K>(), size_t(L), growthLower, &fGrowL, temp_.p, s));
    CSB_TRY(lowerBoundKey(leaves_.as<K>(), size_t(L), growthUpper, &fGrowU, temp_.p, s));
    CSB_TRY(lowerBoundKey(leaves_.as<K>(), size_t(L), focusStart, &fStart, temp_.p, s));
    CSB_TRY(lowerBoundKey(leaves_.as<K>(), size_t(L), focusEnd, &fEnd, temp_.p, s));
    if (fGrowL > fStart)
        CSB_TRY(A::mark(kind_, prefixes_.as<K>(), childOffsets_.as<int>(), centers_.as<T>(), box,
                        leaves_.as<K>() + fStart, int(fGrowL - fStart), 1, macs_.as<char>(), s));
    if (fEnd > fGrowU)
        CSB_TRY(A::mark(kind_, prefixes_.as<K>(), childOffsets_.as<int>(), centers_.as<T>(), box,
                        leaves_.as<K>() + fGrowU, int(fEnd - fGrowU), 1, macs_.as<char>(), s));

    CSB_TRY(leafOps_.reserve((size_t(L) + 1) * 4));
    CSB_TRY(nodeOps_.reserve((size_t(M) + 1) * 4));
    CSB_CUDA(cudaMemsetAsync(leafOps_.as<int>() + L, 0, 4, s));
    CSB_TRY(A::refine(prefixes_.as<K>(), macs_.as<char>(), l2i_.as<int>() + numInternal_, L, int(fStart), int(fEnd),
                      leafOps_.as<int>(), s));
    uint64_t ones = 0;
    CSB_TRY(cstone_b200_count_u32(leafOps_.as<uint32_t>(), size_t(L), 1u, &ones, temp_.p, s));
    const bool converged = ones == uint64_t(L);
    CSB_TRY(cstone_b200_exclusive_scan_u32(leafOps_.as<uint32_t>(), nodeOps_.as<uint32_t>(), size_t(L) + 1, 0u,
                                           temp_.p, s));
    int newL = 0;
    CSB_CUDA(cudaMemcpyAsync(&newL, nodeOps_.as<int>() + L, 4, cudaMemcpyDeviceToHost, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    CSB_TRY(newLeaves_.reserve((size_t(newL) + 1) * sizeof(K)));
    CSB_TRY(A::rebal(leaves_.as<K>(), L, newL, nodeOps_.as<int>(), newLeaves_.as<K>(), s));
    leaves_.swap(newLeaves_);
    numLeaves_ = newL;
    CSB_TRY(link(s));
    // node arrays of the new tree: sized, zeroed (the caller refreshes them with update())
    CSB_TRY(counts_.reserve(size_t(numNodes_) * 4));
    CSB_TRY(macs_.reserve(size_t(numNodes_)));
    CSB_TRY(leafCounts_.reserve(size_t(numLeaves_) * 4));
    CSB_CUDA(cudaMemsetAsync(counts_.p, 0, size_t(numNodes_) * 4, s));
    CSB_CUDA(cudaMemsetAsync(macs_.p, 0, size_t(numNodes_), s));
    CSB_CUDA(cudaMemsetAsync(leafCounts_.p, 0, size_t(numLeaves_) * 4, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    if (convergedHost) *convergedHost = converged ? 1 : 0;
    return 0;
}

} // namespace csb

