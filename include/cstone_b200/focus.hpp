/*
 * cstone_b200/focus.hpp — target groups and the single-rank focus tree over libcstone_b200.so, with the reference's
 * names and argument order:
 *   GroupView, GroupData<GpuTag>                      [ref: traversal/groups.hpp:19-55]
 *   computeFixedGroups, computeGroupSplits            [ref: traversal/groups_gpu.h:60-88]
 *   FocusedOctreeSingleNode<KeyType>                  [ref: focus/octree_focus.hpp:349-427]
 *   macRefine step on the same tree                   [ref: focus/octree_focus.hpp:318-342 macRefineGpu]
 * Particle keys, coordinates and trees are device arrays (cstone::DeviceVector or raw device pointers).
 */
#pragma once

#include <cstdint>
#include <stdexcept>
#include <type_traits>
#include <vector>

#include "../cstone_b200.h"
#include "box.hpp"
#include "device_vector.hpp"

namespace cstone
{

//! @brief view of groups of consecutive target particles, group i = [groupStart[i], groupEnd[i])
struct GroupView
{
    LocalIndex firstBody, lastBody;
    LocalIndex numGroups;
    const LocalIndex* groupStart;
    const LocalIndex* groupEnd;
};

template<class Accelerator>
class GroupData;

template<>
class GroupData<GpuTag>
{
public:
    GroupData()                 = default;
    GroupData(const GroupView&) = delete;
    GroupView view() const { return {firstBody, lastBody, numGroups, groupStart, groupEnd}; }

    DeviceVector<LocalIndex> data;
    LocalIndex firstBody{0}, lastBody{0};
    LocalIndex numGroups{0};
    LocalIndex* groupStart{nullptr};
    LocalIndex* groupEnd{nullptr};

    friend void swap(GroupData& lhs, GroupData& rhs)
    {
        lhs.data.swap(rhs.data);
        std::swap(lhs.firstBody, rhs.firstBody);
        std::swap(lhs.lastBody, rhs.lastBody);
        std::swap(lhs.numGroups, rhs.numGroups);
        std::swap(lhs.groupStart, rhs.groupStart);
        std::swap(lhs.groupEnd, rhs.groupEnd);
    }
};

//! @brief groups of @p groupSize consecutive particles
inline void computeFixedGroups(LocalIndex first, LocalIndex last, unsigned groupSize, GroupData<GpuTag>& groups)
{
    LocalIndex numGroups = groupSize ? (last - first + groupSize - 1) / groupSize : 0;
    groups.data.resize(numGroups + 1);
    detail::check(cstone_b200_compute_fixed_groups(first, last, groupSize, groups.data.data(), nullptr),
                  "computeFixedGroups");
    detail::check(cstone_b200_device_synchronize(), "computeFixedGroups");
    groups.firstBody  = first;
    groups.lastBody   = last;
    groups.numGroups  = numGroups;
    groups.groupStart = groups.data.data();
    groups.groupEnd   = groups.data.data() + 1;
}

namespace detail
{
template<class Tc, class T, class K>
struct GroupSplitsApi;
#define CSB_GROUP_SPLITS_API(KS, K, TS, TC, TH)                                                                        \
    template<>                                                                                                         \
    struct GroupSplitsApi<TC, TH, K>                                                                                   \
    {                                                                                                                  \
        static constexpr auto run = cstone_b200_compute_group_splits_##KS##_##TS;                                      \
    };
CSB_GROUP_SPLITS_API(k32, uint32_t, f64_f64, double, double)
CSB_GROUP_SPLITS_API(k32, uint32_t, f64_f32, double, float)
CSB_GROUP_SPLITS_API(k32, uint32_t, f32_f32, float, float)
CSB_GROUP_SPLITS_API(k64, uint64_t, f64_f64, double, double)
CSB_GROUP_SPLITS_API(k64, uint64_t, f64_f32, double, float)
CSB_GROUP_SPLITS_API(k64, uint64_t, f32_f32, float, float)
#undef CSB_GROUP_SPLITS_API
} // namespace detail

/*! @brief groups of at most @p groupSize (32 or 64) consecutive particles, cut where neighbours along the SFC are
 *         further apart than tolFactor x the leaf edge
 *
 * @param[-]   numSplitsPerGroup   scratch (the reference's temporary); resized, content unspecified
 * @param[out] groups              group start indices, numGroups + 1 entries, the last one is @p last
 */
template<class Tc, class T, class KeyType>
void computeGroupSplits(LocalIndex first,
                        LocalIndex last,
                        const Tc* x,
                        const Tc* y,
                        const Tc* z,
                        const T* h,
                        const KeyType* leaves,
                        TreeNodeIndex numLeaves,
                        const LocalIndex* layout,
                        const Box<Tc> box,
                        unsigned groupSize,
                        float tolFactor,
                        DeviceVector<LocalIndex>& numSplitsPerGroup,
                        DeviceVector<LocalIndex>& groups)
{
    LocalIndex n = last > first ? last - first : 0;
    if (groupSize != 32 && groupSize != 64) { throw std::runtime_error("Unsupported spatial group size\n"); }
    std::uint64_t tempBytes = cstone_b200_group_splits_temp_bytes(n, groupSize);
    numSplitsPerGroup.resize(std::size_t(tempBytes / sizeof(LocalIndex)) + 64);
    groups.resize(std::size_t(n) + 1);
    uint32_t numGroups = 0;
    detail::check(detail::GroupSplitsApi<Tc, T, KeyType>::run(
                      first, last, x, y, z, h, leaves, numLeaves, layout, box.c(), groupSize, tolFactor, groups.data(),
                      groups.size(), &numGroups, numSplitsPerGroup.data(),
                      std::uint64_t(numSplitsPerGroup.size()) * sizeof(LocalIndex), nullptr),
                  "computeGroupSplits");
    detail::check(cstone_b200_device_synchronize(), "computeGroupSplits");
    groups.resize(std::size_t(numGroups) + 1);
}

namespace detail
{
template<class K>
struct FocusTreeApi;
#define CSB_FOCUS_TREE_FACADE(KS, K)                                                                                   \
    template<>                                                                                                         \
    struct FocusTreeApi<K>                                                                                             \
    {                                                                                                                  \
        static constexpr auto create    = cstone_b200_focus_tree_create_##KS;                                          \
        static constexpr auto destroy   = cstone_b200_focus_tree_destroy_##KS;                                         \
        static constexpr auto view      = cstone_b200_focus_tree_view_##KS;                                            \
        static constexpr auto setLeaves = cstone_b200_focus_tree_set_leaves_##KS;                                      \
        static int update(void* h, const cstone_box_f32* b, const K* k, std::uint64_t n, K fs, K fe, const K* m,       \
                          int nm, int32_t* c)                                                                          \
        {                                                                                                              \
            return cstone_b200_focus_tree_update_##KS##_f32(h, b, k, n, fs, fe, m, nm, c, nullptr);                    \
        }                                                                                                              \
        static int update(void* h, const cstone_box_f64* b, const K* k, std::uint64_t n, K fs, K fe, const K* m,       \
                          int nm, int32_t* c)                                                                          \
        {                                                                                                              \
            return cstone_b200_focus_tree_update_##KS##_f64(h, b, k, n, fs, fe, m, nm, c, nullptr);                    \
        }                                                                                                              \
        static int macRefine(void* h, const cstone_box_f32* b, K ofs, K ofe, K fs, K fe, float it, int32_t* c)         \
        {                                                                                                              \
            return cstone_b200_focus_tree_mac_refine_##KS##_f32(h, b, ofs, ofe, fs, fe, it, c, nullptr);               \
        }                                                                                                              \
        static int macRefine(void* h, const cstone_box_f64* b, K ofs, K ofe, K fs, K fe, float it, int32_t* c)         \
        {                                                                                                              \
            return cstone_b200_focus_tree_mac_refine_##KS##_f64(h, b, ofs, ofe, fs, fe, it, c, nullptr);               \
        }                                                                                                              \
    };
CSB_FOCUS_TREE_FACADE(k32, uint32_t)
CSB_FOCUS_TREE_FACADE(k64, uint64_t)
#undef CSB_FOCUS_TREE_FACADE
} // namespace detail

/*! @brief A fully traversable octree, locally focused w.r.t. a MinMac criterion (single rank), device resident
 *
 * update() = one step of the reference class: rebalance by particle counts inside [focusStart, focusEnd) and by the
 * MAC outside, enforce the focus boundaries and the mandatory keys, recompute MACs and counts on the new tree.
 */
template<class KeyType>
class FocusedOctreeSingleNode
{
    using Api = detail::FocusTreeApi<KeyType>;

public:
    FocusedOctreeSingleNode(unsigned bucketSize, float theta)
        : h_(Api::create(CSTONE_SFC_HILBERT, bucketSize, theta))
    {
        if (!h_) throw std::runtime_error("FocusedOctreeSingleNode: invalid arguments");
    }
    ~FocusedOctreeSingleNode() { Api::destroy(h_); }
    FocusedOctreeSingleNode(const FocusedOctreeSingleNode&)            = delete;
    FocusedOctreeSingleNode& operator=(const FocusedOctreeSingleNode&) = delete;

    //! @brief one update step; @p particleKeys sorted, on the device; returns true if the tree did not change
    template<class T>
    bool update(const Box<T>& box,
                const KeyType* particleKeys,
                std::size_t numParticles,
                KeyType focusStart,
                KeyType focusEnd,
                const std::vector<KeyType>& mandatoryKeys = {})
    {
        int32_t converged = 0;
        detail::check(Api::update(h_, box.c(), particleKeys, numParticles, focusStart, focusEnd, mandatoryKeys.data(),
                                  int(mandatoryKeys.size()), &converged),
                      "FocusedOctreeSingleNode::update");
        return converged != 0;
    }

    template<class T>
    bool update(const Box<T>& box,
                const DeviceVector<KeyType>& particleKeys,
                KeyType focusStart,
                KeyType focusEnd,
                const std::vector<KeyType>& mandatoryKeys = {})
    {
        return update(box, particleKeys.data(), particleKeys.size(), focusStart, focusEnd, mandatoryKeys);
    }

    //! @brief one step of macRefineGpu after the focus moved; returns true if no leaf had to be split
    template<class T>
    bool macRefine(const Box<T>& box,
                   KeyType oldFocusStart,
                   KeyType oldFocusEnd,
                   KeyType focusStart,
                   KeyType focusEnd,
                   float invTheta)
    {
        int32_t converged = 0;
        detail::check(Api::macRefine(h_, box.c(), oldFocusStart, oldFocusEnd, focusStart, focusEnd, invTheta, &converged),
                      "FocusedOctreeSingleNode::macRefine");
        return converged != 0;
    }

    //! @brief replace the leaf array (device pointer, numLeafNodes + 1 keys)
    void setLeaves(const KeyType* leaves, TreeNodeIndex numLeafNodes)
    {
        detail::check(Api::setLeaves(h_, leaves, numLeafNodes, nullptr), "FocusedOctreeSingleNode::setLeaves");
    }

    DeviceSpan<KeyType> treeLeaves() const
    {
        auto v = view();
        return {static_cast<const KeyType*>(v.leaves), std::size_t(v.num_leaf_nodes) + 1};
    }
    DeviceSpan<unsigned> leafCounts() const
    {
        auto v = view();
        return {v.leaf_counts, v.leaf_counts ? std::size_t(v.num_leaf_nodes) : 0};
    }
    //! @brief the linked tree (device pointers)
    cstone_b200_focus_tree_view view() const
    {
        cstone_b200_focus_tree_view v;
        detail::check(Api::view(h_, &v), "FocusedOctreeSingleNode::view");
        return v;
    }

private:
    void* h_;
};

} // namespace cstone
