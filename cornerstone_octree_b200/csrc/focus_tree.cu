/*
 * focus_tree.cu — a19 (second part): the single-rank focus tree, device resident.
 *
 * Mirrors FocusedOctreeSingleNode<KeyType> [ref: focus/octree_focus.hpp:349-427] and the combined update it drives,
 * CombinedUpdate::updateFocus [ref: focus/octree_focus.hpp:83-135; GPU twin :154-215]:
 *
 *   update(box, particleKeys, focusStart, focusEnd, mandatoryKeys):
 *     1. op per node from counts inside the focus / MAC outside      rebalance_decision_essential   (focus.cu)
 *     2. focusStart, focusEnd and the mandatory keys are enforced    enforce_keys                   (focus.cu)
 *     3. ancestors of kept nodes are protected                       protect_ancestors              (focus.cu)
 *     4. leaf ops = nodeOps[leafToInternal], scan, new leaf array    gather, scan, rebalance_tree   (csarray.cu)
 *     5. keys that could not be introduced are injected              sort_keys, count/fill_sfc_gaps (primitives.cu)
 *     6. internal tree rebuilt                                       build_octree                   (octree.cu)
 *     7. MAC spheres (1/theta + 0.5), MAC marks w.r.t. the focus     geo_mac_spheres, mark_macs     (focus.cu)
 *     8. leaf counts of the particle keys, scattered + summed up     compute_node_counts, upsweep   (csarray/octree.cu)
 *   returns converged (the tree structure did not change), with the reference's rules for the ResolutionStatus.
 *
 * Host side: orchestration only, all arrays stay in HBM; per update the host reads back three words (status,
 * protect-ancestors count, new leaf count) and the level ranges.  The pieces are this library's own C entry points.
 * The multi-rank FocusedOctree (peer treelet exchange, focus/octree_focus_mpi.hpp) is not part of this file.
 */
#include "focus_tree.cuh"

extern "C" {

#define CSB_FOCUS_TREE_C(KS, K)                                                                                        \
    void* cstone_b200_focus_tree_create_##KS(int sfcKind, uint32_t bucketSize, float theta)                            \
    {                                                                                                                  \
        if (theta <= 0.0f || (sfcKind != CSTONE_SFC_MORTON && sfcKind != CSTONE_SFC_HILBERT)) return nullptr;          \
        return new (std::nothrow) csb::FocusTree<K>(sfcKind, bucketSize, theta);                                       \
    }                                                                                                                  \
    void cstone_b200_focus_tree_destroy_##KS(void* handle) { delete static_cast<csb::FocusTree<K>*>(handle); }         \
    int cstone_b200_focus_tree_update_##KS##_f32(void* handle, const cstone_box_f32* box, const K* particleKeys,       \
                                                 uint64_t numKeys, K focusStart, K focusEnd,                           \
                                                 const K* mandatoryKeysHost, int32_t numMandatory,                     \
                                                 int32_t* convergedHost, void* stream)                                 \
    {                                                                                                                  \
        if (!handle || !box || (numKeys && !particleKeys) || numMandatory < 0 || (numMandatory && !mandatoryKeysHost)) \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        return static_cast<csb::FocusTree<K>*>(handle)->update<float>(box, particleKeys, numKeys, focusStart,          \
                                                                      focusEnd, mandatoryKeysHost, numMandatory,       \
                                                                      convergedHost, csb::asStream(stream));           \
    }                                                                                                                  \
    int cstone_b200_focus_tree_update_##KS##_f64(void* handle, const cstone_box_f64* box, const K* particleKeys,       \
                                                 uint64_t numKeys, K focusStart, K focusEnd,                           \
                                                 const K* mandatoryKeysHost, int32_t numMandatory,                     \
                                                 int32_t* convergedHost, void* stream)                                 \
    {                                                                                                                  \
        if (!handle || !box || (numKeys && !particleKeys) || numMandatory < 0 || (numMandatory && !mandatoryKeysHost)) \
            return CSTONE_E_INVALID_ARGUMENT;                                                                          \
        return static_cast<csb::FocusTree<K>*>(handle)->update<double>(box, particleKeys, numKeys, focusStart,         \
                                                                       focusEnd, mandatoryKeysHost, numMandatory,      \
                                                                       convergedHost, csb::asStream(stream));          \
    }                                                                                                                  \
    int cstone_b200_focus_tree_set_leaves_##KS(void* handle, const K* leaves, int32_t numLeafNodes, void* stream)      \
    {                                                                                                                  \
        if (!handle || !leaves || numLeafNodes < 1) return CSTONE_E_INVALID_ARGUMENT;                                  \
        return static_cast<csb::FocusTree<K>*>(handle)->setLeaves(leaves, numLeafNodes, csb::asStream(stream));        \
    }                                                                                                                  \
    int cstone_b200_focus_tree_mac_refine_##KS##_f32(void* handle, const cstone_box_f32* box, K oldFocusStart,         \
                                                     K oldFocusEnd, K focusStart, K focusEnd, float invTheta,          \
                                                     int32_t* convergedHost, void* stream)                             \
    {                                                                                                                  \
        if (!handle || !box) return CSTONE_E_INVALID_ARGUMENT;                                                         \
        return static_cast<csb::FocusTree<K>*>(handle)->macRefine<float>(                                              \
            box, oldFocusStart, oldFocusEnd, focusStart, focusEnd, invTheta, convergedHost, csb::asStream(stream));    \
    }                                                                                                                  \
    int cstone_b200_focus_tree_mac_refine_##KS##_f64(void* handle, const cstone_box_f64* box, K oldFocusStart,         \
                                                     K oldFocusEnd, K focusStart, K focusEnd, float invTheta,          \
                                                     int32_t* convergedHost, void* stream)                             \
    {                                                                                                                  \
        if (!handle || !box) return CSTONE_E_INVALID_ARGUMENT;                                                         \
        return static_cast<csb::FocusTree<K>*>(handle)->macRefine<double>(                                             \
            box, oldFocusStart, oldFocusEnd, focusStart, focusEnd, invTheta, convergedHost, csb::asStream(stream));    \
    }                                                                                                                  \
    int cstone_b200_focus_tree_view_##KS(void* handle, cstone_b200_focus_tree_view* out)                               \
    {                                                                                                                  \
        if (!handle || !out) return CSTONE_E_INVALID_ARGUMENT;                                                         \
        auto* t = static_cast<csb::FocusTree<K>*>(handle);                                                             \
        if (!t->ready_)                                                                                                \
        {                                                                                                              \
            int st = t->init(0);                                                                                       \
            if (st) return st;                                                                                         \
        }                                                                                                              \
        out->num_leaf_nodes     = t->numLeaves_;                                                                       \
        out->num_internal_nodes = t->numInternal_;                                                                     \
        out->num_nodes          = t->numNodes_;                                                                        \
        out->reserved           = 0;                                                                                   \
        out->leaves             = t->leaves_.p;                                                                        \
        out->leaf_counts        = t->leafCounts_.as<uint32_t>();                                                       \
        out->prefixes           = t->prefixes_.p;                                                                      \
        out->child_offsets      = t->childOffsets_.as<int32_t>();                                                      \
        out->parents            = t->parents_.as<int32_t>();                                                           \
        out->level_range        = t->levelRangeDev_.as<int32_t>();                                                     \
        out->internal_to_leaf   = t->i2l_.as<int32_t>();                                                               \
        out->leaf_to_internal   = t->l2i_.as<int32_t>();                                                               \
        out->counts             = t->counts_.as<uint32_t>();                                                           \
        out->macs               = t->macs_.as<char>();                                                                 \
        return 0;                                                                                                      \
    }
CSB_FOCUS_TREE_C(k32, uint32_t)
CSB_FOCUS_TREE_C(k64, uint64_t)

} // extern "C"
