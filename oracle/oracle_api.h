/*
 * oracle_api.h — TEST INFRASTRUCTURE ONLY.
 *
 * C API shared by the two CPU checkers of the hot path:
 *   - liboracle_port.so  (oracle/cstone_oracle.cpp): our own CPU restatement of the
 *     reference algorithms, prefix  cstone_oracle_
 *   - oracle/_ref/liboracle_ref.so (oracle/ref_wrap.cpp): the UNMODIFIED reference
 *     headers from /root/reference/include compiled behind the same API,
 *     prefix  cstone_ref_
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load these libraries.  The product (libcstone_b200.so) never does.
 *
 * All pointers are HOST pointers.  Shapes follow include/cstone_b200.h without the
 * stream / temp arguments.
 */
#ifndef CSTONE_ORACLE_API_H
#define CSTONE_ORACLE_API_H

#include <stddef.h>
#include <stdint.h>
#include "../include/cstone_b200.h" /* cstone_box_f32 / cstone_box_f64, CSTONE_SFC_* */

#ifndef ORACLE_PREFIX
#error "define ORACLE_PREFIX (cstone_oracle_ or cstone_ref_) before including oracle_api.h"
#endif
#define ORC_CAT2(a, b) a##b
#define ORC_CAT(a, b) ORC_CAT2(a, b)
#define ORC(name) ORC_CAT(ORACLE_PREFIX, name)

#ifdef __cplusplus
extern "C" {
#endif

int ORC(num_threads)(void);
void ORC(set_num_threads)(int n);

/* --- SFC --- */
uint32_t ORC(encode_k32)(int kind, unsigned ix, unsigned iy, unsigned iz);
uint64_t ORC(encode_k64)(int kind, unsigned ix, unsigned iy, unsigned iz);
void ORC(decode_k32)(int kind, uint32_t key, unsigned* xyz);
void ORC(decode_k64)(int kind, uint64_t key, unsigned* xyz);
void ORC(compute_sfc_keys_k32_f32)(int kind, const float* x, const float* y, const float* z, uint32_t* keys,
                                   size_t n, const cstone_box_f32* box);
void ORC(compute_sfc_keys_k32_f64)(int kind, const double* x, const double* y, const double* z, uint32_t* keys,
                                   size_t n, const cstone_box_f64* box);
void ORC(compute_sfc_keys_k64_f32)(int kind, const float* x, const float* y, const float* z, uint64_t* keys,
                                   size_t n, const cstone_box_f32* box);
void ORC(compute_sfc_keys_k64_f64)(int kind, const double* x, const double* y, const double* z, uint64_t* keys,
                                   size_t n, const cstone_box_f64* box);
/* IBox {xmin,xmax,ymin,ymax,zmin,zmax} of the node starting at `key` at `level` (Hilbert / Morton) */
void ORC(sfc_ibox_k32)(int kind, uint32_t key, unsigned level, int* ibox);
void ORC(sfc_ibox_k64)(int kind, uint64_t key, unsigned level, int* ibox);

/* --- sort / gather --- */
void ORC(sort_by_key_k32)(uint32_t* keys, uint32_t* values, size_t n);
void ORC(sort_by_key_k64)(uint64_t* keys, uint32_t* values, size_t n);
void ORC(gather)(const uint32_t* map, size_t n, const void* src, void* dst, int elem_bytes);

/* --- cornerstone leaf array --- */
void ORC(compute_node_counts_k32)(const uint32_t* tree, uint32_t* counts, int32_t num_nodes, const uint32_t* first_key,
                                  const uint32_t* last_key, uint32_t max_count, int use_counts_as_guess);
void ORC(compute_node_counts_k64)(const uint64_t* tree, uint32_t* counts, int32_t num_nodes, const uint64_t* first_key,
                                  const uint64_t* last_key, uint32_t max_count, int use_counts_as_guess);
/* per-leaf ops (un-scanned, num_nodes entries); returns converged */
int ORC(rebalance_decision_k32)(const uint32_t* tree, const uint32_t* counts, int32_t num_nodes, uint32_t bucket_size,
                                int32_t* node_ops);
int ORC(rebalance_decision_k64)(const uint64_t* tree, const uint32_t* counts, int32_t num_nodes, uint32_t bucket_size,
                                int32_t* node_ops);
/* node_ops (num_nodes+1, un-scanned on input, scanned on output); new_tree capacity >= sum(ops)+1; returns new #nodes */
int32_t ORC(rebalance_tree_k32)(const uint32_t* tree, int32_t num_nodes, int32_t* node_ops, uint32_t* new_tree,
                                int32_t capacity);
int32_t ORC(rebalance_tree_k64)(const uint64_t* tree, int32_t num_nodes, int32_t* node_ops, uint64_t* new_tree,
                                int32_t capacity);
/* one updateOctree step on tree/counts held in caller buffers of `capacity` nodes; returns converged (0/1) or -1 if
 * capacity is insufficient; *num_nodes in/out */
int ORC(update_octree_k32)(const uint32_t* first_key, const uint32_t* last_key, uint32_t bucket_size,
                           uint32_t max_count, uint32_t* tree, uint32_t* counts, int32_t capacity, int32_t* num_nodes);
int ORC(update_octree_k64)(const uint64_t* first_key, const uint64_t* last_key, uint32_t bucket_size,
                           uint32_t max_count, uint64_t* tree, uint32_t* counts, int32_t capacity, int32_t* num_nodes);
/* full build; returns number of leaf nodes L (tree has L+1 entries) or -(needed) if capacity < L */
int32_t ORC(compute_octree_k32)(const uint32_t* first_key, const uint32_t* last_key, uint32_t bucket_size,
                                uint32_t max_count, uint32_t* tree, uint32_t* counts, int32_t capacity);
int32_t ORC(compute_octree_k64)(const uint64_t* first_key, const uint64_t* last_key, uint32_t bucket_size,
                                uint32_t max_count, uint64_t* tree, uint32_t* counts, int32_t capacity);

/* --- internal tree --- */
void ORC(build_octree_k32)(const uint32_t* leaves, int32_t num_leaf_nodes, uint32_t* prefixes, int32_t* child_offsets,
                           int32_t* parents, int32_t* level_range, int32_t* internal_to_leaf,
                           int32_t* leaf_to_internal);
void ORC(build_octree_k64)(const uint64_t* leaves, int32_t num_leaf_nodes, uint64_t* prefixes, int32_t* child_offsets,
                           int32_t* parents, int32_t* level_range, int32_t* internal_to_leaf,
                           int32_t* leaf_to_internal);
/* cornerstone keys spanning [a, b) with the fewest nodes; out may be NULL (count only)
 * [ref: sfc/common.hpp:392-438 spanSfcRange; used by countSfcGapsGpu / fillSfcGapsGpu, tree/csarray_gpu.cu:231-265] */
int ORC(span_sfc_range_k32)(uint32_t a, uint32_t b, uint32_t* out);
int ORC(span_sfc_range_k64)(uint64_t a, uint64_t b, uint64_t* out);
void ORC(upsweep_sum)(int32_t num_levels, const int32_t* level_range, const int32_t* child_offsets, uint32_t* counts);
void ORC(geo_centers_k32_f32)(int kind, const uint32_t* prefixes, int32_t n, float* centers, float* sizes,
                              const cstone_box_f32* box);
void ORC(geo_centers_k32_f64)(int kind, const uint32_t* prefixes, int32_t n, double* centers, double* sizes,
                              const cstone_box_f64* box);
void ORC(geo_centers_k64_f32)(int kind, const uint64_t* prefixes, int32_t n, float* centers, float* sizes,
                              const cstone_box_f32* box);
void ORC(geo_centers_k64_f64)(int kind, const uint64_t* prefixes, int32_t n, double* centers, double* sizes,
                              const cstone_box_f64* box);

/* --- halos --- */
void ORC(halo_radii_f32)(const float* h, const uint32_t* segments, size_t num_segments, float scale, float* radii);
void ORC(halo_radii_f64)(const double* h, const uint32_t* segments, size_t num_segments, float scale, float* radii);
void ORC(find_halos_k32_f32)(int kind, const uint32_t* prefixes, const int32_t* child_offsets,
                             const int32_t* internal_to_leaf, const uint32_t* leaves, const float* radii,
                             const cstone_box_f32* box, int32_t first_node, int32_t last_node, int32_t* flags);
void ORC(find_halos_k32_f64)(int kind, const uint32_t* prefixes, const int32_t* child_offsets,
                             const int32_t* internal_to_leaf, const uint32_t* leaves, const float* radii,
                             const cstone_box_f64* box, int32_t first_node, int32_t last_node, int32_t* flags);
void ORC(find_halos_k64_f32)(int kind, const uint64_t* prefixes, const int32_t* child_offsets,
                             const int32_t* internal_to_leaf, const uint64_t* leaves, const float* radii,
                             const cstone_box_f32* box, int32_t first_node, int32_t last_node, int32_t* flags);
void ORC(find_halos_k64_f64)(int kind, const uint64_t* prefixes, const int32_t* child_offsets,
                             const int32_t* internal_to_leaf, const uint64_t* leaves, const float* radii,
                             const cstone_box_f64* box, int32_t first_node, int32_t last_node, int32_t* flags);
void ORC(compute_node_layout)(const uint32_t* counts, const int32_t* flags, int32_t first_assigned,
                              int32_t last_assigned, int32_t num_leaf_nodes, uint32_t* layout);

/* --- neighbours --- */
void ORC(find_neighbors_f32)(const float* x, const float* y, const float* z, const float* h, uint32_t first_id,
                             uint32_t last_id, const cstone_box_f32* box, const int32_t* child_offsets,
                             const int32_t* internal_to_leaf, const uint32_t* layout, const float* centers,
                             const float* sizes, float search_ext_factor, uint32_t ngmax, uint32_t* neighbors,
                             uint32_t* neighbors_count);
void ORC(find_neighbors_f64)(const double* x, const double* y, const double* z, const double* h, uint32_t first_id,
                             uint32_t last_id, const cstone_box_f64* box, const int32_t* child_offsets,
                             const int32_t* internal_to_leaf, const uint32_t* layout, const double* centers,
                             const double* sizes, float search_ext_factor, uint32_t ngmax, uint32_t* neighbors,
                             uint32_t* neighbors_count);

/* --- domain decomposition, host logic (reference build only: liboracle_ref.so)
 * [ref: domain/domaindecomp.hpp:48-71 uniformBins, :115-124 makeSfcAssignment, :139-166 limitBoundaryShifts,
 *  sfc/box.hpp:415-431 limitBoxShrinking] --- */
void cstone_ref_assignment_k32(int nranks, const uint32_t* leaves, const uint32_t* counts, int32_t num_leaves,
                               const uint32_t* old_boundaries, uint32_t* boundaries, uint32_t* rank_counts);
void cstone_ref_assignment_k64(int nranks, const uint64_t* leaves, const uint32_t* counts, int32_t num_leaves,
                               const uint64_t* old_boundaries, uint64_t* boundaries, uint32_t* rank_counts);
void cstone_ref_uniform_bins(const uint32_t* counts, int32_t num_leaves, int32_t num_bins, int32_t* bins,
                             uint32_t* bin_counts);
void cstone_ref_limit_boundary_shifts_k32(int nranks, const uint32_t* old_boundaries, uint32_t* boundaries,
                                          const uint32_t* leaves, const uint32_t* counts, int32_t num_leaves,
                                          uint32_t* rank_counts);
void cstone_ref_limit_boundary_shifts_k64(int nranks, const uint64_t* old_boundaries, uint64_t* boundaries,
                                          const uint64_t* leaves, const uint32_t* counts, int32_t num_leaves,
                                          uint32_t* rank_counts);
void cstone_ref_limit_box_shrinking_f32(const float* fitting6, const float* previous6, float* out6);
void cstone_ref_limit_box_shrinking_f64(const double* fitting6, const double* previous6, double* out6);

#ifdef __cplusplus
}
#endif
#endif
