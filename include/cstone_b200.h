/*
 * cstone_b200.h — C ABI of libcstone_b200.so
 *
 * B200 (sm_100a) implementation of the cornerstone-octree domain hot path:
 * SFC keys -> key sort + gather -> cornerstone leaf array count/rebalance ->
 * internal-tree linking -> halo discovery -> radius neighbour search.
 *
 * Every entry point below replaces one `extern template` GPU function of the
 * reference library `cstone_gpu` (the only FFI-like boundary the reference
 * has).  The reference declaration each one stands in for is
 * cited as  [ref: path:line]  relative to the reference's include/cstone/.
 *
 * Conventions
 *  - plain C: raw DEVICE pointers and sizes, no C++ / torch types.
 *  - key type suffix  _k32 (uint32_t) / _k64 (uint64_t); coordinate type
 *    suffix _f32 / _f64.  `sfc_kind` selects the curve at run time
 *    (CSTONE_SFC_MORTON / CSTONE_SFC_HILBERT) where the reference selects it
 *    through the MortonKey<> / HilbertKey<> strong types.
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *  - return value: 0 on success, a positive cudaError_t, or a negative
 *    CSTONE_E_* logic error.  Nothing is printed and the process is never
 *    terminated by the library (the reference wrapper in INTEGRATION.md maps
 *    non-zero to the reference behaviour: print + exit, or std::runtime_error).
 *  - the callee never allocates caller-visible memory; scratch space is
 *    caller-provided and sized by the *_temp_bytes functions.
 *  - functions that return a scalar result through a HOST pointer
 *    synchronise `stream` before returning (same as the reference, which
 *    returns those values by value).
 *  - there is no CPU fallback: without a CUDA device every call returns the
 *    CUDA error.
 */
#ifndef CSTONE_B200_H
#define CSTONE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CSTONE_B200_ABI_VERSION 1

#define CSTONE_SFC_MORTON 0
#define CSTONE_SFC_HILBERT 1

/* boundary types, same values as cstone::BoundaryType [ref: sfc/box.hpp:89-94] */
#define CSTONE_BOUNDARY_OPEN 0
#define CSTONE_BOUNDARY_PERIODIC 1
#define CSTONE_BOUNDARY_FIXED 2

/* logic errors (negative so that they can not collide with cudaError_t) */
#define CSTONE_E_INVALID_ARGUMENT (-1)
#define CSTONE_E_TEMP_TOO_SMALL (-2) /* [ref: primitives/primitives_gpu.cu:353] throws runtime_error */
#define CSTONE_E_TOO_LARGE (-3)
#define CSTONE_E_STACK_OVERFLOW (-4) /* traversal stack exhausted [ref: traversal/traversal.hpp:102] */
#define CSTONE_E_CAPACITY (-5)

/* Memory layout identical to cstone::Box<T> [ref: sfc/box.hpp:111-191]:
 * limits[6] = xmin,xmax,ymin,ymax,zmin,zmax; lengths; inverse lengths (computed
 * ONCE on the host as T(1)/(max-min), never recomputed on the device);
 * boundaries[3].  A `const cstone::Box<T>&` may be reinterpret_cast to it. */
typedef struct cstone_box_f32
{
    float limits[6];
    float lengths[3];
    float inv_lengths[3];
    char boundaries[3];
} cstone_box_f32;

typedef struct cstone_box_f64
{
    double limits[6];
    double lengths[3];
    double inv_lengths[3];
    char boundaries[3];
} cstone_box_f64;

int cstone_b200_abi_version(void);
/* human readable text for a status returned by any function below */
const char* cstone_b200_status_string(int status);
/* number of kernels launched by this library in the calling process so far */
uint64_t cstone_b200_launch_count(void);

/* device memory behind cstone::DeviceVector<T> [ref: cuda/device_vector.h:24-62, cuda/device_vector.cu], so that a
 * host program built with a plain C++ compiler needs no CUDA headers.  kind: 0 = host->device, 1 = device->host,
 * 2 = device->device, on the legacy stream, complete on return. */
int cstone_b200_device_malloc(void** ptr, size_t bytes);
int cstone_b200_device_free(void* ptr);
int cstone_b200_device_memcpy(void* dst, const void* src, size_t bytes, int kind);
int cstone_b200_device_memset(void* dst, int value, size_t bytes);
int cstone_b200_device_synchronize(void);
/* one process per GPU: select the device of the calling thread (cudaSetDevice) */
int cstone_b200_set_device(int device);

/* ------------------------------------------------------------------------
 * a1/a2  SFC keys
 * [ref: sfc/sfc_gpu.h:38-39  computeSfcKeysGpu<KeyType,T>(x,y,z,keys,n,box)]
 * keys[i] = sfc3D(x[i],y[i],z[i],box) unless keys[i] == removeKey (1<<3*maxLevel)
 * ---------------------------------------------------------------------- */
int cstone_b200_compute_sfc_keys_k32_f32(int sfc_kind, const float* x, const float* y, const float* z,
                                         uint32_t* keys, size_t n, const cstone_box_f32* box, void* stream);
int cstone_b200_compute_sfc_keys_k32_f64(int sfc_kind, const double* x, const double* y, const double* z,
                                         uint32_t* keys, size_t n, const cstone_box_f64* box, void* stream);
int cstone_b200_compute_sfc_keys_k64_f32(int sfc_kind, const float* x, const float* y, const float* z,
                                         uint64_t* keys, size_t n, const cstone_box_f32* box, void* stream);
int cstone_b200_compute_sfc_keys_k64_f64(int sfc_kind, const double* x, const double* y, const double* z,
                                         uint64_t* keys, size_t n, const cstone_box_f64* box, void* stream);

/* integer-coordinate encode / decode of n points (used by the known-answer
 * tests; [ref: sfc/morton.hpp:111-168, sfc/hilbert.hpp:58-109,145-188]) */
int cstone_b200_encode_sfc_k32(int sfc_kind, const uint32_t* ix, const uint32_t* iy, const uint32_t* iz,
                               uint32_t* keys, size_t n, void* stream);
int cstone_b200_encode_sfc_k64(int sfc_kind, const uint32_t* ix, const uint32_t* iy, const uint32_t* iz,
                               uint64_t* keys, size_t n, void* stream);
int cstone_b200_decode_sfc_k32(int sfc_kind, const uint32_t* keys, uint32_t* ix, uint32_t* iy, uint32_t* iz,
                               size_t n, void* stream);
int cstone_b200_decode_sfc_k64(int sfc_kind, const uint64_t* keys, uint32_t* ix, uint32_t* iy, uint32_t* iz,
                               size_t n, void* stream);

/* ------------------------------------------------------------------------
 * a3  coordinate extrema   [ref: primitives/primitives_gpu.h:58-62 MinMaxGpu<T>]
 * One fused pass over x,y,z (any of which may be NULL); out_minmax is a HOST
 * array {xmin,xmax,ymin,ymax,zmin,zmax}; synchronises the stream.
 * `temp` = device scratch of cstone_b200_minmax_temp_bytes() bytes.
 * ---------------------------------------------------------------------- */
uint64_t cstone_b200_minmax_temp_bytes(void);
int cstone_b200_minmax3_f32(const float* x, const float* y, const float* z, size_t n, float* out_minmax,
                            void* temp, void* stream);
int cstone_b200_minmax3_f64(const double* x, const double* y, const double* z, size_t n, double* out_minmax,
                            void* temp, void* stream);

/* ------------------------------------------------------------------------
 * a4  stable key sort carrying a 32-bit permutation (hand-written onesweep
 * LSD radix sort, 8-bit digits, full key width).
 * [ref: primitives/primitives_gpu.h:104-110 sortByKeyTempStorage / sortByKeyGpu,
 *       primitives/gather.cuh:80-108 GpuSfcSorter]
 *  keys[n], values[n]   in/out (sorted result always lands here)
 *  key_buf[n], value_buf[n]  caller-provided alternate buffers
 *  temp / temp_bytes    >= cstone_b200_sort_by_key_temp_bytes_*(n)
 *  values_are_iota != 0: `values` is treated as 0,1,2,... on input without being
 *  read (fuses sequenceGpu [ref: primitives/primitives_gpu.h:87] into pass 0).
 * n must be < 2^30 per call (CSTONE_E_TOO_LARGE otherwise).
 * ---------------------------------------------------------------------- */
uint64_t cstone_b200_sort_by_key_temp_bytes_k32(uint64_t n);
uint64_t cstone_b200_sort_by_key_temp_bytes_k64(uint64_t n);
int cstone_b200_sort_by_key_k32(uint32_t* keys, uint32_t* values, uint32_t* key_buf, uint32_t* value_buf, size_t n,
                                void* temp, uint64_t temp_bytes, int values_are_iota, void* stream);
int cstone_b200_sort_by_key_k64(uint64_t* keys, uint32_t* values, uint64_t* key_buf, uint32_t* value_buf, size_t n,
                                void* temp, uint64_t temp_bytes, int values_are_iota, void* stream);
/* [ref: primitives/primitives_gpu.h:86-87 sequenceGpu] out[i] = init + i */
int cstone_b200_sequence_u32(uint32_t* out, size_t n, uint32_t init, void* stream);

/* ------------------------------------------------------------------------
 * a5  gather / scatter through a LocalIndex map
 * [ref: primitives/primitives_gpu.h:48-56 gatherGpu / scatterGpu]
 *  dst[i] = src[map[i]]  (gather),  dst[map[i]] = src[i]  (scatter)
 * elem_bytes in {4,8,12,16,24,32}.  gather_multi moves `num_arrays` arrays with ONE
 * read of the map (src/dst/elem_bytes are HOST arrays of length num_arrays <= 16).
 * ---------------------------------------------------------------------- */
int cstone_b200_gather(const uint32_t* map, size_t n, const void* src, void* dst, int elem_bytes, void* stream);
int cstone_b200_scatter(const uint32_t* map, size_t n, const void* src, void* dst, int elem_bytes, void* stream);
int cstone_b200_gather_multi(const uint32_t* map, size_t n, int num_arrays, const void* const* src, void* const* dst,
                             const int* elem_bytes, void* stream);

/* ------------------------------------------------------------------------
 * scans  [ref: primitives/primitives_gpu.h:112-122 exclusiveScanGpu / inclusiveScanGpu]
 * single-pass decoupled look-back scan of n u32 values, out may alias in.
 * temp >= cstone_b200_scan_temp_bytes(n).
 * ---------------------------------------------------------------------- */
uint64_t cstone_b200_scan_temp_bytes(uint64_t n);
int cstone_b200_exclusive_scan_u32(const uint32_t* in, uint32_t* out, size_t n, uint32_t init, void* temp,
                                   void* stream);
int cstone_b200_inclusive_scan_u32(const uint32_t* in, uint32_t* out, size_t n, void* temp, void* stream);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as cstone_b200_exclusive_scan_u32, but the scratch lives in a file-scope device array inside the
 * library (65 792 bytes, held by every GPU that loads the library until the process exits), so the call allocates
 * nothing: no cudaMalloc / cudaFree / allocating library routine.
 * Limit: n <= CSTONE_B200_SCAN_NOALLOC_MAX_N, CSTONE_E_TOO_LARGE otherwise (nothing is launched).
 * All calls on one GPU share the array; each call clears the part it uses on `stream` first.  Calls that overlap on
 * the GPU (streams without an ordering between them) race on it: order them. */
#define CSTONE_B200_SCAN_NOALLOC_MAX_N 16777216u
int cstone_b200_exclusive_scan_u32_noalloc(const uint32_t* in, uint32_t* out, size_t n, uint32_t init, void* stream);

/* ------------------------------------------------------------------------
 * a6  leaf particle counts
 * [ref: tree/csarray_gpu.h:55-62 computeNodeCountsGpu<KeyType>]
 * counts[i] = min(#keys in [tree[i], tree[i+1]), max_count).  The result does not
 * depend on use_counts_as_guess (the guess only steers the reference's search).
 * ---------------------------------------------------------------------- */
int cstone_b200_compute_node_counts_k32(const uint32_t* tree, uint32_t* counts, int32_t num_nodes,
                                        const uint32_t* first_key, const uint32_t* last_key, uint32_t max_count,
                                        int use_counts_as_guess, void* stream);
int cstone_b200_compute_node_counts_k64(const uint64_t* tree, uint32_t* counts, int32_t num_nodes,
                                        const uint64_t* first_key, const uint64_t* last_key, uint32_t max_count,
                                        int use_counts_as_guess, void* stream);

/* ------------------------------------------------------------------------
 * a7  rebalance decision + exclusive scan of the node ops
 * [ref: tree/csarray_gpu.h:75-77 computeNodeOpsGpu<KeyType>]
 * node_ops has num_nodes+1 entries and holds the exclusive scan on return;
 * *new_num_nodes (HOST) = node_ops[num_nodes]; *converged (HOST, may be NULL)
 * = 1 iff every op was 1.  temp >= cstone_b200_scan_temp_bytes(num_nodes+1).
 * ---------------------------------------------------------------------- */
int cstone_b200_compute_node_ops_k32(const uint32_t* tree, int32_t num_nodes, const uint32_t* counts,
                                     uint32_t bucket_size, int32_t* node_ops, int32_t* new_num_nodes,
                                     int32_t* converged, void* temp, void* stream);
int cstone_b200_compute_node_ops_k64(const uint64_t* tree, int32_t num_nodes, const uint32_t* counts,
                                     uint32_t bucket_size, int32_t* node_ops, int32_t* new_num_nodes,
                                     int32_t* converged, void* temp, void* stream);
/* the un-scanned per-leaf decision only (0 merge / 1 keep / 8,64,512,4096 split)
 * [ref: tree/csarray.hpp:286-348 calculateNodeOp / rebalanceDecision] */
int cstone_b200_rebalance_decision_k32(const uint32_t* tree, int32_t num_nodes, const uint32_t* counts,
                                       uint32_t bucket_size, int32_t* node_ops, void* stream);
int cstone_b200_rebalance_decision_k64(const uint64_t* tree, int32_t num_nodes, const uint32_t* counts,
                                       uint32_t bucket_size, int32_t* node_ops, void* stream);

/* ------------------------------------------------------------------------
 * a8  emit the rebalanced leaf array
 * [ref: tree/csarray_gpu.h:79-84 rebalanceTreeGpu<KeyType>]
 * new_tree has new_num_nodes+1 entries; node_ops is the scanned array.
 * ---------------------------------------------------------------------- */
int cstone_b200_rebalance_tree_k32(const uint32_t* tree, int32_t num_nodes, int32_t new_num_nodes,
                                   const int32_t* node_ops, uint32_t* new_tree, void* stream);
int cstone_b200_rebalance_tree_k64(const uint64_t* tree, int32_t num_nodes, int32_t new_num_nodes,
                                   const int32_t* node_ops, uint64_t* new_tree, void* stream);

/* ------------------------------------------------------------------------
 * a9  one update step / full build of the cornerstone leaf array
 * [ref: tree/update_gpu.cuh:58-82 updateOctreeGpu; tree/csarray.hpp:429-465
 *       updateOctree / computeOctree]
 * tree / counts are ping-pong device buffers of `capacity` (+1 for trees) entries:
 *   tree_a[capacity+1] in/out, tree_b[capacity+1] scratch, counts[capacity] in/out,
 *   node_ops[capacity+1] scratch, temp >= cstone_b200_scan_temp_bytes(capacity+1).
 * update_octree: *num_nodes (HOST, in/out); *converged (HOST out).  The updated
 *   tree is left in tree_a.  Returns CSTONE_E_CAPACITY if the new tree does not fit
 *   (*num_nodes then holds the required node count, tree_a/counts are unchanged).
 * compute_octree: starts from the root {0, 2^(3*maxLevel)} and iterates
 *   update_octree until converged; *num_nodes / *num_iterations are HOST outputs.
 * ---------------------------------------------------------------------- */
int cstone_b200_update_octree_k32(const uint32_t* first_key, const uint32_t* last_key, uint32_t bucket_size,
                                  uint32_t max_count, uint32_t* tree_a, uint32_t* tree_b, uint32_t* counts,
                                  int32_t* node_ops, int32_t capacity, int32_t* num_nodes, int32_t* converged,
                                  void* temp, void* stream);
int cstone_b200_update_octree_k64(const uint64_t* first_key, const uint64_t* last_key, uint32_t bucket_size,
                                  uint32_t max_count, uint64_t* tree_a, uint64_t* tree_b, uint32_t* counts,
                                  int32_t* node_ops, int32_t capacity, int32_t* num_nodes, int32_t* converged,
                                  void* temp, void* stream);
int cstone_b200_compute_octree_k32(const uint32_t* first_key, const uint32_t* last_key, uint32_t bucket_size,
                                   uint32_t max_count, uint32_t* tree_a, uint32_t* tree_b, uint32_t* counts,
                                   int32_t* node_ops, int32_t capacity, int32_t* num_nodes, int32_t* num_iterations,
                                   void* temp, void* stream);
int cstone_b200_compute_octree_k64(const uint64_t* first_key, const uint64_t* last_key, uint32_t bucket_size,
                                   uint32_t max_count, uint64_t* tree_a, uint64_t* tree_b, uint32_t* counts,
                                   int32_t* node_ops, int32_t capacity, int32_t* num_nodes, int32_t* num_iterations,
                                   void* temp, void* stream);

/* ------------------------------------------------------------------------
 * a10  internal-tree linking
 * [ref: tree/octree_gpu.h:49-50 buildOctreeGpu<KeyType>(cstoneTree, OctreeView d);
 *       OctreeView tree/octree.hpp:279-293; sizes OctreeData::resize :327-342]
 * With L = num_leaf_nodes, I = (L-1)/7, M = L+I the caller provides
 *   prefixes[M], child_offsets[M(+1)], parents[max(1,(M-1)/8)],
 *   level_range[max_tree_level+2] (12 for k32, 23 for k64), internal_to_leaf[M],
 *   leaf_to_internal[M];  temp >= cstone_b200_build_octree_temp_bytes_*(L).
 * ---------------------------------------------------------------------- */
uint64_t cstone_b200_build_octree_temp_bytes_k32(int32_t num_leaf_nodes);
uint64_t cstone_b200_build_octree_temp_bytes_k64(int32_t num_leaf_nodes);
int cstone_b200_build_octree_k32(const uint32_t* leaves, int32_t num_leaf_nodes, uint32_t* prefixes,
                                 int32_t* child_offsets, int32_t* parents, int32_t* level_range,
                                 int32_t* internal_to_leaf, int32_t* leaf_to_internal, void* temp,
                                 uint64_t temp_bytes, void* stream);
int cstone_b200_build_octree_k64(const uint64_t* leaves, int32_t num_leaf_nodes, uint64_t* prefixes,
                                 int32_t* child_offsets, int32_t* parents, int32_t* level_range,
                                 int32_t* internal_to_leaf, int32_t* leaf_to_internal, void* temp,
                                 uint64_t temp_bytes, void* stream);
/* [ref: tree/octree_gpu.h:52-53 upsweepSumGpu] saturating child sum, level_range is a HOST pointer */
int cstone_b200_upsweep_sum(int32_t num_levels, const int32_t* level_range_host, const int32_t* child_offsets,
                            uint32_t* counts, void* stream);

/* ------------------------------------------------------------------------
 * a11  geometric node centres and half sizes
 * [ref: focus/source_center_gpu.h computeGeoCentersGpu<KeyType,T>(prefixes,numNodes,centers,sizes,box);
 *       sfc/box.hpp:334-351 centerAndSize]   centers/sizes are Vec3<T>[num_nodes] (AoS, 3 T each)
 * ---------------------------------------------------------------------- */
int cstone_b200_compute_geo_centers_k32_f32(int sfc_kind, const uint32_t* prefixes, int32_t num_nodes, float* centers,
                                            float* sizes, const cstone_box_f32* box, void* stream);
int cstone_b200_compute_geo_centers_k32_f64(int sfc_kind, const uint32_t* prefixes, int32_t num_nodes,
                                            double* centers, double* sizes, const cstone_box_f64* box, void* stream);
int cstone_b200_compute_geo_centers_k64_f32(int sfc_kind, const uint64_t* prefixes, int32_t num_nodes, float* centers,
                                            float* sizes, const cstone_box_f32* box, void* stream);
int cstone_b200_compute_geo_centers_k64_f64(int sfc_kind, const uint64_t* prefixes, int32_t num_nodes,
                                            double* centers, double* sizes, const cstone_box_f64* box, void* stream);

/* ------------------------------------------------------------------------
 * a12  halo search radius per leaf
 * [ref: primitives/primitives_gpu.h:84-85 segmentMax + scaleGpu as used by
 *       halos/halos.hpp:148-160 Halos::discover]
 * radii[i] = float(max(0, max h[segments[i] .. segments[i+1]))) * scale  for
 * i in [0,num_segments); `segments` has num_segments+1 entries.
 * ---------------------------------------------------------------------- */
int cstone_b200_halo_radii_f32(const float* h, const uint32_t* segments, size_t num_segments, float scale,
                               float* radii, void* stream);
int cstone_b200_halo_radii_f64(const double* h, const uint32_t* segments, size_t num_segments, float scale,
                               float* radii, void* stream);

/* ------------------------------------------------------------------------
 * a13  halo discovery by 3D box collision
 * [ref: traversal/collisions_gpu.h:58-67 findHalosGpu<KeyType,RadiusType,T>]
 * flags[num_leaf_nodes] must be zeroed by the caller; flags of colliding leaves
 * outside [first_node,last_node) are set to 1.  The reference hard-wires the
 * Hilbert curve here (SfcKind); sfc_kind = CSTONE_SFC_MORTON is an extension.
 * ---------------------------------------------------------------------- */
int cstone_b200_find_halos_k32_f32(int sfc_kind, const uint32_t* prefixes, const int32_t* child_offsets,
                                   const int32_t* internal_to_leaf, const uint32_t* leaves, const float* radii,
                                   const cstone_box_f32* box, int32_t first_node, int32_t last_node, int32_t* flags,
                                   void* stream);
int cstone_b200_find_halos_k32_f64(int sfc_kind, const uint32_t* prefixes, const int32_t* child_offsets,
                                   const int32_t* internal_to_leaf, const uint32_t* leaves, const float* radii,
                                   const cstone_box_f64* box, int32_t first_node, int32_t last_node, int32_t* flags,
                                   void* stream);
int cstone_b200_find_halos_k64_f32(int sfc_kind, const uint64_t* prefixes, const int32_t* child_offsets,
                                   const int32_t* internal_to_leaf, const uint64_t* leaves, const float* radii,
                                   const cstone_box_f32* box, int32_t first_node, int32_t last_node, int32_t* flags,
                                   void* stream);
int cstone_b200_find_halos_k64_f64(int sfc_kind, const uint64_t* prefixes, const int32_t* child_offsets,
                                   const int32_t* internal_to_leaf, const uint64_t* leaves, const float* radii,
                                   const cstone_box_f64* box, int32_t first_node, int32_t last_node, int32_t* flags,
                                   void* stream);

/* ------------------------------------------------------------------------
 * a14  particle offsets of the present (assigned or halo) leaves
 * [ref: domain/layout.hpp:150-164 computeNodeLayout]
 * layout[i] = exclusive scan of ((first<=i<last) || flags[i]) ? counts[i] : 0,
 * layout has num_leaf_nodes+1 entries.  temp >= cstone_b200_scan_temp_bytes(num_leaf_nodes+1)
 * ---------------------------------------------------------------------- */
int cstone_b200_compute_node_layout(const uint32_t* counts, const int32_t* flags, int32_t first_assigned,
                                    int32_t last_assigned, int32_t num_leaf_nodes, uint32_t* layout, void* temp,
                                    void* stream);

/* ------------------------------------------------------------------------
 * a15  radius neighbour search, CPU semantics and output layout
 * [ref: findneighbors.hpp:167-178 findNeighbors(x,y,z,h,firstId,lastId,box,OctreeNsView,ngmax,
 *       neighbors,neighborsCount); OctreeNsView tree/octree.hpp:296-317]
 * (the reference library has no host entry point for its GPU traversal,
 *  traversal/find_neighbors.cuh:435-446 is a device function; this one mirrors
 *  the CPU driver.)  neighbors[(i-first_id)*ngmax + k] in the reference's
 * depth-first visiting order, j != i, strict d2 < 4 h_i^2; neighbors_count holds
 * the full count even when it exceeds ngmax.  centers/sizes: Vec3<T>[numNodes].
 * ---------------------------------------------------------------------- */
int cstone_b200_find_neighbors_f32(const float* x, const float* y, const float* z, const float* h, uint32_t first_id,
                                   uint32_t last_id, const cstone_box_f32* box, const int32_t* child_offsets,
                                   const int32_t* internal_to_leaf, const uint32_t* layout, const float* centers,
                                   const float* sizes, float search_ext_factor, uint32_t ngmax, uint32_t* neighbors,
                                   uint32_t* neighbors_count, void* stream);
int cstone_b200_find_neighbors_f64(const double* x, const double* y, const double* z, const double* h,
                                   uint32_t first_id, uint32_t last_id, const cstone_box_f64* box,
                                   const int32_t* child_offsets, const int32_t* internal_to_leaf,
                                   const uint32_t* layout, const double* centers, const double* sizes,
                                   float search_ext_factor, uint32_t ngmax, uint32_t* neighbors,
                                   uint32_t* neighbors_count, void* stream);

/* a15 with caller-provided workspace (same results; the preferred entry point): temp holds one 8-byte record
 * per tree node (first child / particle range of a leaf), rebuilt on every call, so that staging a node during the
 * walk is one round of independent loads [replaces the OctreeNsView member walk of findneighbors.hpp:120-141].
 * temp >= cstone_b200_find_neighbors_temp_bytes(num_leaf_nodes), num_leaf_nodes = OctreeNsView::numLeafNodes. */
uint64_t cstone_b200_find_neighbors_temp_bytes(int32_t num_leaf_nodes);
int cstone_b200_find_neighbors_ws_f32(const float* x, const float* y, const float* z, const float* h,
                                      uint32_t first_id, uint32_t last_id, const cstone_box_f32* box,
                                      const int32_t* child_offsets, const int32_t* internal_to_leaf,
                                      const uint32_t* layout, const float* centers, const float* sizes,
                                      float search_ext_factor, uint32_t ngmax, uint32_t* neighbors,
                                      uint32_t* neighbors_count, int32_t num_leaf_nodes, void* temp, void* stream);
int cstone_b200_find_neighbors_ws_f64(const double* x, const double* y, const double* z, const double* h,
                                      uint32_t first_id, uint32_t last_id, const cstone_box_f64* box,
                                      const int32_t* child_offsets, const int32_t* internal_to_leaf,
                                      const uint32_t* layout, const double* centers, const double* sizes,
                                      float search_ext_factor, uint32_t ngmax, uint32_t* neighbors,
                                      uint32_t* neighbors_count, int32_t num_leaf_nodes, void* temp, void* stream);

/* ------------------------------------------------------------------------
 * small primitives used by the callers of the hot path between its stages
 * [ref: primitives/primitives_gpu.h:39-125; tree/csarray_gpu.h:86-89;
 *       halos/gather_halos_gpu.h:21-27]
 * Results returned through a HOST pointer synchronise the stream (the
 * reference returns them by value).  temp >= cstone_b200_reduce_temp_bytes().
 * ---------------------------------------------------------------------- */
uint64_t cstone_b200_reduce_temp_bytes(void);
/* fillGpu(first, first+n, value): elem_bytes in {1,4,8}, value_host points to one element */
int cstone_b200_fill(void* first, size_t n, const void* value_host, int elem_bytes, void* stream);
/* scaleGpu: first[i] *= value */
int cstone_b200_scale_f32(float* first, size_t n, float value, void* stream);
int cstone_b200_scale_f64(double* first, size_t n, double value, void* stream);
/* incrementGpu: out[i] = in[i] + value (wrapping) */
int cstone_b200_increment_u32(const uint32_t* in, uint32_t* out, size_t n, uint32_t value, void* stream);
int cstone_b200_increment_u64(const uint64_t* in, uint64_t* out, size_t n, uint64_t value, void* stream);
/* maxNormSquareGpu: max_i (x*x + y*y + z*z), 0 for n = 0 */
int cstone_b200_max_norm_square_f32(const float* x, const float* y, const float* z, size_t n, float* out_host,
                                    void* temp, void* stream);
int cstone_b200_max_norm_square_f64(const double* x, const double* y, const double* z, size_t n, double* out_host,
                                    void* temp, void* stream);
/* reduceGpu<unsigned,size_t>: init + sum of in[0..n) in 64 bits */
int cstone_b200_reduce_sum_u32(const uint32_t* in, size_t n, uint64_t init, uint64_t* out_host, void* temp,
                               void* stream);
/* countGpu: number of elements equal to value (int arrays: pass the bit pattern to _u32) */
int cstone_b200_count_u32(const uint32_t* first, size_t n, uint32_t value, uint64_t* out_host, void* temp,
                          void* stream);
int cstone_b200_count_u64(const uint64_t* first, size_t n, uint64_t value, uint64_t* out_host, void* temp,
                          void* stream);
/* lowerBoundGpu(first, last, value) -> index of the first element >= value of a sorted array */
int cstone_b200_lower_bound_u32(const uint32_t* first, size_t n, uint32_t value, uint64_t* out_host, void* temp,
                                void* stream);
int cstone_b200_lower_bound_u64(const uint64_t* first, size_t n, uint64_t value, uint64_t* out_host, void* temp,
                                void* stream);
int cstone_b200_lower_bound_i32(const int32_t* first, size_t n, int32_t value, uint64_t* out_host, void* temp,
                                void* stream);
int cstone_b200_lower_bound_i64(const int64_t* first, size_t n, int64_t value, uint64_t* out_host, void* temp,
                                void* stream);
int cstone_b200_lower_bound_f32(const float* first, size_t n, float value, uint64_t* out_host, void* temp,
                                void* stream);
/* lowerBoundGpu(first, last, valueFirst, valueLast, result): one search per value, result index type u32 / u64
 * [used by createSendRangesGpu, domain/domaindecomp_gpu.cuh:55-69] */
int cstone_b200_lower_bound_many_k32_u32(const uint32_t* first, size_t n, const uint32_t* values, size_t num_values,
                                         uint32_t* result, void* stream);
int cstone_b200_lower_bound_many_k64_u32(const uint64_t* first, size_t n, const uint64_t* values, size_t num_values,
                                         uint32_t* result, void* stream);
int cstone_b200_lower_bound_many_k32_u64(const uint32_t* first, size_t n, const uint32_t* values, size_t num_values,
                                         uint64_t* result, void* stream);
int cstone_b200_lower_bound_many_k64_u64(const uint64_t* first, size_t n, const uint64_t* values, size_t num_values,
                                         uint64_t* result, void* stream);
/* segmentMax<Tin,Tout,IndexType>: out[s] = Tout(max(0, max in[segments[s] .. segments[s+1])));
 * index_bytes = sizeof(IndexType) in {4,8}; num_elements = segments[num_segments] (a hint that only selects the
 * thread mapping, the result does not depend on it) */
int cstone_b200_segment_max_f32_f32(const float* in, const void* segments, int index_bytes, size_t num_segments,
                                    size_t num_elements, float* out, void* stream);
int cstone_b200_segment_max_f64_f32(const double* in, const void* segments, int index_bytes, size_t num_segments,
                                    size_t num_elements, float* out, void* stream);
int cstone_b200_segment_max_f64_f64(const double* in, const void* segments, int index_bytes, size_t num_segments,
                                    size_t num_elements, double* out, void* stream);
/* sortGpu(first, last, keyBuf): ascending key-only sort (onesweep with a scratch permutation in temp) */
uint64_t cstone_b200_sort_keys_temp_bytes_k32(uint64_t n);
uint64_t cstone_b200_sort_keys_temp_bytes_k64(uint64_t n);
int cstone_b200_sort_keys_k32(uint32_t* keys, uint32_t* key_buf, size_t n, void* temp, uint64_t temp_bytes,
                              void* stream);
int cstone_b200_sort_keys_k64(uint64_t* keys, uint64_t* key_buf, size_t n, void* temp, uint64_t temp_bytes,
                              void* stream);
/* countSfcGapsGpu / fillSfcGapsGpu: node_ops[i] = number of cornerstone nodes needed to span [tree[i], tree[i+1])
 * [ref: sfc/common.hpp:392-438 spanSfcRange]; after an exclusive scan of node_ops (num_nodes+1 entries),
 * fill writes those nodes to new_tree + node_ops[i] and new_tree[node_ops[num_nodes]] = tree[num_nodes] */
int cstone_b200_count_sfc_gaps_k32(const uint32_t* tree, int32_t num_nodes, int32_t* node_ops, void* stream);
int cstone_b200_count_sfc_gaps_k64(const uint64_t* tree, int32_t num_nodes, int32_t* node_ops, void* stream);
int cstone_b200_fill_sfc_gaps_k32(const uint32_t* tree, int32_t num_nodes, const int32_t* node_ops,
                                  uint32_t* new_tree, void* stream);
int cstone_b200_fill_sfc_gaps_k64(const uint64_t* tree, int32_t num_nodes, const int32_t* node_ops,
                                  uint64_t* new_tree, void* stream);
/* gatherRanges<T,IndexType>: buffer[t] = src[range_offsets[r] + t - range_scan[r]] for the range r with
 * range_scan[r] <= t < range_scan[r+1] (range_scan has num_ranges entries, the last range ends at buffer_size);
 * index_bytes in {4,8}, elem_bytes in {4,8,12,16}  [halo send-buffer packing, halos/exchange_halos_gpu.cuh:71] */
int cstone_b200_gather_ranges(const void* range_scan, const void* range_offsets, int index_bytes, int32_t num_ranges,
                              const void* src, void* buffer, size_t buffer_size, int elem_bytes, void* stream);

/* ------------------------------------------------------------------------
 * a19 (device kernels of the focus-tree update)
 * [ref: focus/rebalance_gpu.h:39-85, traversal/collisions_gpu.h:69-77 markMacsGpu,
 *       focus/source_center_gpu.h:72-89 geoMacSpheresGpu / setMacGpu / moveCenters]
 * Tree arrays as in OctreeView (prefixes[M], child_offsets[M], parents[(M-1)/8]); macs / markings are char[M];
 * MAC spheres / source centres are Vec4<T>[M] = (x, y, z, mac^2).  temp >= 4 bytes of device scratch.
 * ---------------------------------------------------------------------- */
#define CSTONE_B200_FOCUS_DECL(KS, K)                                                                                  \
    /* rebalanceDecisionEssentialGpu: node_ops[i] in {0 merge, 1 keep, 8 split} for every node */                     \
    int cstone_b200_rebalance_decision_essential_##KS(const K* prefixes, const int32_t* child_offsets,                 \
                                                      const int32_t* parents, const uint32_t* counts,                  \
                                                      const char* macs, K focus_start, K focus_end,                    \
                                                      uint32_t bucket_size, int32_t* node_ops, int32_t num_nodes,      \
                                                      void* stream);                                                   \
    /* macRefineDecisionGpu: per LEAF, 8 if outside [focus_first, focus_last) and the MAC failed, else 1;           \
     * leaf_to_internal = the OctreeView span (build_octree's array + num_internal_nodes) */                          \
    int cstone_b200_mac_refine_decision_##KS(const K* prefixes, const char* macs, const int32_t* leaf_to_internal,     \
                                             int32_t num_leaf_nodes, int32_t focus_first, int32_t focus_last,          \
                                             int32_t* node_ops, void* stream);                                         \
    /* protectAncestorsGpu: *converged_host = 1 iff every resulting op is 1 */                                         \
    int cstone_b200_protect_ancestors_##KS(const K* prefixes, const int32_t* parents, int32_t* node_ops,               \
                                           int32_t num_nodes, int32_t* converged_host, void* temp, void* stream);      \
    /* enforceKeysGpu: *status_host = ResolutionStatus (0 converged, 1 cancelMerge, 2 rebalance, 3 failed) */          \
    int cstone_b200_enforce_keys_##KS(const K* forced_keys, int32_t num_forced_keys, const K* node_keys,               \
                                      const int32_t* child_offsets, const int32_t* parents, int32_t* node_ops,         \
                                      int32_t* status_host, void* temp, void* stream);                                 \
    /* rangeCountGpu: counts_focus[idx] = saturated sum of the global counts inside focus leaf idx, for idx in         \
     * leaves_focus_idx; num_leaf_keys = number of keys of the global leaf array (nodes + 1) */                         \
    int cstone_b200_range_count_##KS(const K* leaves, int32_t num_leaf_keys, const uint32_t* counts,                   \
                                     const K* leaves_focus, const int32_t* leaves_focus_idx, int32_t num_idx,          \
                                     uint32_t* counts_focus, void* stream);
CSTONE_B200_FOCUS_DECL(k32, uint32_t)
CSTONE_B200_FOCUS_DECL(k64, uint64_t)

#define CSTONE_B200_MAC_DECL(KS, K, TS, T)                                                                             \
    /* geoMacSpheresGpu: centres4[i] = (geometric centre, (2 max(size) inv_theta)^2) */                                \
    int cstone_b200_geo_mac_spheres_##KS##_##TS(int sfc_kind, const K* prefixes, int32_t num_nodes, T* centers4,       \
                                                float inv_theta, const cstone_box_##TS* box, void* stream);            \
    /* setMacGpu: mac_spheres4[i][3] = (2 max(size) inv_theta + |centre - geometric centre|)^2 unless it was 0 */      \
    int cstone_b200_set_mac_##KS##_##TS(int sfc_kind, const K* prefixes, int32_t num_nodes, T* mac_spheres4,           \
                                        float inv_theta, const cstone_box_##TS* box, void* stream);                    \
    /* markMacsGpu: markings[i] = 1 for every node outside the focus that fails the MAC w.r.t. one of the              \
     * focus leaves focus_nodes[0..num_focus_nodes] (cornerstone keys, num_focus_nodes + 1 entries) */                 \
    int cstone_b200_mark_macs_##KS##_##TS(int sfc_kind, const K* prefixes, const int32_t* child_offsets,               \
                                          const T* centers4, const cstone_box_##TS* box, const K* focus_nodes,         \
                                          int32_t num_focus_nodes, int limit_source, char* markings, void* stream);
CSTONE_B200_MAC_DECL(k32, uint32_t, f32, float)
CSTONE_B200_MAC_DECL(k32, uint32_t, f64, double)
CSTONE_B200_MAC_DECL(k64, uint64_t, f32, float)
CSTONE_B200_MAC_DECL(k64, uint64_t, f64, double)
/* moveCenters: dst4[i] = (src3[i], 1) */
int cstone_b200_move_centers_f32(const float* src3, int32_t num_nodes, float* dst4, void* stream);
int cstone_b200_move_centers_f64(const double* src3, int32_t num_nodes, double* dst4, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Expansion (mass) centres [ref: focus/source_center_gpu.h:51-74 computeLeafSourceCenterGpu, upsweepCentersGpu;
 * CPU: focus/source_center.hpp:44-126].  centers4 = num_nodes x (x, y, z, mass); suffixes = coordinate type, mass
 * type, centre type — the reference's three instantiations.  leaf_to_internal = the OctreeView span (L entries);
 * layout[i] = index of the first particle of leaf i (L + 1 entries).  Sums run in the reference's order (particle
 * by particle inside a leaf, child 0..7 inside a node): results are bit-identical to the CPU functions.
 * upsweep: level_range is a HOST pointer (num_levels + 1 entries), as in the reference. */
int cstone_b200_compute_leaf_source_center_f64_f64_f64(const double* x, const double* y, const double* z,
                                                       const double* m, const int32_t* leaf_to_internal,
                                                       int32_t num_leaves, const uint32_t* layout, double* centers4,
                                                       void* stream);
int cstone_b200_compute_leaf_source_center_f64_f32_f64(const double* x, const double* y, const double* z,
                                                       const float* m, const int32_t* leaf_to_internal,
                                                       int32_t num_leaves, const uint32_t* layout, double* centers4,
                                                       void* stream);
int cstone_b200_compute_leaf_source_center_f32_f32_f32(const float* x, const float* y, const float* z, const float* m,
                                                       const int32_t* leaf_to_internal, int32_t num_leaves,
                                                       const uint32_t* layout, float* centers4, void* stream);
int cstone_b200_upsweep_centers_f32(int32_t num_levels, const int32_t* level_range_host, const int32_t* child_offsets,
                                    float* centers4, void* stream);
int cstone_b200_upsweep_centers_f64(int32_t num_levels, const int32_t* level_range_host, const int32_t* child_offsets,
                                    double* centers4, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Target groups [ref: traversal/groups_gpu.h:60-88 computeFixedGroups, computeGroupSplits; GroupView
 * traversal/groups.hpp:19-27].  groups = group start indices, num_groups + 1 entries, groups[num_groups] = last
 * (GroupView::groupStart = groups, groupEnd = groups + 1).
 * fixed groups: num_groups = ceil((last-first)/group_size).
 * group splits: group_size 32 or 64 (else CSTONE_E_INVALID_ARGUMENT, the reference throws); suffixes = coordinate
 * type, smoothing-length type; h may be NULL (unused by the split criterion, as in the reference).  *num_groups_host
 * receives the number of groups; CSTONE_E_CAPACITY (with *num_groups_host set) if capacity < num_groups + 1 —
 * (last-first) + 1 entries always suffice.  temp >= cstone_b200_group_splits_temp_bytes(last-first, group_size).
 * The reference has no CPU twin; behaviour is pinned on test/unit_cuda/traversal/groups.cu. */
int cstone_b200_compute_fixed_groups(uint32_t first, uint32_t last, uint32_t group_size, uint32_t* groups,
                                     void* stream);
uint64_t cstone_b200_group_splits_temp_bytes(uint32_t num_particles, uint32_t group_size);
#define CSTONE_B200_GROUPS_DECL(KS, K, TS, TC, TH)                                                                     \
    int cstone_b200_compute_group_splits_##KS##_##TS(                                                                  \
        uint32_t first, uint32_t last, const TC* x, const TC* y, const TC* z, const TH* h, const K* leaves,            \
        int32_t num_leaves, const uint32_t* layout, const cstone_box_##TC##_t* box, uint32_t group_size,               \
        float tol_factor, uint32_t* groups, uint64_t capacity, uint32_t* num_groups_host, void* temp,                  \
        uint64_t temp_bytes, void* stream);
typedef cstone_box_f64 cstone_box_double_t;
typedef cstone_box_f32 cstone_box_float_t;
CSTONE_B200_GROUPS_DECL(k32, uint32_t, f64_f64, double, double)
CSTONE_B200_GROUPS_DECL(k32, uint32_t, f64_f32, double, float)
CSTONE_B200_GROUPS_DECL(k32, uint32_t, f32_f32, float, float)
CSTONE_B200_GROUPS_DECL(k64, uint64_t, f64_f64, double, double)
CSTONE_B200_GROUPS_DECL(k64, uint64_t, f64_f32, double, float)
CSTONE_B200_GROUPS_DECL(k64, uint64_t, f32_f32, float, float)

/* ------------------------------------------------------------------------------------------------------------
 * Single-rank focus tree, device resident [ref: focus/octree_focus.hpp:349-427 FocusedOctreeSingleNode<KeyType>;
 * CombinedUpdate::updateFocus :83-135].  create: bucket_size = maximum particle count of a leaf inside the focus,
 * theta = opening angle outside (NULL on invalid arguments).  update = one step of the reference's update():
 * rebalance by counts inside [focus_start, focus_end) and by MAC outside, focus boundaries and mandatory keys (HOST
 * array, may be unsorted / contain duplicates) enforced or injected, then MAC marks and particle counts of the new
 * tree from the sorted particle keys (DEVICE).  *converged_host = 1 iff the tree structure did not change.
 * view: device pointers owned by the tree, valid until the next update / destroy; leaf_to_internal is the full
 * array of build_octree (the OctreeView span starts at + num_internal_nodes). */
typedef struct cstone_b200_focus_tree_view
{
    int32_t num_leaf_nodes, num_internal_nodes, num_nodes, reserved;
    const void* leaves;          /* K[num_leaf_nodes + 1] */
    const uint32_t* leaf_counts; /* [num_leaf_nodes], NULL before the first update */
    const void* prefixes;        /* K[num_nodes] */
    const int32_t* child_offsets;
    const int32_t* parents;
    const int32_t* level_range; /* device copy, max_tree_level + 2 entries */
    const int32_t* internal_to_leaf;
    const int32_t* leaf_to_internal;
    const uint32_t* counts; /* [num_nodes] */
    const char* macs;       /* [num_nodes] */
} cstone_b200_focus_tree_view;
#define CSTONE_B200_FOCUS_TREE_DECL(KS, K)                                                                             \
    void* cstone_b200_focus_tree_create_##KS(int sfc_kind, uint32_t bucket_size, float theta);                         \
    void cstone_b200_focus_tree_destroy_##KS(void* handle);                                                            \
    int cstone_b200_focus_tree_update_##KS##_f32(void* handle, const cstone_box_f32* box, const K* particle_keys,      \
                                                 uint64_t num_keys, K focus_start, K focus_end,                        \
                                                 const K* mandatory_keys_host, int32_t num_mandatory,                  \
                                                 int32_t* converged_host, void* stream);                               \
    int cstone_b200_focus_tree_update_##KS##_f64(void* handle, const cstone_box_f64* box, const K* particle_keys,      \
                                                 uint64_t num_keys, K focus_start, K focus_end,                        \
                                                 const K* mandatory_keys_host, int32_t num_mandatory,                  \
                                                 int32_t* converged_host, void* stream);                               \
    /* replace the leaf array (device, num_leaf_nodes + 1 keys) and relink; counts and macs are zeroed */             \
    int cstone_b200_focus_tree_set_leaves_##KS(void* handle, const K* leaves, int32_t num_leaf_nodes, void* stream);   \
    /* one step of macRefineGpu [ref: focus/octree_focus.hpp:297-342]: the focus moved from [old_focus_start,          \
     * old_focus_end) to [focus_start, focus_end); leaves outside the focus failing the min-distance MAC w.r.t. the    \
     * newly covered focus leaves are split by one level.  *converged_host = 1 iff nothing was split. */               \
    int cstone_b200_focus_tree_mac_refine_##KS##_f32(void* handle, const cstone_box_f32* box, K old_focus_start,       \
                                                     K old_focus_end, K focus_start, K focus_end, float inv_theta,     \
                                                     int32_t* converged_host, void* stream);                           \
    int cstone_b200_focus_tree_mac_refine_##KS##_f64(void* handle, const cstone_box_f64* box, K old_focus_start,       \
                                                     K old_focus_end, K focus_start, K focus_end, float inv_theta,     \
                                                     int32_t* converged_host, void* stream);                           \
    int cstone_b200_focus_tree_view_##KS(void* handle, cstone_b200_focus_tree_view* out);
CSTONE_B200_FOCUS_TREE_DECL(k32, uint32_t)
CSTONE_B200_FOCUS_TREE_DECL(k64, uint64_t)

#ifdef __cplusplus
}
#endif
#endif /* CSTONE_B200_H */
