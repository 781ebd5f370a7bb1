/*
 * cstone_b200_domain.h — C ABI of the domain engine of libcstone_b200.so
 *
 * The engine is what stands behind cstone::Domain<KeyType,T,GpuTag>
 * [ref: domain/domain.hpp:66-699].  The C++ facade include/cstone_b200/domain.hpp
 * keeps the reference's template signatures (sync / exchangeHalos / startIndex /
 * endIndex / nParticles / nParticlesWithHalos / box) and forwards to the functions
 * below; all particle data stay on the device.
 *
 * sync is split in two calls because the CALLER owns and resizes the particle arrays:
 *   n_new = ..._sync_begin(d, n_in, x, y, z, h, nprops, props, prop_bytes)   // may communicate
 *   (caller resizes keys, x, y, z, h and every property array to n_new elements; contents may be discarded)
 *   ..._sync_finish(d, keys, x, y, z, h, props)                              // writes the SFC-ordered arrays
 * Postconditions are the reference's [ref: domain/domain.hpp:144-179]: x,y,z,h,keys hold assigned particles in
 * [start_index, end_index) and halos elsewhere, everything in SFC order; property arrays are valid in the assigned
 * range and are completed by ..._exchange_halos.
 *
 * Multi-GPU: one process per GPU.  The caller creates the NCCL communicator (e.g. unique id broadcast through
 * torch.distributed or MPI) and hands it over with ..._set_comm; with nranks == 1 no communicator is needed.
 * All functions return 0, a positive cudaError_t, a negative CSTONE_E_* code, or CSTONE_E_NCCL.
 */
#ifndef CSTONE_B200_DOMAIN_H
#define CSTONE_B200_DOMAIN_H

#include "cstone_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

#define CSTONE_E_NCCL (-6)
#define CSTONE_E_LET (-8) /* reference halo mode: peers disagree on node counts / a halo leaf belongs to no peer [ref: halos/halos.hpp:58-100] */
#define CSTONE_E_PEER_TIMEOUT (-9) /* pack-and-push halo exchange: a peer's data did not arrive within 20 s */
#define CSTONE_E_SIZE_MISMATCH (-7) /* array length differs from the previous sync [ref: domain/domain.hpp:130-133] */

typedef struct cstone_b200_domain cstone_b200_domain;

/* NCCL bootstrap helpers (thin wrappers so that a host language without NCCL bindings can drive them) */
int cstone_b200_nccl_unique_id(void* id128 /* 128 bytes, HOST */);
int cstone_b200_nccl_comm_init(const void* id128, int nranks, int rank, void** comm_out);
int cstone_b200_nccl_comm_destroy(void* comm);

/* In-process transport for single-GPU checks of the multi-rank path (no counterpart in the reference, whose ranks are
 * MPI processes): P ranks = P engine instances of ONE process, each driven by its own host thread on the same GPU;
 * messages are device-to-device copies, collectives meet at a host barrier.  Create one world, hand it to all P
 * engines with ..._set_local_world instead of ..._set_comm, then call sync_begin / exchange_halos of the P engines
 * concurrently from P threads.  Host-synchronous: a checking tool, not a fast path. */
void* cstone_b200_local_world_create(int nranks);
void cstone_b200_local_world_destroy(void* world);

#define CSB_DOMAIN_DECL(KS, K, TS, T)                                                                                  \
    /* [ref: domain/domain.hpp:95-113 Domain(rank, nRanks, bucketSize, bucketSizeFocus, theta, box)] */               \
    cstone_b200_domain* cstone_b200_domain_create_##KS##_##TS(int rank, int nranks, unsigned bucket_size,              \
                                                              unsigned bucket_size_focus, float theta,                 \
                                                              const cstone_box_##TS* box);                             \
    void cstone_b200_domain_destroy_##KS##_##TS(cstone_b200_domain* d);                                                \
    int cstone_b200_domain_set_comm_##KS##_##TS(cstone_b200_domain* d, void* nccl_comm);                               \
    int cstone_b200_domain_set_local_world_##KS##_##TS(cstone_b200_domain* d, void* local_world);                      \
    int cstone_b200_domain_set_fields_ready_event_##KS##_##TS(cstone_b200_domain* d, void* cuda_event);                \
    /* CSTONE_SFC_HILBERT (default, what the reference's Domain hard-wires: sfc/sfc.hpp:54-55) or CSTONE_SFC_MORTON;  \
     * call before the first sync */                                                                                   \
    int cstone_b200_domain_set_sfc_kind_##KS##_##TS(cstone_b200_domain* d, int sfc_kind);                              \
    /* [ref: domain/domain.hpp:396 setEndIndex] particles the caller appended behind the assigned range            \
     * (start <= end <= n_total) take part in the next sync */                                                         \
    int cstone_b200_domain_set_end_index_##KS##_##TS(cstone_b200_domain* d, uint64_t end);                             \
    /* [ref: domain/domain.hpp:412 setHaloFactor; halos/halos.hpp:162] halo radius = 2 * factor * max h per leaf */   \
    int cstone_b200_domain_set_halo_factor_##KS##_##TS(cstone_b200_domain* d, float factor);                           \
    /* transport of the halo exchange on nranks > 1.  1, the default: pack-and-push -- one kernel gathers the outgoing  \
     * elements of all arrays and stores them into the receivers' windows over NVLink (CUDA IPC peer memory), one      \
     * kernel per receiver waits for the senders' flags and fills the halo ranges; 0: one gather per array + a grouped \
     * ncclSend/ncclRecv.  What it replaces: halos/exchange_halos_gpu.cuh:34-119 haloExchangeGpu.  The environment    \
     * variable CSB_HALO_PUSH=0 selects 0 for engines that were not told otherwise */                                   \
    int cstone_b200_domain_set_halo_push_##KS##_##TS(cstone_b200_domain* d, int on);                                   \
    /* how halos are found on nranks > 1 (before the first sync).  0, the default: the fast path -- surface boxes of  \
     * all ranks against the leaves of the assigned particles; the halo set is sufficient for 2h searches and a       \
     * subset of the reference's.  1: the reference's own procedure [ref: domain/domain.hpp:217-243] -- peers by      \
     * findPeersMac, the multi-rank FocusedOctree (focus/octree_focus_mpi.hpp:108-273), Halos::discover and           \
     * computeLayout (halos/halos.hpp:127-223) -- halo flags, layout(), startIndex/endIndex, nParticlesWithHalos and   \
     * the tree of octreeProperties() are then the reference's, bit for bit */                                         \
    int cstone_b200_domain_set_halo_mode_##KS##_##TS(cstone_b200_domain* d, int mode);                                 \
    /* halo mode 1, after a sync: HOST views of the focus tree of this rank, valid until the next sync: leaves        \
     * K[L+1], leaf counts u32[L], halo flags i32[L], layout u32[L+1]; assignment2P receives the leaf index range      \
     * [start, end) of every rank in this tree (0,0 for non-peers) [ref: focus/octree_focus_mpi.hpp:560             \
     * assignment()], peers (capacity nranks) the peer ranks */                                                       \
    int cstone_b200_domain_let_view_##KS##_##TS(const cstone_b200_domain* d, int32_t* num_leaves, const K** leaves,    \
                                                const uint32_t** counts, const int32_t** flags,                        \
                                                const uint32_t** layout, int32_t* assignment2P, int32_t* peers,        \
                                                int32_t* num_peers);                                                   \
    /* [ref: domain/domain.hpp:404-406 startCell / endCell] first and one-past-last search-tree leaf of the assigned  \
     * SFC range; start_end2 is a HOST array; synchronises the stream */                                               \
    int cstone_b200_domain_cells_##KS##_##TS(cstone_b200_domain* d, int32_t* start_end2);                              \
    /* [ref: domain/domain.hpp:196-243 sync]; props: HOST array of nprops device pointers, element sizes 4, 8, 12 or  \
     * 16 bytes [ref: domain/buffer_description.hpp:98-110] */                                                          \
    int64_t cstone_b200_domain_sync_begin_##KS##_##TS(cstone_b200_domain* d, uint64_t n_in, const T* x, const T* y,    \
                                                      const T* z, const T* h, int nprops, const void* const* props,    \
                                                      const int* prop_bytes);                                          \
    /* [ref: domain/domain.hpp:245-325 syncGrav] sync for N-body clients: the focus tree is refined by the vector MAC   \
     * around the expansion (mass) centres of its nodes, which are made consistent over the ranks (own leaves from the  \
     * particles, cells above the local domains through the replicated global tree, peers' cells from their owners)    \
     * [ref: focus/octree_focus_mpi.hpp:368-525 updateCenters / updateMacs]; leaves that fail the MAC become halos.   \
     * props[0] MUST be the masses (element type T); halo mode 1 and nranks > 1 are required, otherwise               \
     * CSTONE_E_INVALID_ARGUMENT.  Finish with sync_finish as after sync_begin */                                      \
    int64_t cstone_b200_domain_sync_grav_begin_##KS##_##TS(cstone_b200_domain* d, uint64_t n_in, const T* x, const T* y,    \
                                                      const T* z, const T* h, int nprops, const void* const* props,    \
                                                      const int* prop_bytes);                                          \
    /* globalFocusExchange for node quantities of a client (multipoles, any per-cell data of 4 - 32 bytes), halo mode 1,   \
     * after a sync [ref: focus/octree_focus_mpi.hpp:276-300 peerExchange, :288-366 populateGlobal / gatherGlobalLeaves /  \
     * extractGlobal, :762-784 globalFocusExchange].  node_q: DEVICE array, one element per node of the focus tree      \
     * (order of tree_view's prefixes), valid for the cells inside this rank's SFC range.                               \
     *   focus_peer_exchange   cells in a peer's range take the owner's values                                          \
     *   focus_global_gather   global_leaf_q (DEVICE, one element per leaf of the global tree): this rank's leaves from  \
     *                         node_q, all others from their owners                                                     \
     *   global_octree         the global tree linked on the device, for the CLIENT's upsweep over the gathered leaves:  \
     *                         ptrs6 = prefixes K[M], child_offsets i32[M+1], parents, level_range (device),            \
     *                         internal_to_leaf, leaf_to_internal i32[M]; level_range_host i32[maxLevel+2]               \
     *   focus_global_extract  cells of the focus tree that belong to no peer read global_node_q (DEVICE, one element    \
     *                         per node of the global octree)                                                            \
     * All ranks call them together (the first two communicate).  CSTONE_E_LET if a cell has no counterpart. */          \
    int cstone_b200_domain_focus_peer_exchange_##KS##_##TS(cstone_b200_domain* d, void* node_q, int elem_bytes);       \
    int cstone_b200_domain_focus_global_gather_##KS##_##TS(cstone_b200_domain* d, const void* node_q, int elem_bytes,  \
                                                           void* global_leaf_q);                                       \
    int cstone_b200_domain_focus_global_extract_##KS##_##TS(cstone_b200_domain* d, const void* global_node_q,          \
                                                            int elem_bytes, void* node_q);                             \
    int cstone_b200_domain_global_octree_##KS##_##TS(const cstone_b200_domain* d, int32_t* num_leaves,                 \
                                                     int32_t* num_nodes, const void** ptrs6,                           \
                                                     int32_t* level_range_host);                                       \
    /* after sync_grav_begin: DEVICE views of the expansion centres (x, y, z, MAC radius^2) T[4 M] and the MAC marks   \
     * char[M] of the focus tree, M = *num_nodes nodes in the order of tree_view's prefixes */                         \
    int cstone_b200_domain_let_centers_##KS##_##TS(const cstone_b200_domain* d, const T** centers4, const char** macs, \
                                                   int32_t* num_nodes);                                                \
    int cstone_b200_domain_sync_finish_##KS##_##TS(cstone_b200_domain* d, K* keys, T* x, T* y, T* z, T* h,             \
                                                   void* const* props);                                                \
    /* [ref: domain/domain.hpp:381-386 exchangeHalos] arrays: HOST array of device pointers of length n_new */        \
    int cstone_b200_domain_exchange_halos_##KS##_##TS(cstone_b200_domain* d, int narrays, void* const* arrays,         \
                                                      const int* elem_bytes);                                          \
    /* [ref: domain/domain.hpp:334-378 reapplySync] the particle exchange and re-ordering of the last sync applied to  \
     * further arrays (device arrays here; the reference supports host arrays only).  in: HOST array of device        \
     * pointers to arrays in the layout the caller's arrays had BEFORE that sync; out: HOST array of device pointers   \
     * to arrays of the new size, whose assigned range [start, end) is written; element sizes 4, 8, 12 or 16 bytes.    \
     * Collective over the ranks of the domain. */                                                                     \
    int cstone_b200_domain_reapply_sync_##KS##_##TS(cstone_b200_domain* d, int narrays, const void* const* in,         \
                                                    const int* elem_bytes, void* const* out);                          \
    /* [ref: domain/domain.hpp:389-437] accessors; info[8] = start, end, n_total, n_leaves(focus), n_nodes(focus),    \
     * n_leaves(global), num halo send particles, num halo recv particles */                                           \
    int cstone_b200_domain_info_##KS##_##TS(const cstone_b200_domain* d, int64_t* info8);                              \
    int cstone_b200_domain_box_##KS##_##TS(const cstone_b200_domain* d, cstone_box_##TS* box_out);                     \
    /* device time (ms, CUDA events) of the 10 stages of the last sync_begin: box, sfc keys, sort, global tree +      \
     * assignment, particle exchange (pack gather + NCCL), re-key + re-sort of the assigned set, halo discovery,       \
     * final gather + halo exchange + halo keys, search-tree update, linking + centres + layout */                     \
    int cstone_b200_domain_stage_ms_##KS##_##TS(const cstone_b200_domain* d, float* ms10);                     \
    /* device views of the search tree over assigned + halo particles, valid until the next sync:                     \
     * ptrs[10] = leaves K[L+1], counts u32[L], layout u32[L+1], prefixes K[M], child_offsets i32[M+1],                \
     *            internal_to_leaf i32[M], leaf_to_internal i32[M], level_range i32[maxLevel+2], centers T[3M],        \
     *            sizes T[3M]   [ref: tree/octree.hpp:296-317 OctreeNsView, domain/domain.hpp:400-420] */              \
    int cstone_b200_domain_tree_view_##KS##_##TS(const cstone_b200_domain* d, const void** ptrs10);                    \
    /* global (replicated) tree on the HOST: leaves K[Lg+1], counts u32[Lg], rank boundaries K[nranks+1] */            \
    int cstone_b200_domain_global_tree_##KS##_##TS(const cstone_b200_domain* d, const K** leaves,                      \
                                                   const uint32_t** counts, const K** assignment);                     \
    /* findNeighbors over the assigned range with the domain's own tree [ref: findneighbors.hpp:167-178] */           \
    int cstone_b200_domain_find_neighbors_##KS##_##TS(const cstone_b200_domain* d, const T* x, const T* y,             \
                                                      const T* z, const T* h, uint32_t ngmax, uint32_t* neighbors,     \
                                                      uint32_t* neighbors_count);

CSB_DOMAIN_DECL(k64, uint64_t, f64, double)
CSB_DOMAIN_DECL(k64, uint64_t, f32, float)
CSB_DOMAIN_DECL(k32, uint32_t, f64, double)
CSB_DOMAIN_DECL(k32, uint32_t, f32, float)

/* ---- host-side decomposition logic, exported for CPU tests (no GPU needed) ------------------------------
 * [ref: domain/domaindecomp.hpp:48-71 uniformBins, :115-124 makeSfcAssignment, :139-166 limitBoundaryShifts]
 * leaves[L+1], counts[L] -> boundaries[nranks+1] (keys) and rank_counts[nranks];
 * old_boundaries may be NULL (first call). */
int cstone_b200_host_assignment_k64(int nranks, const uint64_t* leaves, const uint32_t* counts, int32_t num_leaves,
                                    const uint64_t* old_boundaries, uint64_t* boundaries, uint32_t* rank_counts);
int cstone_b200_host_assignment_k32(int nranks, const uint32_t* leaves, const uint32_t* counts, int32_t num_leaves,
                                    const uint32_t* old_boundaries, uint32_t* boundaries, uint32_t* rank_counts);
/* [ref: tree/csarray.hpp:477-488 updateTreelet, as called by focus/exchange_focus.hpp:380-405 focusTransfer] HOST
 * function: leaves [first, last) of a full cornerstone leaf array rebalanced once with their counts (sibling groups may
 * reach outside of the range, as in the reference); out receives the new keys incl. the upper boundary leaves[last].
 * Returns the number of keys written (> 0), CSTONE_E_CAPACITY if out is too small */
int cstone_b200_host_update_treelet_k32(const uint32_t* leaves, const uint32_t* counts, int32_t first, int32_t last,
                                        uint32_t bucket, uint32_t* out, int32_t capacity);
int cstone_b200_host_update_treelet_k64(const uint64_t* leaves, const uint32_t* counts, int32_t first, int32_t last,
                                        uint32_t bucket, uint64_t* out, int32_t capacity);
/* [ref: traversal/peers.hpp:62-117 findPeersMac] HOST function: ranks owning a global leaf that fails the mutual
 * minimum-distance / vector MAC (inv_theta_eff) against a global leaf of my_rank, by a dual traversal of the global
 * tree given as its leaf keys (num_leaves + 1 keys); boundaries: nranks + 1 keys; is_peer: nranks flags */
#define CSTONE_B200_HOST_PEERS_DECL(KS, K, TS)                                                                         \
    int cstone_b200_host_find_peers_##KS##_##TS(int sfc_kind, int my_rank, int nranks, const K* boundaries,            \
                                                const K* leaves, int32_t num_leaves, const cstone_box_##TS* box,       \
                                                float inv_theta_eff, int32_t* is_peer);
CSTONE_B200_HOST_PEERS_DECL(k32, uint32_t, f32)
CSTONE_B200_HOST_PEERS_DECL(k32, uint32_t, f64)
CSTONE_B200_HOST_PEERS_DECL(k64, uint64_t, f32)
CSTONE_B200_HOST_PEERS_DECL(k64, uint64_t, f64)
/* [ref: domain/domaindecomp.hpp:48-71 uniformBins] bins[num_bins+1] (leaf indices), bin_counts[num_bins] */
int cstone_b200_host_uniform_bins(const uint32_t* counts, int32_t num_leaves, int32_t num_bins, int32_t* bins,
                                  uint32_t* bin_counts);
/* [ref: domain/domaindecomp.hpp:139-166 limitBoundaryShifts] boundaries / rank_counts are updated in place */
int cstone_b200_host_limit_boundary_shifts_k64(int nranks, const uint64_t* old_boundaries, uint64_t* boundaries,
                                               const uint64_t* leaves, const uint32_t* counts, int32_t num_leaves,
                                               uint32_t* rank_counts);
int cstone_b200_host_limit_boundary_shifts_k32(int nranks, const uint32_t* old_boundaries, uint32_t* boundaries,
                                               const uint32_t* leaves, const uint32_t* counts, int32_t num_leaves,
                                               uint32_t* rank_counts);
/* [ref: sfc/box.hpp:415-431 limitBoxShrinking] limits are {xmin,xmax,ymin,ymax,zmin,zmax} */
void cstone_b200_host_limit_box_shrinking_f64(const double* fitting6, const double* previous6, double* out6);
void cstone_b200_host_limit_box_shrinking_f32(const float* fitting6, const float* previous6, float* out6);

#ifdef __cplusplus
}
#endif
#endif
