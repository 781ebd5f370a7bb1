/*
 * ref_wrap.cpp — TEST INFRASTRUCTURE ONLY.
 *
 * Thin extern "C" wrappers around the UNMODIFIED reference CPU headers
 * (/root/reference/include/cstone/...).  No reference source is copied: this file only
 * #includes the headers where they lie and forwards calls.  Built by oracle/Makefile
 * into oracle/_ref/liboracle_ref.so (git-ignored) with
 *   g++ -std=c++20 -O3 -fopenmp -ffp-contract=off -I/root/reference/include
 * It is the strongest checker we have: the reference's own code for every function of
 * the hot path, used to pin the port (oracle/cstone_oracle.cpp) and the CUDA kernels.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <utility>
#include <vector>
#include <omp.h>

#include "cstone/sfc/sfc.hpp"
#include "cstone/primitives/gather.hpp"
#include "cstone/tree/csarray.hpp"
#include "cstone/tree/octree.hpp"
#include "cstone/focus/source_center.hpp"
#include "cstone/traversal/collisions.hpp"
#include "cstone/domain/layout.hpp"
#include "cstone/findneighbors.hpp"
#include "cstone/domain/domaindecomp.hpp"

#define ORACLE_PREFIX cstone_ref_
#include "oracle_api.h"

using namespace cstone;

static_assert(sizeof(Box<float>) == sizeof(cstone_box_f32), "Box<float> layout");
static_assert(sizeof(Box<double>) == sizeof(cstone_box_f64), "Box<double> layout");

template<class T, class CB>
static const Box<T>& asBox(const CB* b)
{
    return *reinterpret_cast<const Box<T>*>(b);
}

[[noreturn]] static void unsupported(const char* what)
{
    std::fprintf(stderr, "cstone_ref: %s is hard-wired to the Hilbert curve in the reference\n", what);
    std::abort();
}

template<class K, class T, class CB>
static void sfcKeys(int kind, const T* x, const T* y, const T* z, K* keys, size_t n, const CB* box)
{
    if (kind == CSTONE_SFC_MORTON)
        computeSfcKeys(x, y, z, reinterpret_cast<MortonKey<K>*>(keys), n, asBox<T>(box));
    else
        computeSfcKeys(x, y, z, reinterpret_cast<HilbertKey<K>*>(keys), n, asBox<T>(box));
}

template<class K>
static K encode(int kind, unsigned ix, unsigned iy, unsigned iz)
{
    return kind == CSTONE_SFC_MORTON ? iMorton<K>(ix, iy, iz) : iHilbert<K>(ix, iy, iz);
}

template<class K>
static void decode(int kind, K key, unsigned* xyz)
{
    auto t = (kind == CSTONE_SFC_MORTON) ? decodeMorton<K>(key) : decodeHilbert<K>(key);
    xyz[0] = util::get<0>(t);
    xyz[1] = util::get<1>(t);
    xyz[2] = util::get<2>(t);
}

template<class K>
static void ibox(int kind, K key, unsigned level, int* out)
{
    IBox b = (kind == CSTONE_SFC_MORTON) ? mortonIBox<K>(key, level) : hilbertIBox<K>(key, level);
    out[0] = b.xmin(), out[1] = b.xmax(), out[2] = b.ymin(), out[3] = b.ymax(), out[4] = b.zmin(), out[5] = b.zmax();
}

template<class K>
static int32_t rebalanceTreeW(const K* tree, int32_t numNodes, int32_t* nodeOps, K* newTree, int32_t capacity)
{
    gsl::span<const K> treeSpan(tree, numNodes + 1);
    std::vector<K> out;
    rebalanceTree(treeSpan, out, nodeOps);
    int32_t newNumNodes = int32_t(out.size()) - 1;
    if (newNumNodes > capacity) return -newNumNodes;
    std::copy(out.begin(), out.end(), newTree);
    return newNumNodes;
}

template<class K>
static int updateOctreeW(const K* first, const K* last, unsigned bucket, unsigned maxCount, K* tree, unsigned* counts,
                         int32_t capacity, int32_t* numNodes)
{
    std::vector<K> t(tree, tree + *numNodes + 1);
    std::vector<unsigned> c(counts, counts + *numNodes);
    bool converged = updateOctree(first, last, bucket, t, c, maxCount);
    int32_t newN   = int32_t(nNodes(t));
    if (newN > capacity)
    {
        *numNodes = newN;
        return -1;
    }
    std::copy(t.begin(), t.end(), tree);
    std::copy(c.begin(), c.end(), counts);
    *numNodes = newN;
    return converged ? 1 : 0;
}

template<class K>
static int32_t computeOctreeW(const K* first, const K* last, unsigned bucket, unsigned maxCount, K* tree,
                              unsigned* counts, int32_t capacity)
{
    auto [t, c]  = computeOctree(first, last, bucket, maxCount);
    int32_t newN = int32_t(nNodes(t));
    if (newN > capacity) return -newN;
    std::copy(t.begin(), t.end(), tree);
    std::copy(c.begin(), c.end(), counts);
    return newN;
}

template<class K>
static void buildOctreeW(const K* leaves, int32_t L, K* prefixes, int32_t* childOffsets, int32_t* parents,
                         int32_t* levelRange, int32_t* internalToLeaf, int32_t* leafToInternal)
{
    buildOctreeCpu(leaves, L, (L - 1) / 7, prefixes, childOffsets, parents, levelRange, internalToLeaf, leafToInternal);
}

template<class K, class T, class CB>
static void geoCentersW(int kind, const K* prefixes, int32_t n, T* centers, T* sizes, const CB* box)
{
    if (kind != CSTONE_SFC_HILBERT) unsupported("nodeFpCenters");
    nodeFpCenters<K>(gsl::span<const K>(prefixes, n), reinterpret_cast<Vec3<T>*>(centers),
                     reinterpret_cast<Vec3<T>*>(sizes), asBox<T>(box));
}

//! halo radius as computed by Halos::discover, CPU branch (halos/halos.hpp:171-183), for arbitrary segments
template<class Th>
static void haloRadiiW(const Th* h, const uint32_t* seg, size_t numSegments, float scale, float* radii)
{
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < numSegments; ++i)
    {
        radii[i] = 0.0f;
        // reference GPU branch (segmentMax + scaleGpu): float(max) * scale; CPU branch: max * 2 * ext narrowed to float.
        // Both agree for scale = 2 (power of two); we follow the GPU order of operations.
        if (seg[i + 1] > seg[i]) { radii[i] = float(*std::max_element(h + seg[i], h + seg[i + 1])) * scale; }
    }
}

template<class K, class T, class CB>
static void findHalosW(int kind, const K* prefixes, const int32_t* childOffsets, const int32_t* internalToLeaf,
                       const K* leaves, const float* radii, const CB* box, int32_t first, int32_t last, int32_t* flags)
{
    if (kind != CSTONE_SFC_HILBERT) unsupported("findHalos");
    findHalos(prefixes, childOffsets, internalToLeaf, leaves, radii, asBox<T>(box), first, last, flags);
}

template<class T, class CB>
static void findNeighborsW(const T* x, const T* y, const T* z, const T* h, uint32_t firstId, uint32_t lastId,
                           const CB* box, const int32_t* childOffsets, const int32_t* internalToLeaf,
                           const uint32_t* layout, const T* centers, const T* sizes, float ext, uint32_t ngmax,
                           uint32_t* neighbors, uint32_t* neighborsCount)
{
    OctreeNsView<T, uint64_t> view{0,
                                   nullptr,
                                   childOffsets,
                                   internalToLeaf,
                                   nullptr,
                                   nullptr,
                                   layout,
                                   reinterpret_cast<const Vec3<T>*>(centers),
                                   reinterpret_cast<const Vec3<T>*>(sizes),
                                   ext};
    findNeighbors(x, y, z, h, firstId, lastId, asBox<T>(box), view, ngmax, neighbors, neighborsCount);
}

extern "C" {

int cstone_ref_num_threads(void) { return omp_get_max_threads(); }
void cstone_ref_set_num_threads(int n) { omp_set_num_threads(n); }

uint32_t cstone_ref_encode_k32(int kind, unsigned ix, unsigned iy, unsigned iz) { return encode<uint32_t>(kind, ix, iy, iz); }
uint64_t cstone_ref_encode_k64(int kind, unsigned ix, unsigned iy, unsigned iz) { return encode<uint64_t>(kind, ix, iy, iz); }
void cstone_ref_decode_k32(int kind, uint32_t key, unsigned* xyz) { decode<uint32_t>(kind, key, xyz); }
void cstone_ref_decode_k64(int kind, uint64_t key, unsigned* xyz) { decode<uint64_t>(kind, key, xyz); }

void cstone_ref_compute_sfc_keys_k32_f32(int kind, const float* x, const float* y, const float* z, uint32_t* keys,
                                         size_t n, const cstone_box_f32* box) { sfcKeys(kind, x, y, z, keys, n, box); }
void cstone_ref_compute_sfc_keys_k32_f64(int kind, const double* x, const double* y, const double* z, uint32_t* keys,
                                         size_t n, const cstone_box_f64* box) { sfcKeys(kind, x, y, z, keys, n, box); }
void cstone_ref_compute_sfc_keys_k64_f32(int kind, const float* x, const float* y, const float* z, uint64_t* keys,
                                         size_t n, const cstone_box_f32* box) { sfcKeys(kind, x, y, z, keys, n, box); }
void cstone_ref_compute_sfc_keys_k64_f64(int kind, const double* x, const double* y, const double* z, uint64_t* keys,
                                         size_t n, const cstone_box_f64* box) { sfcKeys(kind, x, y, z, keys, n, box); }
void cstone_ref_sfc_ibox_k32(int kind, uint32_t key, unsigned level, int* b) { ibox<uint32_t>(kind, key, level, b); }
void cstone_ref_sfc_ibox_k64(int kind, uint64_t key, unsigned level, int* b) { ibox<uint64_t>(kind, key, level, b); }

void cstone_ref_sort_by_key_k32(uint32_t* keys, uint32_t* values, size_t n) { sort_by_key(keys, keys + n, values); }
void cstone_ref_sort_by_key_k64(uint64_t* keys, uint32_t* values, size_t n) { sort_by_key(keys, keys + n, values); }

void cstone_ref_gather(const uint32_t* map, size_t n, const void* src, void* dst, int elemBytes)
{
    gsl::span<const uint32_t> m(map, n);
    switch (elemBytes)
    {
        case 4: gather<uint32_t>(m, (const uint32_t*)src, (uint32_t*)dst); break;
        case 8: gather<uint32_t>(m, (const uint64_t*)src, (uint64_t*)dst); break;
        case 12: gather<uint32_t>(m, (const util::array<uint32_t, 3>*)src, (util::array<uint32_t, 3>*)dst); break;
        case 16: gather<uint32_t>(m, (const util::array<uint64_t, 2>*)src, (util::array<uint64_t, 2>*)dst); break;
        case 24: gather<uint32_t>(m, (const util::array<uint64_t, 3>*)src, (util::array<uint64_t, 3>*)dst); break;
        case 32: gather<uint32_t>(m, (const util::array<uint64_t, 4>*)src, (util::array<uint64_t, 4>*)dst); break;
        default: std::abort();
    }
}

void cstone_ref_compute_node_counts_k32(const uint32_t* tree, uint32_t* counts, int32_t n, const uint32_t* f,
                                        const uint32_t* l, uint32_t maxCount, int guess)
{
    computeNodeCounts(tree, counts, n, f, l, maxCount, guess != 0);
}
void cstone_ref_compute_node_counts_k64(const uint64_t* tree, uint32_t* counts, int32_t n, const uint64_t* f,
                                        const uint64_t* l, uint32_t maxCount, int guess)
{
    computeNodeCounts(tree, counts, n, f, l, maxCount, guess != 0);
}
int cstone_ref_rebalance_decision_k32(const uint32_t* tree, const uint32_t* counts, int32_t n, uint32_t bucket,
                                      int32_t* ops) { return rebalanceDecision(tree, counts, n, bucket, ops); }
int cstone_ref_rebalance_decision_k64(const uint64_t* tree, const uint32_t* counts, int32_t n, uint32_t bucket,
                                      int32_t* ops) { return rebalanceDecision(tree, counts, n, bucket, ops); }
int32_t cstone_ref_rebalance_tree_k32(const uint32_t* tree, int32_t n, int32_t* ops, uint32_t* newTree, int32_t cap)
{ return rebalanceTreeW(tree, n, ops, newTree, cap); }
int32_t cstone_ref_rebalance_tree_k64(const uint64_t* tree, int32_t n, int32_t* ops, uint64_t* newTree, int32_t cap)
{ return rebalanceTreeW(tree, n, ops, newTree, cap); }
int cstone_ref_update_octree_k32(const uint32_t* f, const uint32_t* l, uint32_t bucket, uint32_t maxCount,
                                 uint32_t* tree, uint32_t* counts, int32_t cap, int32_t* n)
{ return updateOctreeW(f, l, bucket, maxCount, tree, counts, cap, n); }
int cstone_ref_update_octree_k64(const uint64_t* f, const uint64_t* l, uint32_t bucket, uint32_t maxCount,
                                 uint64_t* tree, uint32_t* counts, int32_t cap, int32_t* n)
{ return updateOctreeW(f, l, bucket, maxCount, tree, counts, cap, n); }
int32_t cstone_ref_compute_octree_k32(const uint32_t* f, const uint32_t* l, uint32_t bucket, uint32_t maxCount,
                                      uint32_t* tree, uint32_t* counts, int32_t cap)
{ return computeOctreeW(f, l, bucket, maxCount, tree, counts, cap); }
int32_t cstone_ref_compute_octree_k64(const uint64_t* f, const uint64_t* l, uint32_t bucket, uint32_t maxCount,
                                      uint64_t* tree, uint32_t* counts, int32_t cap)
{ return computeOctreeW(f, l, bucket, maxCount, tree, counts, cap); }

void cstone_ref_build_octree_k32(const uint32_t* leaves, int32_t L, uint32_t* p, int32_t* co, int32_t* pa, int32_t* lr,
                                 int32_t* i2l, int32_t* l2i) { buildOctreeW(leaves, L, p, co, pa, lr, i2l, l2i); }
void cstone_ref_build_octree_k64(const uint64_t* leaves, int32_t L, uint64_t* p, int32_t* co, int32_t* pa, int32_t* lr,
                                 int32_t* i2l, int32_t* l2i) { buildOctreeW(leaves, L, p, co, pa, lr, i2l, l2i); }
int cstone_ref_span_sfc_range_k32(uint32_t a, uint32_t b, uint32_t* out)
{ return out ? spanSfcRange(a, b, out) : spanSfcRange(a, b); }
int cstone_ref_span_sfc_range_k64(uint64_t a, uint64_t b, uint64_t* out)
{ return out ? spanSfcRange(a, b, out) : spanSfcRange(a, b); }
void cstone_ref_upsweep_sum(int32_t numLevels, const int32_t* levelRange, const int32_t* childOffsets, uint32_t* counts)
{
    // as used by FocusedOctree::updateCounts: upsweep(levelRange, childOffsets, counts, NodeCount<unsigned>{})
    upsweep(gsl::span<const int32_t>(levelRange, numLevels + 1), gsl::span<const int32_t>(childOffsets, size_t(levelRange[numLevels])),
            counts, NodeCount<unsigned>{});
}

void cstone_ref_geo_centers_k32_f32(int kind, const uint32_t* p, int32_t n, float* c, float* s, const cstone_box_f32* b)
{ geoCentersW(kind, p, n, c, s, b); }
void cstone_ref_geo_centers_k32_f64(int kind, const uint32_t* p, int32_t n, double* c, double* s, const cstone_box_f64* b)
{ geoCentersW(kind, p, n, c, s, b); }
void cstone_ref_geo_centers_k64_f32(int kind, const uint64_t* p, int32_t n, float* c, float* s, const cstone_box_f32* b)
{ geoCentersW(kind, p, n, c, s, b); }
void cstone_ref_geo_centers_k64_f64(int kind, const uint64_t* p, int32_t n, double* c, double* s, const cstone_box_f64* b)
{ geoCentersW(kind, p, n, c, s, b); }

void cstone_ref_halo_radii_f32(const float* h, const uint32_t* seg, size_t n, float scale, float* r) { haloRadiiW(h, seg, n, scale, r); }
void cstone_ref_halo_radii_f64(const double* h, const uint32_t* seg, size_t n, float scale, float* r) { haloRadiiW(h, seg, n, scale, r); }

void cstone_ref_find_halos_k32_f32(int kind, const uint32_t* p, const int32_t* co, const int32_t* i2l, const uint32_t* lv,
                                   const float* r, const cstone_box_f32* b, int32_t f, int32_t l, int32_t* flags)
{ findHalosW<uint32_t, float>(kind, p, co, i2l, lv, r, b, f, l, flags); }
void cstone_ref_find_halos_k32_f64(int kind, const uint32_t* p, const int32_t* co, const int32_t* i2l, const uint32_t* lv,
                                   const float* r, const cstone_box_f64* b, int32_t f, int32_t l, int32_t* flags)
{ findHalosW<uint32_t, double>(kind, p, co, i2l, lv, r, b, f, l, flags); }
void cstone_ref_find_halos_k64_f32(int kind, const uint64_t* p, const int32_t* co, const int32_t* i2l, const uint64_t* lv,
                                   const float* r, const cstone_box_f32* b, int32_t f, int32_t l, int32_t* flags)
{ findHalosW<uint64_t, float>(kind, p, co, i2l, lv, r, b, f, l, flags); }
void cstone_ref_find_halos_k64_f64(int kind, const uint64_t* p, const int32_t* co, const int32_t* i2l, const uint64_t* lv,
                                   const float* r, const cstone_box_f64* b, int32_t f, int32_t l, int32_t* flags)
{ findHalosW<uint64_t, double>(kind, p, co, i2l, lv, r, b, f, l, flags); }

void cstone_ref_compute_node_layout(const uint32_t* counts, const int32_t* flags, int32_t first, int32_t last,
                                    int32_t L, uint32_t* layout)
{
    computeNodeLayout(gsl::span<const unsigned>(counts, L), gsl::span<const int>(flags, L), first, last,
                      gsl::span<LocalIndex>(layout, L + 1));
}

void cstone_ref_find_neighbors_f32(const float* x, const float* y, const float* z, const float* h, uint32_t f, uint32_t l,
                                   const cstone_box_f32* box, const int32_t* co, const int32_t* i2l, const uint32_t* layout,
                                   const float* c, const float* s, float ext, uint32_t ngmax, uint32_t* nb, uint32_t* nc)
{ findNeighborsW(x, y, z, h, f, l, box, co, i2l, layout, c, s, ext, ngmax, nb, nc); }
void cstone_ref_find_neighbors_f64(const double* x, const double* y, const double* z, const double* h, uint32_t f, uint32_t l,
                                   const cstone_box_f64* box, const int32_t* co, const int32_t* i2l, const uint32_t* layout,
                                   const double* c, const double* s, float ext, uint32_t ngmax, uint32_t* nb, uint32_t* nc)
{ findNeighborsW(x, y, z, h, f, l, box, co, i2l, layout, c, s, ext, ngmax, nb, nc); }

} // extern "C"

template<class K>
static void assignmentW(int nranks, const K* leaves, const uint32_t* counts, int32_t L, const K* oldB, K* boundaries,
                        uint32_t* rankCounts)
{
    std::vector<unsigned> c(counts, counts + L);
    SfcAssignment<K> a = makeSfcAssignment(nranks, c, leaves);
    if (oldB)
    {
        SfcAssignment<K> old(nranks);
        for (int r = 0; r <= nranks; ++r)
            old.set(r, oldB[r], 0);
        limitBoundaryShifts<K>(old, a, gsl::span<const K>(leaves, L + 1), gsl::span<const unsigned>(counts, L));
    }
    for (int r = 0; r <= nranks; ++r)
        boundaries[r] = a[r];
    for (int r = 0; r < nranks; ++r)
        rankCounts[r] = a.totalCount(r);
}
template<class T>
static void limitBoxW(const T* f, const T* p, T* o)
{
    Box<T> fit(f[0], f[1], f[2], f[3], f[4], f[5]), prev(p[0], p[1], p[2], p[3], p[4], p[5]);
    Box<T> r = limitBoxShrinking(fit, prev);
    o[0] = r.xmin(), o[1] = r.xmax(), o[2] = r.ymin(), o[3] = r.ymax(), o[4] = r.zmin(), o[5] = r.zmax();
}

extern "C" {
void cstone_ref_assignment_k32(int n, const uint32_t* l, const uint32_t* c, int32_t L, const uint32_t* ob, uint32_t* b,
                               uint32_t* rc)
{ assignmentW(n, l, c, L, ob, b, rc); }
void cstone_ref_assignment_k64(int n, const uint64_t* l, const uint32_t* c, int32_t L, const uint64_t* ob, uint64_t* b,
                               uint32_t* rc)
{ assignmentW(n, l, c, L, ob, b, rc); }
void cstone_ref_limit_box_shrinking_f32(const float* f, const float* p, float* o) { limitBoxW(f, p, o); }
void cstone_ref_limit_box_shrinking_f64(const double* f, const double* p, double* o) { limitBoxW(f, p, o); }
}

template<class K>
static void limitShiftsW(int nranks, const K* oldB, K* boundaries, const K* leaves, const uint32_t* counts, int32_t L,
                         uint32_t* rankCounts)
{
    SfcAssignment<K> old(nranks), cur(nranks);
    for (int r = 0; r <= nranks; ++r)
    {
        old.set(r, oldB[r], 0);
        cur.set(r, boundaries[r], r < nranks ? rankCounts[r] : 0);
    }
    limitBoundaryShifts<K>(old, cur, gsl::span<const K>(leaves, L + 1), gsl::span<const unsigned>(counts, L));
    for (int r = 0; r <= nranks; ++r)
        boundaries[r] = cur[r];
    for (int r = 0; r < nranks; ++r)
        rankCounts[r] = cur.totalCount(r);
}

extern "C" {
void cstone_ref_uniform_bins(const uint32_t* counts, int32_t L, int32_t numBins, int32_t* bins, uint32_t* binCounts)
{
    std::vector<unsigned> c(counts, counts + L);
    uniformBins(c, gsl::span<TreeNodeIndex>(bins, numBins + 1), gsl::span<LocalIndex>(binCounts, numBins));
}
void cstone_ref_limit_boundary_shifts_k32(int n, const uint32_t* ob, uint32_t* b, const uint32_t* l, const uint32_t* c,
                                          int32_t L, uint32_t* rc)
{ limitShiftsW(n, ob, b, l, c, L, rc); }
void cstone_ref_limit_boundary_shifts_k64(int n, const uint64_t* ob, uint64_t* b, const uint64_t* l, const uint32_t* c,
                                          int32_t L, uint32_t* rc)
{ limitShiftsW(n, ob, b, l, c, L, rc); }
}

// ------------------------------------------------------------------------------------------------
// a19 device-kernel counterparts: the reference's CPU functions of the focus-tree update (reference build only)
// ------------------------------------------------------------------------------------------------
#include "cstone/focus/rebalance.hpp"
#include "cstone/traversal/macs.hpp"

template<class K>
static void rebalanceEssentialW(const K* p, const int32_t* co, const int32_t* pa, const uint32_t* c, const char* macs,
                                K fs, K fe, uint32_t bucket, int32_t* ops, int32_t n)
{
    rebalanceDecisionEssential<K>(gsl::span<const K>(p, size_t(n)), co, pa, c, macs, fs, fe, bucket, ops);
}
template<class K>
static void macRefineW(const K* p, const char* macs, const int32_t* l2i, int32_t L, int32_t f0, int32_t f1, int32_t* ops)
{
    // the loop of macRefineDecisionGpu [focus/rebalance_gpu.cu:93-104] over the CPU macRefineOp
    for (int32_t i = 0; i < L; ++i)
        ops[i] = (i < f0 || i >= f1) ? macRefineOp(p[l2i[i]], macs[l2i[i]]) : 1;
}
template<class K, class T, class CB>
static void geoMacSpheresW(const K* p, int32_t n, T* c4, float invTheta, const CB* box)
{
    for (int32_t i = 0; i < n; ++i)
    {
        Vec4<T> v = computeMinMacR2(p[i], invTheta, asBox<T>(box));
        for (int d = 0; d < 4; ++d)
            c4[4 * i + d] = v[d];
    }
}
template<class K, class T, class CB>
static void setMacW(const K* p, int32_t n, T* c4, float invTheta, const CB* box)
{
    // setMacKernel [focus/source_center_gpu.cu] over the CPU computeVecMacR2
    for (int32_t i = 0; i < n; ++i)
    {
        Vec3<T> c{c4[4 * i], c4[4 * i + 1], c4[4 * i + 2]};
        T mac         = computeVecMacR2(p[i], c, invTheta, asBox<T>(box));
        c4[4 * i + 3] = (c4[4 * i + 3] != T(0)) ? mac : T(0);
    }
}
template<class K, class T, class CB>
static void markMacsW(const K* p, const int32_t* co, const T* c4, const CB* box, const K* focus, int32_t nf, int limit,
                      char* markings)
{
    markMacs(p, co, reinterpret_cast<const Vec4<T>*>(c4), asBox<T>(box), focus, nf, bool(limit), markings);
}

extern "C" {
#define REF_FOCUS(KS, K)                                                                                               \
    void cstone_ref_rebalance_decision_essential_##KS(const K* p, const int32_t* co, const int32_t* pa,                \
                                                      const uint32_t* c, const char* macs, K fs, K fe,                 \
                                                      uint32_t bucket, int32_t* ops, int32_t n)                        \
    { rebalanceEssentialW<K>(p, co, pa, c, macs, fs, fe, bucket, ops, n); }                                            \
    void cstone_ref_mac_refine_decision_##KS(const K* p, const char* macs, const int32_t* l2i, int32_t L, int32_t f0,  \
                                             int32_t f1, int32_t* ops)                                                 \
    { macRefineW<K>(p, macs, l2i, L, f0, f1, ops); }                                                                   \
    int cstone_ref_protect_ancestors_##KS(const K* p, const int32_t* pa, int32_t* ops, int32_t n)                      \
    { return protectAncestors<K>(gsl::span<const K>(p, size_t(n)), pa, ops) ? 1 : 0; }                                 \
    int cstone_ref_enforce_keys_##KS(const K* keys, int32_t nk, const K* p, const int32_t* co, const int32_t* pa,      \
                                     int32_t* ops)                                                                     \
    { return int(enforceKeys<K>(gsl::span<const K>(keys, size_t(nk)), p, co, pa, ops)); }                              \
    void cstone_ref_range_count_##KS(const K* leaves, int32_t nKeys, const uint32_t* counts, const K* lf, int32_t nf,  \
                                     const int32_t* idx, int32_t ni, uint32_t* cf)                                     \
    {                                                                                                                  \
        rangeCount<K>(gsl::span<const K>(leaves, size_t(nKeys)), gsl::span<const unsigned>(counts, size_t(nKeys - 1)), \
                      gsl::span<const K>(lf, size_t(nf)), gsl::span<const TreeNodeIndex>(idx, size_t(ni)),             \
                      gsl::span<unsigned>(cf, size_t(nf - 1)));                                                        \
    }
REF_FOCUS(k32, uint32_t)
REF_FOCUS(k64, uint64_t)
#define REF_MAC(KS, K, TS, T)                                                                                          \
    void cstone_ref_geo_mac_spheres_##KS##_##TS(const K* p, int32_t n, T* c4, float it, const cstone_box_##TS* box)    \
    { geoMacSpheresW<K, T>(p, n, c4, it, box); }                                                                       \
    void cstone_ref_set_mac_##KS##_##TS(const K* p, int32_t n, T* c4, float it, const cstone_box_##TS* box)            \
    { setMacW<K, T>(p, n, c4, it, box); }                                                                              \
    void cstone_ref_mark_macs_##KS##_##TS(const K* p, const int32_t* co, const T* c4, const cstone_box_##TS* box,      \
                                          const K* focus, int32_t nf, int limit, char* markings)                       \
    { markMacsW<K, T>(p, co, c4, box, focus, nf, limit, markings); }
REF_MAC(k32, uint32_t, f32, float)
REF_MAC(k32, uint32_t, f64, double)
REF_MAC(k64, uint64_t, f32, float)
REF_MAC(k64, uint64_t, f64, double)
}

// ------------------------------------------------------------------------------------------------
// n3 device-kernel counterparts: expansion centres (reference build only)
// ------------------------------------------------------------------------------------------------
template<class Tc, class Tm, class Tf>
static void leafMassCenterW(const Tc* x, const Tc* y, const Tc* z, const Tm* m, size_t n, const int32_t* l2i, int32_t L,
                            const uint32_t* layout, Tf* c4)
{
    computeLeafMassCenter<Tc, Tm, Tf>({x, n}, {y, n}, {z, n}, {m, n}, {l2i, size_t(L)}, layout,
                                      reinterpret_cast<SourceCenterType<Tf>*>(c4));
}
template<class T>
static void upsweepCentersW(int32_t numLevels, const int32_t* levelRange, const int32_t* co, T* c4)
{
    // as FocusedOctree::updateCenters does [focus/octree_focus_mpi.hpp:421]
    upsweep(gsl::span<const TreeNodeIndex>(levelRange, size_t(numLevels + 1)),
            gsl::span<const TreeNodeIndex>(co, size_t(levelRange[numLevels])),
            reinterpret_cast<SourceCenterType<T>*>(c4), CombineSourceCenter<T>{});
}
extern "C" {
void cstone_ref_compute_leaf_source_center_f64_f64_f64(const double* x, const double* y, const double* z,
                                                       const double* m, size_t n, const int32_t* l2i, int32_t L,
                                                       const uint32_t* layout, double* c4)
{ leafMassCenterW<double, double, double>(x, y, z, m, n, l2i, L, layout, c4); }
void cstone_ref_compute_leaf_source_center_f64_f32_f64(const double* x, const double* y, const double* z,
                                                       const float* m, size_t n, const int32_t* l2i, int32_t L,
                                                       const uint32_t* layout, double* c4)
{ leafMassCenterW<double, float, double>(x, y, z, m, n, l2i, L, layout, c4); }
void cstone_ref_compute_leaf_source_center_f32_f32_f32(const float* x, const float* y, const float* z, const float* m,
                                                       size_t n, const int32_t* l2i, int32_t L, const uint32_t* layout,
                                                       float* c4)
{ leafMassCenterW<float, float, float>(x, y, z, m, n, l2i, L, layout, c4); }
void cstone_ref_upsweep_centers_f32(int32_t nl, const int32_t* lr, const int32_t* co, float* c4)
{ upsweepCentersW<float>(nl, lr, co, c4); }
void cstone_ref_upsweep_centers_f64(int32_t nl, const int32_t* lr, const int32_t* co, double* c4)
{ upsweepCentersW<double>(nl, lr, co, c4); }
}

// ------------------------------------------------------------------------------------------------
// a19: the reference's single-rank focus tree (reference build only)
// ------------------------------------------------------------------------------------------------
#include "cstone/focus/octree_focus.hpp"

template<class K, class T, class CB>
static int focusUpdateW(void* h, const CB* box, const K* keys, size_t n, K fs, K fe, const K* mand, int32_t nm)
{
    auto* t = static_cast<FocusedOctreeSingleNode<K>*>(h);
    return t->update(asBox<T>(box), gsl::span<const K>(keys, n), fs, fe, gsl::span<const K>(mand, size_t(nm))) ? 1 : 0;
}
extern "C" {
#define REF_FOCUS_TREE(KS, K)                                                                                          \
    void* cstone_ref_focus_tree_create_##KS(uint32_t bucket, float theta)                                              \
    { return new FocusedOctreeSingleNode<K>(bucket, theta); }                                                          \
    void cstone_ref_focus_tree_destroy_##KS(void* h) { delete static_cast<FocusedOctreeSingleNode<K>*>(h); }           \
    int cstone_ref_focus_tree_update_##KS##_f32(void* h, const cstone_box_f32* box, const K* keys, size_t n, K fs,     \
                                                K fe, const K* mand, int32_t nm)                                       \
    { return focusUpdateW<K, float>(h, box, keys, n, fs, fe, mand, nm); }                                              \
    int cstone_ref_focus_tree_update_##KS##_f64(void* h, const cstone_box_f64* box, const K* keys, size_t n, K fs,     \
                                                K fe, const K* mand, int32_t nm)                                       \
    { return focusUpdateW<K, double>(h, box, keys, n, fs, fe, mand, nm); }                                             \
    int32_t cstone_ref_focus_tree_num_leaves_##KS(void* h)                                                             \
    { return int32_t(static_cast<FocusedOctreeSingleNode<K>*>(h)->treeLeaves().size()) - 1; }                          \
    void cstone_ref_focus_tree_get_##KS(void* h, K* leaves, uint32_t* counts)                                          \
    {                                                                                                                  \
        auto* t = static_cast<FocusedOctreeSingleNode<K>*>(h);                                                         \
        auto l  = t->treeLeaves();                                                                                     \
        auto c  = t->leafCounts();                                                                                     \
        std::copy(l.begin(), l.end(), leaves);                                                                         \
        std::copy(c.begin(), c.end(), counts);                                                                         \
    }
REF_FOCUS_TREE(k32, uint32_t)
REF_FOCUS_TREE(k64, uint64_t)
}

// one step of the reference's macRefine on a leaf array [focus/octree_focus.hpp:260-295]; leaves has room for
// capacity + 1 keys, *numLeaves is updated; returns converged, or -1 if the new tree does not fit
template<class K, class T, class CB>
static int macRefineStepW(K* leavesIo, int32_t* numLeaves, int32_t capacity, K ofs, K ofe, K fs, K fe, float invTheta,
                          const CB* box)
{
    std::vector<K> leaves(leavesIo, leavesIo + *numLeaves + 1);
    OctreeData<K, CpuTag> tree;
    tree.resize(nNodes(leaves));
    updateInternalTree<K>(leaves, tree.data());
    std::vector<SourceCenterType<T>> centers;
    std::vector<char> macs;
    bool conv = macRefine(tree, leaves, centers, macs, ofs, ofe, fs, fe, invTheta, asBox<T>(box));
    if (int32_t(nNodes(leaves)) > capacity) return -1;
    std::copy(leaves.begin(), leaves.end(), leavesIo);
    *numLeaves = int32_t(nNodes(leaves));
    return conv ? 1 : 0;
}
extern "C" {
#define REF_MAC_REFINE(KS, K, TS, T)                                                                                   \
    int cstone_ref_mac_refine_step_##KS##_##TS(K* leaves, int32_t* n, int32_t cap, K ofs, K ofe, K fs, K fe,           \
                                               float invTheta, const cstone_box_##TS* box)                             \
    { return macRefineStepW<K, T>(leaves, n, cap, ofs, ofe, fs, fe, invTheta, box); }
REF_MAC_REFINE(k32, uint32_t, f32, float)
REF_MAC_REFINE(k32, uint32_t, f64, double)
REF_MAC_REFINE(k64, uint64_t, f32, float)
REF_MAC_REFINE(k64, uint64_t, f64, double)
}
