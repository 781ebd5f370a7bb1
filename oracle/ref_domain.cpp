/*
 * ref_domain.cpp — TEST INFRASTRUCTURE ONLY: the multi-rank oracle.
 *
 * Compiles the UNMODIFIED reference class cstone::Domain<KeyType,T,CpuTag> (domain/domain.hpp:66-699) against the
 * thread-rank mpi.h of oracle/mpi_shim, so that P ranks run in one process, driven the way the reference's own
 * integration test drives them (test/integration_mpi/domain_nranks.cpp:103-149).  Built into
 * oracle/_ref/liboracle_refdom.so by oracle/Makefile; reads nothing of the product.
 *
 * C API (ctypes, oracle/loader.py RefDomain): create(P, ...) -> handle; set(rank, arrays); sync() runs one
 * Domain::sync on every rank concurrently; getters return each rank's arrays, buffer description, box, global
 * tree + assignment, focus tree leaves / counts / assignment, layout; find_neighbors runs the reference's
 * findNeighbors over the rank's assigned range; exchange_halos repeats the halo pattern for one more array.
 */
#include <mpi.h>
#include <omp.h>

#include <chrono>
#include <cmath>
#include <cstring>
#include <memory>
#include <tuple>
#include <vector>

#include "cstone/domain/domain.hpp"
#include "cstone/findneighbors.hpp"

using namespace cstone;

namespace
{

template<class K, class T>
struct RankState
{
    std::unique_ptr<Domain<K, T, CpuTag>> domain;
    std::vector<K> keys;
    std::vector<T> x, y, z, h;
    std::vector<T> p0, p1, p2, p3; // up to four conserved properties (0, 1, 2 or 4 are wired)
    std::vector<T> m; // masses (syncGrav)
    std::vector<T> s1, s2, s3;
    double syncSeconds = 0, nbSeconds = 0;
    std::vector<unsigned> nb, nc; // neighbour lists / counts kept between bench steps
    double ncSum = 0;
};

template<class K, class T>
struct World
{
    int P = 1, nprops = 0, ompPerRank = 1;
    std::vector<RankState<K, T>> ranks;
};

template<class K, class T>
void syncRank(int rank, void* arg)
{
    auto* w = static_cast<World<K, T>*>(arg);
    auto& r = w->ranks[rank];
    omp_set_num_threads(w->ompPerRank);
    r.keys.resize(r.x.size());
    MPI_Barrier(MPI_COMM_WORLD);
    auto t0 = std::chrono::steady_clock::now();
    if (w->nprops == 0) r.domain->sync(r.keys, r.x, r.y, r.z, r.h, std::tuple{}, std::tie(r.s1, r.s2, r.s3));
    else if (w->nprops == 1) r.domain->sync(r.keys, r.x, r.y, r.z, r.h, std::tie(r.p0), std::tie(r.s1, r.s2, r.s3));
    else r.domain->sync(r.keys, r.x, r.y, r.z, r.h, std::tie(r.p0, r.p1), std::tie(r.s1, r.s2, r.s3));
    MPI_Barrier(MPI_COMM_WORLD);
    r.syncSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

//! Domain::syncGrav [ref: domain/domain.hpp:245-325] with the masses set by set_mass; one conserved property at most
template<class K, class T>
void syncGravRank(int rank, void* arg)
{
    auto* w = static_cast<World<K, T>*>(arg);
    auto& r = w->ranks[rank];
    omp_set_num_threads(w->ompPerRank);
    r.keys.resize(r.x.size());
    MPI_Barrier(MPI_COMM_WORLD);
    if (w->nprops == 0) r.domain->syncGrav(r.keys, r.x, r.y, r.z, r.h, r.m, std::tuple{}, std::tie(r.s1, r.s2, r.s3));
    else r.domain->syncGrav(r.keys, r.x, r.y, r.z, r.h, r.m, std::tie(r.p0), std::tie(r.s1, r.s2, r.s3));
    MPI_Barrier(MPI_COMM_WORLD);
}

template<class K, class T>
struct HaloArg
{
    World<K, T>* w;
    std::vector<std::vector<T>*> arrays;
};

template<class K, class T>
void haloRank(int rank, void* arg)
{
    auto* a = static_cast<HaloArg<K, T>*>(arg);
    auto& r = a->w->ranks[rank];
    omp_set_num_threads(a->w->ompPerRank);
    r.domain->exchangeHalos(std::tie(*a->arrays[rank]), r.s1, r.s2);
    MPI_Barrier(MPI_COMM_WORLD);
}

template<class K, class T>
struct NbArg
{
    World<K, T>* w;
    unsigned ngmax;
    unsigned** nb;
    unsigned** nc;
};

template<class K, class T>
void nbRank(int rank, void* arg)
{
    auto* a = static_cast<NbArg<K, T>*>(arg);
    auto& r = a->w->ranks[rank];
    omp_set_num_threads(a->w->ompPerRank);
    MPI_Barrier(MPI_COMM_WORLD);
    auto t0 = std::chrono::steady_clock::now();
    findNeighbors(r.x.data(), r.y.data(), r.z.data(), r.h.data(), r.domain->startIndex(), r.domain->endIndex(),
                  r.domain->box(), r.domain->octreeProperties(), a->ngmax, a->nb[rank], a->nc[rank]);
    MPI_Barrier(MPI_COMM_WORLD);
    r.nbSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

/* One step of bench.py's reference arm on one rank: the drift bench.py applies on the GPU (same expressions, same
 * order), Domain::sync, findNeighbors over the assigned range into buffers that live across steps. */
template<class K, class T>
struct StepArg
{
    World<K, T>* w;
    double amp; // signed amplitude of this step
    unsigned ngmax;
};

template<class K, class T>
void stepRank(int rank, void* arg)
{
    auto* a = static_cast<StepArg<K, T>*>(arg);
    auto* w = a->w;
    auto& r = w->ranks[rank];
    omp_set_num_threads(w->ompPerRank);
    MPI_Barrier(MPI_COMM_WORLD);
    auto t0 = std::chrono::steady_clock::now();
    const T wv  = T(2.0 * 3.14159265358979323846 * 37.0);
    const T amp = T(a->amp);
    const std::size_t n = r.x.size();
#pragma omp parallel for
    for (std::size_t i = 0; i < n; ++i)
        r.x[i] += amp * std::sin(r.y[i] * wv);
#pragma omp parallel for
    for (std::size_t i = 0; i < n; ++i)
        r.y[i] += amp * std::sin(r.z[i] * wv);
#pragma omp parallel for
    for (std::size_t i = 0; i < n; ++i)
        r.z[i] += amp * std::sin(r.x[i] * wv);
    r.keys.resize(r.x.size());
    if (w->nprops == 0) r.domain->sync(r.keys, r.x, r.y, r.z, r.h, std::tuple{}, std::tie(r.s1, r.s2, r.s3));
    else if (w->nprops == 1) r.domain->sync(r.keys, r.x, r.y, r.z, r.h, std::tie(r.p0), std::tie(r.s1, r.s2, r.s3));
    else if (w->nprops == 2) r.domain->sync(r.keys, r.x, r.y, r.z, r.h, std::tie(r.p0, r.p1), std::tie(r.s1, r.s2, r.s3));
    else r.domain->sync(r.keys, r.x, r.y, r.z, r.h, std::tie(r.p0, r.p1, r.p2, r.p3), std::tie(r.s1, r.s2, r.s3));
    MPI_Barrier(MPI_COMM_WORLD);
    r.syncSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    auto t1 = std::chrono::steady_clock::now();
    const std::size_t nw = r.domain->nParticles();
    if (r.nb.size() < nw * a->ngmax) r.nb.resize(std::size_t(double(nw) * 1.05) * a->ngmax);
    if (r.nc.size() < nw) r.nc.resize(std::size_t(double(nw) * 1.05));
    findNeighbors(r.x.data(), r.y.data(), r.z.data(), r.h.data(), r.domain->startIndex(), r.domain->endIndex(),
                  r.domain->box(), r.domain->octreeProperties(), a->ngmax, r.nb.data(), r.nc.data());
    MPI_Barrier(MPI_COMM_WORLD);
    r.nbSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();
    double s    = 0;
#pragma omp parallel for reduction(+ : s)
    for (std::size_t i = 0; i < nw; ++i)
        s += r.nc[i];
    r.ncSum = s;
}

template<class T>
Box<T> toBox(const T* lim, const int* pbc)
{
    auto b = [](int p) { return p ? BoundaryType::periodic : BoundaryType::open; };
    return Box<T>(lim[0], lim[1], lim[2], lim[3], lim[4], lim[5], b(pbc[0]), b(pbc[1]), b(pbc[2]));
}

template<class K, class T>
void* create(int P, unsigned bucket, unsigned bucketFocus, float theta, const T* lim, const int* pbc, int nprops,
             int ompPerRank)
{
    auto* w       = new World<K, T>;
    w->P          = P;
    w->nprops     = nprops;
    w->ompPerRank = ompPerRank > 0 ? ompPerRank : 1;
    w->ranks.resize(P);
    for (int r = 0; r < P; ++r)
        w->ranks[r].domain =
            std::make_unique<Domain<K, T, CpuTag>>(r, P, bucket, bucketFocus, theta, toBox<T>(lim, pbc));
    return w;
}

template<class K, class T>
int setRank(void* h, int rank, uint64_t n, const T* x, const T* y, const T* z, const T* hh, const T* p0, const T* p1)
{
    auto* w = static_cast<World<K, T>*>(h);
    auto& r = w->ranks[rank];
    r.x.assign(x, x + n), r.y.assign(y, y + n), r.z.assign(z, z + n), r.h.assign(hh, hh + n);
    if (w->nprops > 0) r.p0.assign(p0, p0 + n);
    if (w->nprops > 1) r.p1.assign(p1, p1 + n);
    if (w->nprops > 2)
    {
        // bench.py's c4 fields: double(globalId) * {1,2,3,4}; p0 carries the id, the others are derived here
        r.p1.resize(n), r.p2.resize(n), r.p3.resize(n);
        for (uint64_t i = 0; i < n; ++i)
            r.p1[i] = T(2) * p0[i], r.p2[i] = T(3) * p0[i], r.p3[i] = T(4) * p0[i];
    }
    return 0;
}

template<class K, class T>
void info(void* h, int rank, int64_t* out, T* box6)
{
    auto* w  = static_cast<World<K, T>*>(h);
    auto& r  = w->ranks[rank];
    auto& d  = *r.domain;
    out[0]   = d.startIndex();
    out[1]   = d.endIndex();
    out[2]   = d.nParticlesWithHalos();
    out[3]   = d.focusTree().octreeViewAcc().numLeafNodes;
    out[4]   = d.globalTree().numLeafNodes();
    out[5]   = d.startCell();
    out[6]   = d.endCell();
    out[7]   = int64_t(r.x.size());
    auto box = d.box();
    box6[0] = box.xmin(), box6[1] = box.xmax(), box6[2] = box.ymin(), box6[3] = box.ymax(), box6[4] = box.zmin(),
    box6[5] = box.zmax();
}

//! which: 0 keys, 1 x, 2 y, 3 z, 4 h, 5 p0, 6 p1 (rank arrays, length info[7])
template<class K, class T>
int getArray(void* h, int rank, int which, void* out)
{
    auto& r = static_cast<World<K, T>*>(h)->ranks[rank];
    auto cp = [out](const auto& v) { std::memcpy(out, v.data(), v.size() * sizeof(v[0])); };
    switch (which)
    {
        case 0: cp(r.keys); break;
        case 1: cp(r.x); break;
        case 2: cp(r.y); break;
        case 3: cp(r.z); break;
        case 4: cp(r.h); break;
        case 5: cp(r.p0); break;
        case 6: cp(r.p1); break;
        case 7: cp(r.m); break;
        default: return 1;
    }
    return 0;
}

//! trees: 0 global leaves (K[Lg+1]), 1 focus leaves (K[Lf+1]), 2 focus leaf counts (u32[Lf]), 3 layout (u32[Lf+1]),
//!        4 focus assignment (int[2P]: start,end per rank)
template<class K, class T>
int getTree(void* h, int rank, int which, void* out)
{
    auto* w = static_cast<World<K, T>*>(h);
    auto& d = *w->ranks[rank].domain;
    switch (which)
    {
        case 0:
        {
            auto l = d.globalTree().treeLeaves();
            std::memcpy(out, l.data(), l.size() * sizeof(K));
            break;
        }
        case 1:
        {
            auto l = d.focusTree().treeLeaves();
            std::memcpy(out, l.data(), l.size() * sizeof(K));
            break;
        }
        case 2:
        {
            auto c = d.focusTree().leafCounts();
            std::memcpy(out, c.data(), c.size() * sizeof(unsigned));
            break;
        }
        case 3:
        {
            auto l = d.layout();
            std::memcpy(out, l.data(), l.size() * sizeof(LocalIndex));
            break;
        }
        case 5: // expansion centres of the focus tree after syncGrav: (x, y, z, mac radius^2) per node
        {
            auto c = d.focusTree().expansionCentersAcc();
            std::memcpy(out, c.data(), c.size() * sizeof(c[0]));
            break;
        }
        case 6: // number of nodes of the focus tree (int64)
        {
            *static_cast<int64_t*>(out) = int64_t(d.focusTree().octreeViewAcc().numNodes);
            break;
        }
        case 4:
        {
            auto a = d.focusTree().assignment();
            int* o = static_cast<int*>(out);
            for (int q = 0; q < w->P; ++q)
                o[2 * q] = a[q].start(), o[2 * q + 1] = a[q].end();
            break;
        }
        default: return 1;
    }
    return 0;
}

//! the reference's findPeersMac [ref: traversal/peers.hpp:62-117] on a global leaf array and rank boundaries
template<class K, class T>
int findPeers(int myRank, int numRanks, const K* boundaries, const K* leaves, int numLeaves, const T* lim, const int* pbc,
              float invThetaEff, int* isPeer)
{
    Octree<K> tree;
    tree.update(leaves, numLeaves);
    SfcAssignment<K> assignment(numRanks);
    for (int r = 0; r <= numRanks; ++r)
        assignment.set(r, boundaries[r], 0);
    std::vector<int> peers = findPeersMac(myRank, assignment, tree, toBox<T>(lim, pbc), invThetaEff);
    for (int r = 0; r < numRanks; ++r)
        isPeer[r] = 0;
    for (int p : peers)
        isPeer[p] = 1;
    return 0;
}

} // namespace

#define REFDOM_DEF(KS, K, TS, T)                                                                                       \
    void* cstone_refdom_create_##KS##_##TS(int P, unsigned bucket, unsigned bucketFocus, float theta, const T* lim,    \
                                           const int* pbc, int nprops, int ompPerRank)                                 \
    {                                                                                                                  \
        return create<K, T>(P, bucket, bucketFocus, theta, lim, pbc, nprops, ompPerRank);                              \
    }                                                                                                                  \
    void cstone_refdom_destroy_##KS##_##TS(void* h) { delete static_cast<World<K, T>*>(h); }                           \
    int cstone_refdom_set_##KS##_##TS(void* h, int rank, uint64_t n, const T* x, const T* y, const T* z, const T* hh,  \
                                      const T* p0, const T* p1)                                                        \
    {                                                                                                                  \
        return setRank<K, T>(h, rank, n, x, y, z, hh, p0, p1);                                                         \
    }                                                                                                                  \
    int cstone_refdom_find_peers_##KS##_##TS(int myRank, int numRanks, const K* boundaries, const K* leaves,           \
                                             int numLeaves, const T* lim, const int* pbc, float invThetaEff,           \
                                             int* isPeer)                                                              \
    {                                                                                                                  \
        return findPeers<K, T>(myRank, numRanks, boundaries, leaves, numLeaves, lim, pbc, invThetaEff, isPeer);        \
    }                                                                                                                  \
    /* updateTreelet [ref: tree/csarray.hpp:477-488] on leaves [first, last) of a full leaf array, as focusTransfer calls \
     * it [ref: focus/exchange_focus.hpp:380-405]; out has room for 4096 * (last - first) + 1 keys; returns the count */ \
    int cstone_refdom_update_treelet_##KS##_##TS(const K* leaves, const unsigned* counts, int first, int last,         \
                                                 unsigned bucket, K* out)                                              \
    {                                                                                                                  \
        auto tl = updateTreelet(gsl::span<const K>(leaves + first, size_t(last - first) + 1),                          \
                                gsl::span<const unsigned>(counts + first, size_t(last - first)), bucket);              \
        std::memcpy(out, tl.data(), tl.size() * sizeof(K));                                                            \
        return int(tl.size());                                                                                         \
    }                                                                                                                  \
    int cstone_refdom_set_mass_##KS##_##TS(void* h, int rank, uint64_t n, const T* m)                                  \
    {                                                                                                                  \
        static_cast<World<K, T>*>(h)->ranks[rank].m.assign(m, m + n);                                                  \
        return 0;                                                                                                      \
    }                                                                                                                  \
    int cstone_refdom_sync_grav_##KS##_##TS(void* h)                                                                   \
    {                                                                                                                  \
        auto* w = static_cast<World<K, T>*>(h);                                                                        \
        return mpi_shim_run(w->P, syncGravRank<K, T>, w);                                                              \
    }                                                                                                                  \
    int cstone_refdom_sync_##KS##_##TS(void* h)                                                                        \
    {                                                                                                                  \
        auto* w = static_cast<World<K, T>*>(h);                                                                        \
        return mpi_shim_run(w->P, syncRank<K, T>, w);                                                                  \
    }                                                                                                                  \
    void cstone_refdom_info_##KS##_##TS(void* h, int rank, int64_t* out8, T* box6)                                     \
    {                                                                                                                  \
        info<K, T>(h, rank, out8, box6);                                                                               \
    }                                                                                                                  \
    int cstone_refdom_get_array_##KS##_##TS(void* h, int rank, int which, void* out)                                   \
    {                                                                                                                  \
        return getArray<K, T>(h, rank, which, out);                                                                    \
    }                                                                                                                  \
    int cstone_refdom_get_tree_##KS##_##TS(void* h, int rank, int which, void* out)                                    \
    {                                                                                                                  \
        return getTree<K, T>(h, rank, which, out);                                                                     \
    }                                                                                                                  \
    int cstone_refdom_find_neighbors_##KS##_##TS(void* h, unsigned ngmax, unsigned** nb, unsigned** nc)                \
    {                                                                                                                  \
        auto* w = static_cast<World<K, T>*>(h);                                                                        \
        NbArg<K, T> a{w, ngmax, nb, nc};                                                                               \
        return mpi_shim_run(w->P, nbRank<K, T>, &a);                                                                   \
    }                                                                                                                  \
    /* arrays[rank]: T[nParticlesWithHalos(rank)], assigned range filled by the caller; halos are filled in place */   \
    int cstone_refdom_exchange_halos_##KS##_##TS(void* h, T** arrays)                                                  \
    {                                                                                                                  \
        auto* w = static_cast<World<K, T>*>(h);                                                                        \
        std::vector<std::vector<T>> vecs(w->P);                                                                        \
        HaloArg<K, T> a{w, {}};                                                                                        \
        for (int r = 0; r < w->P; ++r)                                                                                 \
        {                                                                                                              \
            vecs[r].assign(arrays[r], arrays[r] + w->ranks[r].domain->nParticlesWithHalos());                          \
            a.arrays.push_back(&vecs[r]);                                                                              \
        }                                                                                                              \
        int st = mpi_shim_run(w->P, haloRank<K, T>, &a);                                                               \
        for (int r = 0; r < w->P; ++r)                                                                                 \
            std::memcpy(arrays[r], vecs[r].data(), vecs[r].size() * sizeof(T));                                        \
        return st;                                                                                                     \
    }                                                                                                                  \
    /* bench.py reference arm: drift + sync + findNeighbors on every rank; out3 = assigned particles (sum over ranks),  \
     * sum of neighbour counts, halos (sum over ranks) */                                                              \
    int cstone_refdom_bench_step_##KS##_##TS(void* h, double amp, unsigned ngmax, double* out3)                        \
    {                                                                                                                  \
        auto* w = static_cast<World<K, T>*>(h);                                                                        \
        StepArg<K, T> a{w, amp, ngmax};                                                                                \
        int st = mpi_shim_run(w->P, stepRank<K, T>, &a);                                                               \
        out3[0] = out3[1] = out3[2] = 0;                                                                               \
        for (auto& r : w->ranks)                                                                                       \
        {                                                                                                              \
            out3[0] += double(r.domain->nParticles());                                                                 \
            out3[1] += r.ncSum;                                                                                        \
            out3[2] += double(r.domain->nParticlesWithHalos() - r.domain->nParticles());                               \
        }                                                                                                              \
        return st;                                                                                                     \
    }                                                                                                                  \
    /* neighbour results of the last bench step of one rank (views into the oracle's buffers) */                       \
    void cstone_refdom_bench_neighbors_##KS##_##TS(void* h, int rank, const unsigned** nb, const unsigned** nc)        \
    {                                                                                                                  \
        auto& r = static_cast<World<K, T>*>(h)->ranks[rank];                                                           \
        *nb = r.nb.data(), *nc = r.nc.data();                                                                          \
    }                                                                                                                  \
    void cstone_refdom_seconds_##KS##_##TS(void* h, double* sync, double* nb)                                          \
    {                                                                                                                  \
        auto* w = static_cast<World<K, T>*>(h);                                                                        \
        *sync = 0, *nb = 0;                                                                                            \
        for (auto& r : w->ranks)                                                                                       \
            *sync = std::max(*sync, r.syncSeconds), *nb = std::max(*nb, r.nbSeconds);                                  \
    }

extern "C" {
REFDOM_DEF(k64, uint64_t, f64, double)
REFDOM_DEF(k64, uint64_t, f32, float)
REFDOM_DEF(k32, uint32_t, f64, double)
}
