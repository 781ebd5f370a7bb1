/*
 * cstone_b200/domain.hpp — cstone::Domain<KeyType, T, GpuTag> over libcstone_b200.so
 *
 * Source-compatible with the public surface of the reference class [ref: domain/domain.hpp:66-437] for the hot path:
 *   Domain(rank, nRanks, bucketSize, bucketSizeFocus, theta, box)
 *   sync(keys, x, y, z, h, std::tie(properties...), std::tie(scratch...))
 *   exchangeHalos(std::tie(arrays...), sendBuffer, receiveBuffer)
 *   startIndex / endIndex / nParticles / nParticlesWithHalos / box / octreeProperties / layout
 * All vectors are cstone::DeviceVector; particle data never leave the device.  The scratch-buffer arguments are
 * accepted for source compatibility and left untouched (the engine owns grow-only scratch).
 * Multi-GPU: one process per GPU; hand the engine an NCCL communicator with setNcclComm() (or initNcclComm() with a
 * unique id broadcast through the application's MPI / torch.distributed).
 * Not provided (out of scope): syncGrav, updateExpansionCenters, focusTree(), CpuTag.
 */
#pragma once

#include <array>
#include <cstdint>
#include <tuple>
#include <utility>
#include <vector>

#include "../cstone_b200_domain.h"
#include "box.hpp"
#include "device_vector.hpp"

namespace cstone
{

template<class T>
struct Vec3
{
    T v[3];
    T& operator[](int i) { return v[i]; }
    const T& operator[](int i) const { return v[i]; }
};

//! [ref: tree/octree.hpp:296-317 OctreeNsView] device pointers, valid until the next sync
template<class T, class KeyType>
struct OctreeNsView
{
    TreeNodeIndex numLeafNodes;
    const KeyType* prefixes;
    const TreeNodeIndex* childOffsets;
    const TreeNodeIndex* internalToLeaf;
    const TreeNodeIndex* levelRange;
    const KeyType* leaves;
    //! index of the first particle of each leaf
    const LocalIndex* layout;
    //! geometrical node centres and half sizes
    const Vec3<T>* centers;
    const Vec3<T>* sizes;
    //! enlarges the target boxes for slightly outdated trees; 1.0 for converged trees (the Domain's are)
    float searchExtFactor{1.0};
};

namespace detail
{
template<class K, class T>
struct DomainApi;

#define CSB_DOMAIN_API(KS, K, TS, T)                                                                                   \
    template<>                                                                                                         \
    struct DomainApi<K, T>                                                                                             \
    {                                                                                                                  \
        static constexpr auto create        = cstone_b200_domain_create_##KS##_##TS;                                   \
        static constexpr auto destroy       = cstone_b200_domain_destroy_##KS##_##TS;                                  \
        static constexpr auto setComm       = cstone_b200_domain_set_comm_##KS##_##TS;                                 \
        static constexpr auto setSfcKind    = cstone_b200_domain_set_sfc_kind_##KS##_##TS;                             \
        static constexpr auto setEndIndex   = cstone_b200_domain_set_end_index_##KS##_##TS;                            \
        static constexpr auto setHaloFactor = cstone_b200_domain_set_halo_factor_##KS##_##TS;                          \
        static constexpr auto setHaloMode   = cstone_b200_domain_set_halo_mode_##KS##_##TS;                            \
        static constexpr auto cells         = cstone_b200_domain_cells_##KS##_##TS;                                    \
        static constexpr auto syncBegin     = cstone_b200_domain_sync_begin_##KS##_##TS;                               \
        static constexpr auto syncGravBegin = cstone_b200_domain_sync_grav_begin_##KS##_##TS;                          \
        static constexpr auto syncFinish    = cstone_b200_domain_sync_finish_##KS##_##TS;                              \
        static constexpr auto exchangeHalos = cstone_b200_domain_exchange_halos_##KS##_##TS;                           \
        static constexpr auto reapplySync   = cstone_b200_domain_reapply_sync_##KS##_##TS;                             \
        static constexpr auto info          = cstone_b200_domain_info_##KS##_##TS;                                     \
        static constexpr auto box           = cstone_b200_domain_box_##KS##_##TS;                                      \
        static constexpr auto stageMs       = cstone_b200_domain_stage_ms_##KS##_##TS;                                 \
        static constexpr auto treeView      = cstone_b200_domain_tree_view_##KS##_##TS;                                \
        static constexpr auto globalTree    = cstone_b200_domain_global_tree_##KS##_##TS;                              \
        static constexpr auto findNeighbors = cstone_b200_domain_find_neighbors_##KS##_##TS;                           \
    };
CSB_DOMAIN_API(k64, uint64_t, f64, double)
CSB_DOMAIN_API(k64, uint64_t, f32, float)
CSB_DOMAIN_API(k32, uint32_t, f64, double)
CSB_DOMAIN_API(k32, uint32_t, f32, float)
#undef CSB_DOMAIN_API
} // namespace detail

template<class KeyType, class T, class Accelerator = GpuTag>
class Domain
{
    static_assert(std::is_same_v<Accelerator, GpuTag>, "this library provides the GpuTag Domain only");
    static_assert(std::is_same_v<KeyType, uint32_t> || std::is_same_v<KeyType, uint64_t> ||
                      (std::is_unsigned_v<KeyType> && (sizeof(KeyType) == 4 || sizeof(KeyType) == 8)),
                  "SFC key type needs to be a 32- or 64-bit unsigned integer");
    using K   = std::conditional_t<sizeof(KeyType) == 8, uint64_t, uint32_t>;
    using Api = detail::DomainApi<K, T>;

public:
    using RealType = T;

    Domain(int rank,
           int nRanks,
           unsigned bucketSize,
           unsigned bucketSizeFocus,
           float theta,
           const Box<T>& box = Box<T>{0, 1})
        : myRank_(rank)
        , numRanks_(nRanks)
        , box_(box)
    {
        if (bucketSize < bucketSizeFocus)
        {
            throw std::runtime_error("The bucket size of the global tree must not be smaller than the bucket size"
                                     " of the focused tree\n");
        }
        d_ = Api::create(rank, nRanks, bucketSize, bucketSizeFocus, theta, box.c());
        if (!d_) throw std::runtime_error("Domain: engine creation failed\n");
    }
    Domain(const Domain&)            = delete;
    Domain& operator=(const Domain&) = delete;
    ~Domain()
    {
        Api::destroy(d_);
        if (ownedComm_) cstone_b200_nccl_comm_destroy(ownedComm_);
    }

    //! hand over an existing ncclComm_t (not owned)
    void setNcclComm(void* ncclComm) { detail::check(Api::setComm(d_, ncclComm), "Domain::setNcclComm"); }
    //! extension: CSTONE_SFC_MORTON instead of the Hilbert curve the reference's Domain hard-wires; before sync
    void setSfcKind(int sfcKind) { detail::check(Api::setSfcKind(d_, sfcKind), "Domain::setSfcKind"); }
    //! 128-byte id for ncclCommInitRank, to be created on rank 0 and broadcast by the application
    static std::array<char, 128> ncclUniqueId()
    {
        std::array<char, 128> id{};
        detail::check(cstone_b200_nccl_unique_id(id.data()), "ncclGetUniqueId");
        return id;
    }
    //! collective over all ranks of the domain; the communicator is owned by the Domain
    void initNcclComm(const std::array<char, 128>& id)
    {
        detail::check(cstone_b200_nccl_comm_init(id.data(), numRanks_, myRank_, &ownedComm_), "ncclCommInitRank");
        setNcclComm(ownedComm_);
    }

    /*! [ref: domain/domain.hpp:196-243]  pre- and postconditions as documented there (:118-179):
     *  on return every array has nParticlesWithHalos() elements in SFC order, x,y,z,h,keys valid everywhere,
     *  properties valid in [startIndex(), endIndex()). */
    template<class KeyVec, class VectorX, class VectorH, class... Vectors1, class... Vectors2>
    void sync(KeyVec& particleKeys,
              VectorX& x,
              VectorX& y,
              VectorX& z,
              VectorH& h,
              std::tuple<Vectors1&...> particleProperties,
              std::tuple<Vectors2&...> /*scratchBuffers*/)
    {
        syncImpl(false, particleKeys, x, y, z, h, particleProperties);
    }

    /*! [ref: domain/domain.hpp:245-325]  sync for N-body clients: the focus tree is refined with the vector MAC around
     *  the expansion centres of its cells, leaves that fail it become halos.  m travels like x, y, z, h (valid in the
     *  assigned range on return).  Needs setReferenceHalos(true) and more than one rank. */
    template<class KeyVec, class VectorX, class VectorH, class VectorM, class... Vectors1, class... Vectors2>
    void syncGrav(KeyVec& particleKeys,
                  VectorX& x,
                  VectorX& y,
                  VectorX& z,
                  VectorH& h,
                  VectorM& m,
                  std::tuple<Vectors1&...> particleProperties,
                  std::tuple<Vectors2&...> /*scratchBuffers*/)
    {
        static_assert(std::is_same_v<typename VectorM::value_type, T>, "masses must hold T");
        syncImpl(true, particleKeys, x, y, z, h, std::tuple_cat(std::tie(m), particleProperties));
    }

private:
    template<class KeyVec, class VectorX, class VectorH, class... Vectors1>
    void syncImpl(bool grav, KeyVec& particleKeys, VectorX& x, VectorX& y, VectorX& z, VectorH& h,
                  std::tuple<Vectors1&...> particleProperties)
    {
        static_assert(sizeof(typename KeyVec::value_type) == sizeof(K), "key vector does not match KeyType");
        static_assert(std::is_same_v<typename VectorX::value_type, T>, "coordinate vectors must hold T");
        static_assert(std::is_same_v<typename VectorH::value_type, T>, "h must hold T");
        static_assert(((sizeof(typename Vectors1::value_type) % 4 == 0 && sizeof(typename Vectors1::value_type) <= 16) && ...),
                      "property elements: 4, 8, 12 or 16 bytes (domain/buffer_description.hpp:98-110)");

        const std::size_t n = x.size();
        std::apply([&](auto&... p) { checkSizesEqual(n, y, z, h, p...); }, particleProperties);

        constexpr int np = sizeof...(Vectors1);
        std::array<const void*, np + 1> in{};
        std::array<void*, np + 1> out{};
        std::array<int, np + 1> bytes{};
        int i = 0;
        std::apply([&](auto&... p) { ((in[i] = p.data(), bytes[i] = int(sizeof(*p.data())), ++i), ...); },
                   particleProperties);

        long long nNew = grav ? Api::syncGravBegin(d_, n, x.data(), y.data(), z.data(), h.data(), np, in.data(), bytes.data())
                              : Api::syncBegin(d_, n, x.data(), y.data(), z.data(), h.data(), np, in.data(), bytes.data());
        if (nNew == CSTONE_E_SIZE_MISMATCH) throw std::runtime_error("Domain sync: input array sizes are inconsistent\n");
        detail::check(nNew < 0 ? nNew : 0, "Domain::sync");

        // contents are rewritten by syncFinish; only capacity matters.  Growth policy of the reference's reallocate
        // (util/reallocate.hpp): a vector that has to grow reserves allocGrowthRate_ times the new size
        auto grow = [&](auto& v)
        {
            if (v.capacity() < std::size_t(nNew)) v.reserve(std::size_t(double(nNew) * allocGrowthRate_));
            v.resize(nNew);
        };
        grow(particleKeys), grow(x), grow(y), grow(z), grow(h);
        std::apply([&](auto&... p) { (grow(p), ...); }, particleProperties);
        i = 0;
        std::apply([&](auto&... p) { ((out[i++] = p.data()), ...); }, particleProperties);
        detail::check(Api::syncFinish(d_, reinterpret_cast<K*>(particleKeys.data()), x.data(), y.data(), z.data(),
                                      h.data(), out.data()),
                      "Domain::sync");
        refresh();
    }

public:
    //! [ref: domain/domain.hpp:381-386]  send/receive buffers accepted for source compatibility
    template<class... Vectors, class SendBuffer, class ReceiveBuffer>
    void exchangeHalos(std::tuple<Vectors&...> arrays, SendBuffer&, ReceiveBuffer&) const
    {
        std::apply([this](auto&... a) { checkSizesEqual(info_[2], a...); }, arrays);
        constexpr int na = sizeof...(Vectors);
        std::array<void*, na + 1> ptrs{};
        std::array<int, na + 1> bytes{};
        int i = 0;
        std::apply([&](auto&... a) { ((ptrs[i] = a.data(), bytes[i] = int(sizeof(*a.data())), ++i), ...); }, arrays);
        detail::check(Api::exchangeHalos(d_, na, ptrs.data(), bytes.data()), "Domain::exchangeHalos");
    }

    /*! [ref: domain/domain.hpp:334-378]  reapply the exchange and ordering of the last sync to further arrays, which
     *  still have the layout the particle arrays had BEFORE that sync; on return they have nParticlesWithHalos()
     *  elements with the assigned range valid.  Buffers and ordering are accepted for source compatibility (the
     *  engine keeps the pattern itself); DEVICE vectors here, where the reference supports host vectors only. */
    template<class... Vectors, class SendBuffer, class ReceiveBuffer, class OVec>
    void reapplySync(std::tuple<Vectors&...> arrays, SendBuffer&, ReceiveBuffer&, OVec& /*ordering*/) const
    {
        static_assert(((sizeof(typename Vectors::value_type) % 4 == 0 && sizeof(typename Vectors::value_type) <= 16) && ...));
        constexpr int na = sizeof...(Vectors);
        std::array<const void*, na + 1> in{};
        std::array<void*, na + 1> out{};
        std::array<int, na + 1> bytes{};
        std::tuple<std::remove_reference_t<Vectors>...> fresh;
        int i = 0;
        std::apply([&](auto&... a) { ((in[i] = a.data(), bytes[i] = int(sizeof(*a.data())), ++i), ...); }, arrays);
        i = 0;
        std::apply([&](auto&... f) { ((f.resize(info_[2]), out[i++] = f.data()), ...); }, fresh);
        detail::check(Api::reapplySync(d_, na, in.data(), bytes.data(), out.data()), "Domain::reapplySync");
        swapTuples(arrays, fresh, std::make_index_sequence<sizeof...(Vectors)>{});
    }

    [[nodiscard]] LocalIndex startIndex() const { return LocalIndex(info_[0]); }
    [[nodiscard]] LocalIndex endIndex() const { return LocalIndex(info_[1]); }
    //! [ref: domain/domain.hpp:396] set the index of the last particle (used to increase the number of particles)
    void setEndIndex(const std::size_t i)
    {
        detail::check(Api::setEndIndex(d_, i), "Domain::setEndIndex");
        info_[1] = int64_t(i);
    }
    [[nodiscard]] LocalIndex nParticles() const { return endIndex() - startIndex(); }
    [[nodiscard]] LocalIndex nParticlesWithHalos() const { return LocalIndex(info_[2]); }
    const Box<T>& box() const { return box_; }
    //! [ref: domain/domain.hpp:404-406] first / one-past-last leaf of the search tree inside the assigned SFC range
    TreeNodeIndex startCell() const { return cellRange()[0]; }
    TreeNodeIndex endCell() const { return cellRange()[1]; }
    //! [ref: domain/domain.hpp:408] particle offsets of each leaf of the search tree (device memory, numLeafNodes + 1)
    DeviceSpan<const LocalIndex> layout() const
    {
        const void* p[10];
        detail::check(Api::treeView(d_, p), "Domain::layout");
        return {static_cast<const LocalIndex*>(p[2]), std::size_t(info_[3]) + 1};
    }
    //! [ref: domain/domain.hpp:411] accepted for source compatibility: this engine converges both trees in every sync
    void setTreeConv(bool flag) { convergeTrees_ = flag; }
    //! [ref: domain/domain.hpp:412] halo search radius = 2 * factor * max h of a leaf
    void setHaloFactor(float factor) { detail::check(Api::setHaloFactor(d_, factor), "Domain::setHaloFactor"); }
    /*! halos on more than one rank (call before the first sync).  false, the default: the fast path (halo set a subset
     *  of the reference's, sufficient for 2h searches).  true: the reference's procedure on the multi-rank focus tree
     *  [ref: domain/domain.hpp:217-243]; layout(), startIndex(), endIndex(), nParticlesWithHalos() and
     *  octreeProperties() then equal the reference's bit for bit */
    void setReferenceHalos(bool on) { detail::check(Api::setHaloMode(d_, on ? 1 : 0), "Domain::setReferenceHalos"); }
    //! [ref: domain/domain.hpp:413] over-allocation of the particle arrays when they have to grow
    void setGrowthAllocRate(float factor) { allocGrowthRate_ = factor; }

    //! [ref: domain/domain.hpp:424-437] search tree over assigned + halo particles (device pointers)
    OctreeNsView<T, KeyType> octreeProperties() const
    {
        const void* p[10];
        detail::check(Api::treeView(d_, p), "Domain::octreeProperties");
        TreeNodeIndex L = TreeNodeIndex(info_[3]);
        return {L,
                static_cast<const KeyType*>(p[3]),
                static_cast<const TreeNodeIndex*>(p[4]),
                static_cast<const TreeNodeIndex*>(p[5]),
                static_cast<const TreeNodeIndex*>(p[7]),
                static_cast<const KeyType*>(p[0]),
                static_cast<const LocalIndex*>(p[2]),
                static_cast<const Vec3<T>*>(p[8]),
                static_cast<const Vec3<T>*>(p[9])};
    }
    //! number of particles this rank sends / receives as halos per exchange
    std::size_t numHaloSend() const { return std::size_t(info_[6]); }
    std::size_t numHaloRecv() const { return std::size_t(info_[7]); }

    //! global (replicated) leaves, counts and rank boundaries on the host [ref: domain/domain.hpp:400 globalTree]
    void globalTree(std::vector<KeyType>& leaves, std::vector<unsigned>& counts, std::vector<KeyType>& assignment) const
    {
        const K *l, *a;
        const uint32_t* c;
        detail::check(Api::globalTree(d_, &l, &c, &a), "Domain::globalTree");
        std::size_t Lg = std::size_t(info_[5]);
        leaves.assign(reinterpret_cast<const KeyType*>(l), reinterpret_cast<const KeyType*>(l) + Lg + 1);
        counts.assign(c, c + Lg);
        assignment.assign(reinterpret_cast<const KeyType*>(a), reinterpret_cast<const KeyType*>(a) + numRanks_ + 1);
    }

    /*! neighbour search for the assigned particles with the domain's own tree
     *  [ref: findneighbors.hpp:167-178 findNeighbors]: neighbors[(i - startIndex) * ngmax + k], neighborsCount[i - startIndex];
     *  both vectors are resized here */
    void findNeighbors(const DeviceVector<T>& x,
                       const DeviceVector<T>& y,
                       const DeviceVector<T>& z,
                       const DeviceVector<T>& h,
                       unsigned ngmax,
                       DeviceVector<LocalIndex>& neighbors,
                       DeviceVector<unsigned>& neighborsCount) const
    {
        checkSizesEqual(info_[2], x, y, z, h);
        if (neighbors.size() < std::size_t(nParticles()) * ngmax) neighbors.resize(std::size_t(nParticles()) * ngmax);
        if (neighborsCount.size() < nParticles()) neighborsCount.resize(nParticles());
        detail::check(Api::findNeighbors(d_, x.data(), y.data(), z.data(), h.data(), ngmax, neighbors.data(),
                                         neighborsCount.data()),
                      "Domain::findNeighbors");
    }

    //! device time in ms of the stages of the last sync (see cstone_b200_domain.h)
    std::array<float, 10> stageMs() const
    {
        std::array<float, 10> ms{};
        detail::check(Api::stageMs(d_, ms.data()), "Domain::stageMs");
        return ms;
    }

private:
    template<class... Arrays>
    static void checkSizesEqual(std::size_t value, const Arrays&... arrays)
    {
        const std::size_t sizes[] = {arrays.size()..., value};
        for (std::size_t s : sizes)
            if (s != value) throw std::runtime_error("Domain sync: input array sizes are inconsistent\n");
    }

    template<class A, class B, std::size_t... I>
    static void swapTuples(A& a, B& b, std::index_sequence<I...>)
    {
        (std::get<I>(a).swap(std::get<I>(b)), ...);
    }

    std::array<TreeNodeIndex, 2> cellRange() const
    {
        int32_t c[2];
        detail::check(Api::cells(d_, c), "Domain::startCell");
        return {c[0], c[1]};
    }

    void refresh()
    {
        detail::check(Api::info(d_, info_), "Domain::info");
        typename Box<T>::CType b;
        detail::check(Api::box(d_, &b), "Domain::box");
        box_ = Box<T>(b);
    }

    int myRank_, numRanks_;
    Box<T> box_;
    cstone_b200_domain* d_ = nullptr;
    void* ownedComm_       = nullptr;
    bool convergeTrees_    = false;
    float allocGrowthRate_ = 1.05f;
    int64_t info_[8]       = {0, 0, 0, 0, 0, 0, 0, 0};
};

} // namespace cstone
