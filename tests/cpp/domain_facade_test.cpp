/*
 * C++ client of include/cstone_b200/domain.hpp, shaped like the reference's own integration test
 * [ref: test/integration_mpi/domain_nranks.cpp:58-160, domain_2ranks.cpp]: every rank draws the same global Gaussian
 * cloud, keeps its slice (NOT in SFC order), calls Domain::sync and checks
 *   - the sum of assigned particles (printed; the launcher adds them up),
 *   - keys sorted and consistent with x,y,z through the post-sync box,
 *   - properties travelled with their particles, halos filled by exchangeHalos,
 *   - neighbour counts of the assigned particles against an all-to-all search over the GLOBAL cloud
 *     (proves that the halos are sufficient),
 *   - a second sync with moved particles, the size-mismatch and the bucket-size exceptions.
 * Built with plain g++ (no nvcc, no CUDA headers):
 *   g++ -std=c++17 -O2 -ffp-contract=off -Iinclude tests/cpp/domain_facade_test.cpp -Lcornerstone_octree_b200 -lcstone_b200
 * usage: domain_facade_test <rank> <nranks> <idfile> [numParticles]
 */
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <random>
#include <thread>

#include "cstone_b200/domain.hpp"

using namespace cstone;

static int g_fail = 0;
#define EXPECT(cond)                                                                                                   \
    do                                                                                                                 \
    {                                                                                                                  \
        if (!(cond))                                                                                                   \
        {                                                                                                              \
            std::printf("FAILED %s:%d  %s\n", __FILE__, __LINE__, #cond);                                              \
            ++g_fail;                                                                                                  \
        }                                                                                                              \
    } while (0)

template<class T>
static void globalCloud(std::vector<T>& x, std::vector<T>& y, std::vector<T>& z, std::vector<T>& h, const Box<T>& box,
                        unsigned seed)
{
    std::mt19937 gen(seed);
    auto draw = [&gen](T lo, T hi)
    {
        std::normal_distribution<T> d((hi + lo) / 2, (hi - lo) / 5);
        return std::max(std::min(d(gen), hi), lo);
    };
    for (std::size_t i = 0; i < x.size(); ++i)
    {
        x[i] = draw(box.xmin(), box.xmax());
        y[i] = draw(box.ymin(), box.ymax());
        z[i] = draw(box.zmin(), box.zmax());
        // larger radii away from the centre where the density is low
        h[i] = T(0.05) + T(0.2) * (x[i] * x[i] + y[i] * y[i] + z[i] * z[i]);
    }
}

template<class T>
static T minImage(T d, T l, bool pbc)
{
    if (!pbc) return d;
    if (d > l / 2) return d - l;
    if (d < -l / 2) return d + l;
    return d;
}

template<class KeyType, class T>
static void run(int rank, int nRanks, const std::array<char, 128>& id, LocalIndex numGlobal, bool pbc, bool refHalos = false)
{
    Box<T> box = pbc ? Box<T>(-1, 1, BoundaryType::periodic) : Box<T>(-1, 1);
    Domain<KeyType, T, GpuTag> domain(rank, nRanks, 64, 8, 1.0f, box);
    if (nRanks > 1) domain.initNcclComm(id);
    if (refHalos) domain.setReferenceHalos(true); // the reference's halo procedure on the multi-rank focus tree
    std::fprintf(stderr, "rank %d: run(pbc %d, reference halos %d) starts\n", rank, int(pbc), int(refHalos));

    numGlobal = (numGlobal / nRanks) * nRanks;
    std::vector<T> xg(numGlobal), yg(numGlobal), zg(numGlobal), hg(numGlobal);
    globalCloud(xg, yg, zg, hg, box, 42);

    LocalIndex first = LocalIndex(uint64_t(rank) * numGlobal / nRanks), last = LocalIndex(uint64_t(rank + 1) * numGlobal / nRanks);
    std::vector<T> hx(xg.begin() + first, xg.begin() + last), hy(yg.begin() + first, yg.begin() + last),
        hz(zg.begin() + first, zg.begin() + last), hh(hg.begin() + first, hg.begin() + last);
    // two properties of different width that identify the particle: global id (u32) and x+y+z (T)
    std::vector<unsigned> hid(last - first);
    std::vector<double> hsum(last - first);
    for (LocalIndex i = first; i < last; ++i)
    {
        hid[i - first]  = i;
        hsum[i - first] = double(xg[i]) + double(yg[i]) + double(zg[i]);
    }

    DeviceVector<T> x(hx), y(hy), z(hz), h(hh), s1, s2;
    DeviceVector<unsigned> pid(hid), s3;
    DeviceVector<double> psum(hsum);
    DeviceVector<KeyType> keys;

    for (int step = 0; step < 2; ++step)
    {
        // a copy of one property in the layout BEFORE the sync: reapplySync must take it where sync takes pid
        DeviceVector<unsigned> pidLate(pid);
        std::fprintf(stderr, "rank %d: sync %d\n", rank, step);
        domain.sync(keys, x, y, z, h, std::tie(pid, psum), std::tie(s1, s2, s3));
        std::fprintf(stderr, "rank %d: sync %d done\n", rank, step);
        domain.reapplySync(std::tie(pidLate), s1, s2, s3);
        {
            auto viaSync = toHost(pid), viaReapply = toHost(pidLate);
            EXPECT(viaReapply.size() == viaSync.size());
            bool same = viaReapply.size() == viaSync.size();
            for (LocalIndex q = domain.startIndex(); same && q < domain.endIndex(); ++q)
                same = viaSync[q] == viaReapply[q];
            EXPECT(same);
        }
        domain.exchangeHalos(std::tie(pid, psum), s1, s2);

        LocalIndex n = domain.nParticlesWithHalos(), a = domain.startIndex(), b = domain.endIndex();
        EXPECT(keys.size() == n && x.size() == n && h.size() == n && pid.size() == n && psum.size() == n);
        EXPECT(a <= b && b <= n);
        auto k = toHost(keys);
        auto cx = toHost(x), cy = toHost(y), cz = toHost(z), ch = toHost(h);
        auto cid = toHost(pid);
        auto csum = toHost(psum);
        EXPECT(std::is_sorted(k.begin(), k.end()));
        // keys follow from the coordinates and the post-sync box
        DeviceVector<KeyType> keyCheck(n);
        // entries equal to the remove key are left alone by computeSfcKeys [ref: sfc/sfc_gpu.cu:40-45]: start from zero
        detail::check(cstone_b200_device_memset(keyCheck.data(), 0, n * sizeof(KeyType)), "memset");
        Box<T> nbox = domain.box();
        if constexpr (sizeof(KeyType) == 8 && sizeof(T) == 8)
            detail::check(cstone_b200_compute_sfc_keys_k64_f64(CSTONE_SFC_HILBERT, x.data(), y.data(), z.data(),
                                                               keyCheck.data(), n, nbox.c(), nullptr), "keys");
        else if constexpr (sizeof(KeyType) == 4 && sizeof(T) == 4)
            detail::check(cstone_b200_compute_sfc_keys_k32_f32(CSTONE_SFC_HILBERT, x.data(), y.data(), z.data(),
                                                               keyCheck.data(), n, nbox.c(), nullptr), "keys");
        detail::check(cstone_b200_device_synchronize(), "sync");
        EXPECT(toHost(keyCheck) == k);
        // every particle (assigned AND halo) still carries its own identity and coordinates
        std::size_t wrong = 0;
        for (LocalIndex i = 0; i < n; ++i)
        {
            unsigned g = cid[i];
            if (g >= numGlobal || cx[i] != xg[g] || cy[i] != yg[g] || cz[i] != zg[g] || ch[i] != hg[g] ||
                csum[i] != double(xg[g]) + double(yg[g]) + double(zg[g]))
                ++wrong;
        }
        EXPECT(wrong == 0);

        // neighbour counts against an all-to-all search over the global cloud
        const unsigned ngmax = 400;
        DeviceVector<LocalIndex> nb;
        DeviceVector<unsigned> nc;
        domain.findNeighbors(x, y, z, h, ngmax, nb, nc);
        auto hnc = toHost(nc);
        auto hnb = toHost(nb);
        std::size_t badCount = 0, badList = 0;
        uint64_t ncSum = 0;
        for (LocalIndex i = a; i < b; ++i)
        {
            T r2 = T(2) * ch[i] * T(2) * ch[i];
            std::vector<unsigned> ref;
            for (LocalIndex j = 0; j < numGlobal; ++j)
            {
                if (j == cid[i]) continue;
                T dx = minImage(cx[i] - xg[j], nbox.lx(), pbc), dy = minImage(cy[i] - yg[j], nbox.ly(), pbc),
                  dz = minImage(cz[i] - zg[j], nbox.lz(), pbc);
                if (dx * dx + dy * dy + dz * dz < r2) ref.push_back(j);
            }
            unsigned c = hnc[i - a];
            ncSum += c;
            if (c != ref.size()) { ++badCount; continue; }
            std::vector<unsigned> got(std::min(c, ngmax));
            for (unsigned q = 0; q < got.size(); ++q)
                got[q] = cid[hnb[std::size_t(i - a) * ngmax + q]];
            std::sort(got.begin(), got.end());
            if (c <= ngmax && got != ref) ++badList;
        }
        EXPECT(badCount == 0);
        EXPECT(badList == 0);
        auto view = domain.octreeProperties();
        EXPECT(view.numLeafNodes > 0 && view.searchExtFactor == 1.0f);
        // the reference's accessors: layout(), startCell(), endCell() [ref: domain/domain.hpp:404-408]
        {
            auto lay = domain.layout().toHost();
            EXPECT(lay.size() == std::size_t(view.numLeafNodes) + 1 && lay.front() == 0 && lay.back() == n);
            TreeNodeIndex c0 = domain.startCell(), c1 = domain.endCell();
            EXPECT(0 <= c0 && c0 < c1 && c1 <= view.numLeafNodes);
            EXPECT(lay[c0] == a && lay[c1] == b);
        }
        std::printf("RANK %d step %d assigned %u total %u halosend %zu halorecv %zu leaves %d ncsum %llu\n", rank, step,
                    b - a, n, domain.numHaloSend(), domain.numHaloRecv(), view.numLeafNodes,
                    (unsigned long long)ncSum);

        // move the global cloud a little (same on every rank) and write the assigned part back
        for (LocalIndex j = 0; j < numGlobal; ++j)
        {
            T s   = T(0.01) * T((j % 7) - 3);
            xg[j] = std::max(std::min(xg[j] + s, box.xmax()), box.xmin());
            yg[j] = std::max(std::min(yg[j] - s, box.ymax()), box.ymin());
        }
        for (LocalIndex i = 0; i < n; ++i)
        {
            cx[i] = xg[cid[i]], cy[i] = yg[cid[i]];
            csum[i] = double(xg[cid[i]]) + double(yg[cid[i]]) + double(zg[cid[i]]);
        }
        x = cx, y = cy, psum = csum;
    }

    if (refHalos)
    {
        // [ref: domain/domain.hpp:245-325 syncGrav] masses travel with the particles, keys stay sorted, sizes agree
        DeviceVector<T> m(std::vector<T>(x.size(), T(1) / T(numGlobal)));
        for (int step = 0; step < 2; ++step)
        {
            std::fprintf(stderr, "rank %d: syncGrav %d\n", rank, step);
            domain.syncGrav(keys, x, y, z, h, m, std::tie(pid, psum), std::tie(s1, s2, s3));
            LocalIndex n = domain.nParticlesWithHalos(), a = domain.startIndex(), b = domain.endIndex();
            EXPECT(keys.size() == n && x.size() == n && m.size() == n && pid.size() == n);
            auto k  = toHost(keys);
            auto cm = toHost(m);
            EXPECT(std::is_sorted(k.begin(), k.end()));
            bool massOk = true;
            for (LocalIndex q = a; q < b; ++q)
                massOk = massOk && cm[q] == T(1) / T(numGlobal);
            EXPECT(massOk);
            auto view = domain.octreeProperties();
            EXPECT(view.numLeafNodes > 0);
        }
    }

    // array sizes must not change between calls [ref: domain/domain.hpp:130-133]
    if (nRanks == 1)
    {
        bool thrown = false;
        DeviceVector<T> shortX(x.size() - 1), shortY(x.size() - 1), shortZ(x.size() - 1), shortH(x.size() - 1);
        try
        {
            domain.sync(keys, shortX, shortY, shortZ, shortH, std::tuple<>{}, std::tie(s1, s2, s3));
        }
        catch (std::runtime_error&)
        {
            thrown = true;
        }
        EXPECT(thrown);
        thrown = false;
        try
        {
            Domain<KeyType, T, GpuTag> bad(0, 1, 8, 64, 1.0f);
        }
        catch (std::runtime_error&)
        {
            thrown = true;
        }
        EXPECT(thrown);
    }
}

//! rank 0 draws a NCCL id and leaves it in a file, the others wait for it
static std::array<char, 128> shareId(int rank, int nRanks, const std::string& idFile)
{
    std::array<char, 128> id{};
    if (nRanks < 2) return id;
    if (rank == 0)
    {
        id = Domain<uint64_t, double, GpuTag>::ncclUniqueId();
        std::ofstream(idFile + ".tmp", std::ios::binary).write(id.data(), 128);
        std::rename((idFile + ".tmp").c_str(), idFile.c_str());
    }
    else
    {
        for (int t = 0; t < 600; ++t)
        {
            std::ifstream f(idFile, std::ios::binary);
            if (f && f.read(id.data(), 128)) break;
            std::this_thread::sleep_for(std::chrono::milliseconds(100));
        }
    }
    return id;
}

int main(int argc, char** argv)
{
    if (argc < 4)
    {
        std::printf("usage: %s rank nranks idfile [numParticles]\n", argv[0]);
        return 2;
    }
    int rank = std::atoi(argv[1]), nRanks = std::atoi(argv[2]);
    std::string idFile = argv[3];
    LocalIndex n       = argc > 4 ? LocalIndex(std::atol(argv[4])) : 4000;

    // one process per GPU: LOCAL device = rank (the launcher may narrow CUDA_VISIBLE_DEVICES instead)
    if (const char* dev = std::getenv("CSB_TEST_DEVICE")) { detail::check(cstone_b200_set_device(std::atoi(dev)), "set_device"); }
    else if (nRanks > 1) { detail::check(cstone_b200_set_device(rank), "set_device"); }

    std::array<char, 128> id = shareId(rank, nRanks, idFile);
    run<uint64_t, double>(rank, nRanks, id, n, false);
    if (nRanks > 1)
    {
        // the same checks with the reference's halo procedure, then syncGrav on that domain (a second communicator)
        std::array<char, 128> id2 = shareId(rank, nRanks, idFile + ".let");
        run<uint64_t, double>(rank, nRanks, id2, n, false, true);
    }
    if (nRanks == 1)
    {
        run<uint64_t, double>(rank, nRanks, id, n, false, true); // focus tree of the reference halo mode + syncGrav, one rank
        run<uint64_t, double>(rank, nRanks, id, n, true);
        run<uint32_t, float>(rank, nRanks, id, n / 2, false);
    }
    std::printf(g_fail ? "RANK %d FACADE_FAILED %d\n" : "RANK %d FACADE_OK %d\n", rank, g_fail);
    return g_fail ? 1 : 0;
}
