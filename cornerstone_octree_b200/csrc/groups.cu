/*
 * groups.cu — n2 (first part): target groups for group-wise tree traversal.
 *
 * Replaces computeFixedGroups / computeGroupSplits [ref: traversal/groups_gpu.h:60-88, traversal/groups_gpu.cu:39-149,
 * device functions traversal/groups_gpu.cuh:54-230 findSplits, makeSplits, groupSplitsKernel].  The reference has no
 * CPU version of these; the behaviour is pinned on its CUDA unit test vectors (test/unit_cuda/traversal/groups.cu).
 *
 * computeGroupSplits: the particle range [first, last) is cut into fixed groups of group_size (32 or 64) consecutive
 * particles; each fixed group is cut again wherever two CONSECUTIVE particles are further apart (in box-normalised
 * coordinates) than tol_factor * edge length of the leaf of the group's first particle.
 *
 * Reference pipeline: split kernel (a 21-step leaf search per particle), scan, D2H, length kernel (popcount loops),
 * second scan, H2D of `last`.
 * Here: the leaf level of every group is scattered from the leaf side (groupLeafLevelsKernel, no search), then the
 * split kernel (one warp per fixed group, coalesced independent loads, one ballot per 32 particles), one scan of the split
 * counts, and a start kernel that writes group starts straight from the bit positions — the lengths of a fixed
 * group always add up to group_size, so start(g, j) = first + g*group_size + (position of the j-th split bit) + 1
 * and the second scan is not needed.  Algorithmic bytes: 3 sizeof(Tc) per particle read, 4 B per group written.
 *
 * Arithmetic follows the reference expression by expression without contraction: p = x * inverse length,
 * d = p_next - p, d2 = (dx dx + dy dy) + dz dz, split iff d2 > (cbrt(8 s^3) * tol)^2 with s = half edge of the leaf
 * in the unit box (Th arithmetic, like the reference's Box<T> unitBox).
 */
#include "scan.cuh"

namespace csb
{

constexpr int kGroupThreads = 256;

__global__ void __launch_bounds__(kGroupThreads)
    fixedGroupsKernel(uint32_t first, uint32_t last, uint32_t groupSize, uint32_t* __restrict__ groups, uint32_t numGroups)
{
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g <= numGroups; g += gridDim.x * blockDim.x)
        groups[g] = (g == numGroups) ? last : first + g * groupSize;
}

/*! Leaf level per fixed group, computed from the LEAF side: the reference takes, for every body of the first warp-width
 *  of a group, the leaf upper_bound(layout, layout + numLeaves, body) - 1 and keeps the smallest one.  Leaf i holds the
 *  bodies [layout[i], layout[i+1]) (the last leaf: everything from layout[i] on), so it is a maximum of the leaf levels
 *  scattered to the groups whose first 32 bodies a leaf touches: one coalesced pass over the leaf arrays with an
 *  atomicMax per (leaf, group) pair, and no search at all on the particle side.  levels[] starts at 0 (root, volume 1,
 *  the reference's initial value).  Leaves spanning more than 32 groups are spread over the warp. */
template<class K>
__global__ void __launch_bounds__(kGroupThreads)
    groupLeafLevelsKernel(uint32_t first, uint32_t last, uint32_t groupSize, const K* __restrict__ leaves, int numLeaves,
                          const uint32_t* __restrict__ layout, uint32_t* __restrict__ levels)
{
    const uint32_t lane = laneId();
    for (int base = int((blockIdx.x * blockDim.x + threadIdx.x) & ~31u); base < numLeaves; base += int(gridDim.x * blockDim.x))
    {
        const int i = base + int(lane);
        uint32_t sb = 0, eb = 0, level = 0;
        if (i < numLeaves)
        {
            sb    = (i == 0) ? 0u : layout[i];
            eb    = (i == numLeaves - 1) ? 0xffffffffu : layout[i + 1];
            sb    = max(sb, first);
            eb    = min(eb, last);
            level = treeLevel<K>(K(leaves[i + 1] - leaves[i]));
        }
        const bool active   = sb < eb && level > 0;
        const uint32_t gLo  = active ? (sb - first) / groupSize : 1u;
        const uint32_t gHi  = active ? (eb - 1u - first) / groupSize : 0u;
        const bool big      = active && gHi - gLo > 32u;
        if (active && !big)
        {
            for (uint32_t g = gLo; g <= gHi; ++g)
            {
                const uint32_t lo = first + g * groupSize; // the first warp-width of the group
                if (max(sb, lo) < min(eb, lo + 32u)) atomicMax(&levels[g], level);
            }
        }
        for (unsigned todo = __ballot_sync(0xffffffffu, big); todo; todo &= todo - 1u)
        {
            const int src      = __ffs(int(todo)) - 1;
            const uint32_t lv  = __shfl_sync(0xffffffffu, level, src);
            const uint32_t g0  = __shfl_sync(0xffffffffu, gLo, src), g1 = __shfl_sync(0xffffffffu, gHi, src);
            const uint32_t s0  = __shfl_sync(0xffffffffu, sb, src), e0 = __shfl_sync(0xffffffffu, eb, src);
            for (uint32_t g = g0 + lane; g <= g1; g += 32u)
            {
                const uint32_t lo = first + g * groupSize;
                if (max(s0, lo) < min(e0, lo + 32u)) atomicMax(&levels[g], lv);
            }
        }
    }
}

template<int NWT, class K, class Tc, class Th>
__global__ void __launch_bounds__(kGroupThreads)
    groupSplitsKernel(uint32_t first, uint32_t last, const Tc* __restrict__ x, const Tc* __restrict__ y,
                      const Tc* __restrict__ z, const uint32_t* __restrict__ levels, Tc ilx, Tc ily, Tc ilz,
                      float tolFactor, uint32_t* __restrict__ masks, uint32_t* __restrict__ numSplits, uint32_t numFixed)
{
    constexpr uint32_t groupSize = 32u * NWT;
    const uint32_t lane          = laneId();
    for (uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; g < numFixed; g += (gridDim.x * blockDim.x) >> 5)
    {
        uint32_t body[NWT];
#pragma unroll
        for (int k = 0; k < NWT; ++k)
            body[k] = min(first + g * groupSize + uint32_t(k) * 32u + lane, last - 1u);

        // volume of the smallest leaf under the group's first warp-width of bodies (groupLeafLevelsKernel)
        const unsigned level = levels[g];
        const Th half = Th(int(1u << (KeyTraits<K>::maxLevel - level))) * (Th(0.5) / Th(1u << KeyTraits<K>::maxLevel));
        const Th vol  = Th(8) * half * half * half;
        Tc distCrit = Tc(cbrt(vol) * tolFactor); // Th * float, evaluated in Th
        Tc crit2    = distCrit * distCrit;

        Tc px[NWT], py[NWT], pz[NWT];
#pragma unroll
        for (int k = 0; k < NWT; ++k)
            px[k] = x[body[k]] * ilx, py[k] = y[body[k]] * ily, pz[k] = z[body[k]] * ilz;

        int count = 1;
#pragma unroll
        for (int k = 0; k < NWT; ++k)
        {
            Tc nx = __shfl_down_sync(0xffffffffu, px[k], 1);
            Tc ny = __shfl_down_sync(0xffffffffu, py[k], 1);
            Tc nz = __shfl_down_sync(0xffffffffu, pz[k], 1);
            if (k + 1 < NWT)
            {
                Tc sx = __shfl_sync(0xffffffffu, px[(k + 1) % NWT], 0);
                Tc sy = __shfl_sync(0xffffffffu, py[(k + 1) % NWT], 0);
                Tc sz = __shfl_sync(0xffffffffu, pz[(k + 1) % NWT], 0);
                if (lane == 31u) nx = sx, ny = sy, nz = sz;
            }
            Tc dx = nx - px[k], dy = ny - py[k], dz = nz - pz[k];
            Tc d2 = dx * dx + dy * dy + dz * dz;
            uint32_t m = __ballot_sync(0xffffffffu, d2 > crit2);
            if (lane == 0) masks[size_t(g) * NWT + k] = m;
            count += __popc(m);
        }
        if (lane == 0) numSplits[g] = uint32_t(count);
    }
}

template<int NWT>
__global__ void __launch_bounds__(kGroupThreads)
    groupStartsKernel(const uint32_t* __restrict__ masks, const uint32_t* __restrict__ offsets, uint32_t numFixed,
                      uint32_t first, uint32_t last, uint32_t numGroups, uint32_t* __restrict__ groups)
{
    constexpr uint32_t groupSize = 32u * NWT;
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < numFixed; g += gridDim.x * blockDim.x)
    {
        uint32_t base = first + g * groupSize;
        uint32_t* out = groups + offsets[g];
        *out++        = base;
#pragma unroll
        for (int k = 0; k < NWT; ++k)
        {
            uint32_t m = masks[size_t(g) * NWT + k];
            while (m)
            {
                int b  = __ffs(int(m)) - 1;
                *out++ = base + uint32_t(k) * 32u + uint32_t(b) + 1u;
                m &= m - 1u;
            }
        }
        if (g + 1 == numFixed) groups[numGroups] = last;
    }
}

inline size_t align128(size_t b) { return (b + 127) & ~size_t(127); }

struct GroupTemp
{
    uint32_t *masks, *numSplits, *offsets, *levels, *total;
    void* scan;
    size_t bytes;
};

inline GroupTemp carveGroupTemp(void* temp, size_t numFixed, int nwt)
{
    char* p = reinterpret_cast<char*>(temp);
    GroupTemp t;
    t.masks     = reinterpret_cast<uint32_t*>(p), p += align128(numFixed * nwt * 4);
    t.numSplits = reinterpret_cast<uint32_t*>(p), p += align128(numFixed * 4);
    t.offsets   = reinterpret_cast<uint32_t*>(p), p += align128(numFixed * 4);
    t.levels    = reinterpret_cast<uint32_t*>(p), p += align128(numFixed * 4);
    t.total     = reinterpret_cast<uint32_t*>(p), p += 128;
    t.scan      = p, p += align128(scanTempBytes(numFixed));
    t.bytes     = size_t(p - reinterpret_cast<char*>(temp));
    return t;
}

template<int NWT, class K, class Tc, class Th>
int groupSplitsImpl(uint32_t first, uint32_t last, const Tc* x, const Tc* y, const Tc* z, const K* leaves,
                    int numLeaves, const uint32_t* layout, const typename BoxT<Tc>::type* box, float tolFactor,
                    uint32_t* groups, uint64_t capacity, uint32_t* numGroupsHost, void* temp, uint64_t tempBytes,
                    cudaStream_t s)
{
    uint32_t numFixed = (last - first + 32u * NWT - 1u) / (32u * NWT);
    GroupTemp t       = carveGroupTemp(temp, numFixed, NWT);
    if (tempBytes < t.bytes) return CSTONE_E_TEMP_TOO_SMALL;

    CSB_CUDA(cudaMemsetAsync(t.levels, 0, size_t(numFixed) * 4, s));
    groupLeafLevelsKernel<K><<<gridFor(size_t(numLeaves), kGroupThreads, 8), kGroupThreads, 0, s>>>(
        first, last, 32u * NWT, leaves, numLeaves, layout, t.levels);
    CSB_LAUNCHED();
    groupSplitsKernel<NWT, K, Tc, Th><<<gridFor(size_t(numFixed) * 32, kGroupThreads, 8), kGroupThreads, 0, s>>>(
        first, last, x, y, z, t.levels, Tc(box->inv_lengths[0]), Tc(box->inv_lengths[1]), Tc(box->inv_lengths[2]),
        tolFactor, t.masks, t.numSplits, numFixed);
    CSB_LAUNCHED();
    int st = launchScan<false>(LoadPlain{t.numSplits}, t.offsets, numFixed, 0u, t.scan, t.total, s);
    if (st) return st;
    uint32_t numGroups = 0;
    CSB_CUDA(cudaMemcpyAsync(&numGroups, t.total, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    *numGroupsHost = numGroups;
    if (uint64_t(numGroups) + 1 > capacity) return CSTONE_E_CAPACITY;

    groupStartsKernel<NWT><<<gridFor(size_t(numFixed), kGroupThreads, 8), kGroupThreads, 0, s>>>(
        t.masks, t.offsets, numFixed, first, last, numGroups, groups);
    CSB_LAUNCHED();
    return 0;
}

template<class K, class Tc, class Th>
int groupSplits(uint32_t first, uint32_t last, const Tc* x, const Tc* y, const Tc* z, const Th* /*h*/, const K* leaves,
                int numLeaves, const uint32_t* layout, const typename BoxT<Tc>::type* box, uint32_t groupSize,
                float tolFactor, uint32_t* groups, uint64_t capacity, uint32_t* numGroupsHost, void* temp,
                uint64_t tempBytes, void* stream)
{
    if (!numGroupsHost) return CSTONE_E_INVALID_ARGUMENT;
    *numGroupsHost = 0;
    if (groupSize != 32 && groupSize != 64) return CSTONE_E_INVALID_ARGUMENT; // [ref: groups_gpu.cu:137] throws
    if (last <= first)
    {
        if (capacity < 1 || !groups) return CSTONE_E_CAPACITY;
        CSB_CUDA(cudaMemcpyAsync(groups, &last, sizeof(uint32_t), cudaMemcpyHostToDevice, asStream(stream)));
        CSB_CUDA(cudaStreamSynchronize(asStream(stream)));
        return 0;
    }
    if (!x || !y || !z || !leaves || !layout || !box || !groups || !temp || numLeaves <= 0)
        return CSTONE_E_INVALID_ARGUMENT;
    if (groupSize == 32)
        return groupSplitsImpl<1, K, Tc, Th>(first, last, x, y, z, leaves, numLeaves, layout, box, tolFactor, groups,
                                             capacity, numGroupsHost, temp, tempBytes, asStream(stream));
    return groupSplitsImpl<2, K, Tc, Th>(first, last, x, y, z, leaves, numLeaves, layout, box, tolFactor, groups,
                                         capacity, numGroupsHost, temp, tempBytes, asStream(stream));
}

} // namespace csb

extern "C" {

int cstone_b200_compute_fixed_groups(uint32_t first, uint32_t last, uint32_t groupSize, uint32_t* groups, void* stream)
{
    if (!groups || groupSize == 0 || last < first) return CSTONE_E_INVALID_ARGUMENT;
    uint32_t numGroups = (last - first + groupSize - 1) / groupSize;
    csb::fixedGroupsKernel<<<csb::gridFor(size_t(numGroups) + 1, csb::kGroupThreads, 8), csb::kGroupThreads, 0,
                             csb::asStream(stream)>>>(first, last, groupSize, groups, numGroups);
    CSB_LAUNCHED();
    return 0;
}

uint64_t cstone_b200_group_splits_temp_bytes(uint32_t numParticles, uint32_t groupSize)
{
    if (groupSize != 32 && groupSize != 64) return 0;
    size_t numFixed = (size_t(numParticles) + groupSize - 1) / groupSize;
    return csb::carveGroupTemp(nullptr, numFixed, int(groupSize / 32)).bytes;
}

#define CSB_GROUPS_API(KS, K, TS, Tc, Th)                                                                              \
    int cstone_b200_compute_group_splits_##KS##_##TS(                                                                  \
        uint32_t first, uint32_t last, const Tc* x, const Tc* y, const Tc* z, const Th* h, const K* leaves,            \
        int32_t numLeaves, const uint32_t* layout, const csb::BoxT<Tc>::type* box, uint32_t groupSize,                 \
        float tolFactor, uint32_t* groups, uint64_t capacity, uint32_t* numGroupsHost, void* temp, uint64_t tempBytes, \
        void* stream)                                                                                                  \
    {                                                                                                                  \
        return csb::groupSplits<K, Tc, Th>(first, last, x, y, z, h, leaves, numLeaves, layout, box, groupSize,         \
                                           tolFactor, groups, capacity, numGroupsHost, temp, tempBytes, stream);       \
    }
CSB_GROUPS_API(k32, uint32_t, f64_f64, double, double)
CSB_GROUPS_API(k32, uint32_t, f64_f32, double, float)
CSB_GROUPS_API(k32, uint32_t, f32_f32, float, float)
CSB_GROUPS_API(k64, uint64_t, f64_f64, double, double)
CSB_GROUPS_API(k64, uint64_t, f64_f32, double, float)
CSB_GROUPS_API(k64, uint64_t, f32_f32, float, float)

} // extern "C"
