/*
 * radix_sort.cu — a4: stable LSD radix sort of (key, u32 value) pairs, hand-written "onesweep".
 *
 * Replaces sortByKeyGpu / sortByKeyTempStorage / sequenceGpu [ref: primitives/primitives_gpu.h:86-110,
 * primitives/primitives_gpu.cu:285-392 (cub::DeviceRadixSort::SortPairs, thrust::sort_by_key)] as used by
 * GpuSfcSorter [ref: primitives/gather.cuh:80-108] and buildOctreeGpu [ref: tree/octree_gpu.cu:161].
 * Result = the unique stable ascending order (same as the CPU std::stable_sort path, primitives/gather.hpp:58-90).
 *
 * Algorithm (8-bit digits, full key width: 4 passes for 32-bit, 8 for 64-bit keys)
 *   1. one histogram kernel over the keys fills the digit counts of ALL passes,
 *   2. a tiny kernel turns them into exclusive bin offsets,
 *   3. per pass ONE kernel ("onesweep"): every block takes a tile (atomic ticket), ranks its keys by digit with
 *      warp-level match_any/popc (stable), publishes the per-digit tile counts and resolves the counts of all
 *      earlier tiles by decoupled look-back (one status word per tile and digit), reorders the tile through
 *      shared memory and writes keys and values in digit-contiguous runs.
 *   Per pass each key and value is read once and written once: 2*(sizeof(K)+4) bytes per element.
 *   With values_are_iota the first pass synthesises the value from the element index (no value read).
 */
#include "common.cuh"

namespace csb
{

constexpr int kRadixBits    = 8;
constexpr int kRadix        = 1 << kRadixBits;
constexpr int kSortThreads  = 256; // == kRadix: thread d owns digit d in the look-back
constexpr int kSortWarps    = kSortThreads / 32;
#ifndef RS_ITEMS
#define RS_ITEMS 20 // measured at 32 M pairs (tools/sort_variants.sh): u64 12/16/20/24 items = 2.14/2.07/1.98/2.00 ms
#endif
#ifndef RS_MINBLOCKS
#define RS_MINBLOCKS(K) (sizeof(K) == 8 ? 2 : 3) // 91 KB (u64) / 71 KB (u32) of shared memory per block
#endif
constexpr int kSortItems    = RS_ITEMS;
constexpr int kLookBack     = 4;  // predecessor tiles requested before the reorder
constexpr int kLookBackWide = 16; // ... and per round afterwards
constexpr int kSortTile     = kSortThreads * kSortItems;
constexpr uint32_t kStFlagShift     = 30;
constexpr uint32_t kStValueMask     = (1u << kStFlagShift) - 1u;
constexpr uint32_t kStAggregate     = 1u << kStFlagShift;
constexpr uint32_t kStInclusive     = 2u << kStFlagShift;

template<class K>
__host__ __device__ constexpr int numPasses()
{
    return (KeyTraits<K>::keyBits + 1 + kRadixBits - 1) / kRadixBits; // +1: keys may carry the bit above keyBits
}
// 32-bit: 31 bits -> 4 passes; 64-bit: 64 bits -> 8 passes (removeKey = 2^63 and placeholder-bit keys use bit 63)

struct SortTemp
{
    uint32_t* hist;    // [passes][256] counts, then exclusive offsets
    uint32_t* tickets; // [passes] (stride 32 words)
    uint32_t* status;  // [passes][numTiles][256]
    size_t bytes;
};

inline size_t sortNumTiles(size_t n) { return (n + kSortTile - 1) / kSortTile; }

template<class K>
uint64_t sortTempBytes(uint64_t n)
{
    uint64_t P = numPasses<K>();
    return P * kRadix * 4 + P * 32 * 4 + P * sortNumTiles(n) * kRadix * 4 + 256;
}

template<class K>
SortTemp carveSortTemp(void* temp, size_t n)
{
    constexpr int P = numPasses<K>();
    SortTemp t;
    t.hist    = reinterpret_cast<uint32_t*>(temp);
    t.tickets = t.hist + P * kRadix;
    t.status  = t.tickets + P * 32;
    t.bytes   = size_t(P) * kRadix * 4 + size_t(P) * 32 * 4 + size_t(P) * sortNumTiles(n) * kRadix * 4;
    return t;
}

__device__ inline uint32_t ldRelaxed(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ inline void stRelaxed(uint32_t* p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

/* Lanes of the warp that hold the same 8-bit digit.  Eight ballots instead of match.any: MATCH runs on the ADU pipe at
 * ~1 warp instruction per 64 cycles per SM (measured: 72 % ADU utilisation was the limiter of the whole pass), a
 * VOTE + LOP3 pair per bit is an order of magnitude cheaper. */
template<int Bit>
__device__ inline unsigned sameBitLanes(unsigned d)
{
    // lanes whose bit `Bit` of d equals mine: ballot of the bit, inverted where my bit is clear (LOP3.P, VOTE, @!p LOP3)
    unsigned m;
    asm("{\n\t.reg .pred p;\n\t.reg .b32 t;\n\tand.b32 t, %1, %2;\n\tsetp.ne.u32 p, t, 0;\n\t"
        "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t@!p not.b32 %0, %0;\n\t}"
        : "=r"(m)
        : "r"(d), "n"(1u << Bit));
    return m;
}
__device__ inline unsigned matchDigit(unsigned d)
{
    static_assert(kRadixBits == 8, "eight ballots");
    return (sameBitLanes<0>(d) & sameBitLanes<1>(d) & sameBitLanes<2>(d)) &
           (sameBitLanes<3>(d) & sameBitLanes<4>(d) & sameBitLanes<5>(d)) & (sameBitLanes<6>(d) & sameBitLanes<7>(d));
}

// ------------------------------------------------------------------------------------------------
// 1. digit histograms of all passes in one sweep over the keys
// ------------------------------------------------------------------------------------------------
template<class K>
__global__ void __launch_bounds__(512) histogramKernel(const K* __restrict__ keys, size_t n, uint32_t* __restrict__ hist)
{
    constexpr int P = numPasses<K>();
    __shared__ uint32_t sHist[P * kRadix];
    for (int i = threadIdx.x; i < P * kRadix; i += blockDim.x)
        sHist[i] = 0;
    __syncthreads();

    /* SFC keys of a time-stepping code arrive almost sorted: the upper digits of 32 consecutive keys are then
     * identical and a plain atomicAdd would serialise 32 ways on one address.  One vote on the upper half of the
     * key decides for P/2 digits at once: if it is warp-uniform, lane p adds the warp's count to digit p.
     * (Full warps only; the ragged tail takes the plain path.) */
    auto addKey = [&](K key, bool fullWarp)
    {
        constexpr int PH    = P / 2;
        const unsigned top  = unsigned(key >> (PH * kRadixBits));
        bool uniform        = false;
        if (fullWarp) uniform = __all_sync(0xffffffffu, top == __shfl_sync(0xffffffffu, top, 0));
#pragma unroll
        for (int p = 0; p < PH; ++p)
            atomicAdd(&sHist[p * kRadix + (unsigned(key >> (p * kRadixBits)) & (kRadix - 1))], 1u);
        if (uniform)
        {
            const unsigned l = threadIdx.x & 31u;
            if (l < unsigned(P - PH))
                atomicAdd(&sHist[(PH + l) * kRadix + ((top >> (l * kRadixBits)) & (kRadix - 1))], 32u);
        }
        else
        {
#pragma unroll
            for (int p = PH; p < P; ++p)
                atomicAdd(&sHist[p * kRadix + (unsigned(key >> (p * kRadixBits)) & (kRadix - 1))], 1u);
        }
    };
    // kHistUnroll independent coalesced loads per thread are in flight before the first shared-memory atomic
    constexpr int kHistUnroll = 4;
    const size_t chunk        = size_t(blockDim.x) * kHistUnroll;
    const size_t stride       = size_t(gridDim.x) * chunk;
    size_t base               = size_t(blockIdx.x) * chunk;
    for (; base + chunk <= n; base += stride)
    {
        K key[kHistUnroll];
#pragma unroll
        for (int k = 0; k < kHistUnroll; ++k)
            key[k] = keys[base + size_t(k) * blockDim.x + threadIdx.x];
#pragma unroll
        for (int k = 0; k < kHistUnroll; ++k)
            addKey(key[k], true);
    }
    if (base < n) // the ragged last chunk (one block only)
    {
        for (size_t i = base + threadIdx.x; i < n; i += blockDim.x)
            addKey(keys[i], __activemask() == 0xffffffffu);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < P * kRadix; i += blockDim.x)
    {
        uint32_t c = sHist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// 2. exclusive scan of each pass's 256 bins (one block of 256 threads per pass)
/* Also notes in tickets[32 * pass + 1] whether every key carries the same digit in this pass (one bin holds all n keys):
 * such a pass is the identity permutation.  A pair of passes (2q, 2q+1), q >= 1, that are both uniform is skipped
 * altogether (the ping-pong buffers keep their roles); a single uniform pass copies its tile straight across.  The
 * Warren-Salmon prefixes buildOctree sorts occupy 3 * depth + 1 low bits, so 4 - 5 of their 8 passes fall away. */
__global__ void __launch_bounds__(kRadix) binOffsetsKernel(uint32_t* __restrict__ hist, uint32_t* __restrict__ tickets,
                                                           uint32_t n)
{
    __shared__ uint32_t sWarp[kRadix / 32];
    uint32_t* h   = hist + blockIdx.x * kRadix;
    uint32_t c    = h[threadIdx.x];
    if (c == n) tickets[blockIdx.x * 32 + 1] = 1u;
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (int(laneId()) >= o) incl += t;
    }
    if (laneId() == 31) sWarp[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (int w = 0; w < int(threadIdx.x >> 5); ++w)
        off += sWarp[w];
    h[threadIdx.x] = off + incl - c;
}

// ------------------------------------------------------------------------------------------------
// 3. one onesweep pass
// ------------------------------------------------------------------------------------------------
template<class K>
struct SortSmem
{
    K keys[kSortTile];
    uint32_t vals[kSortTile];
    uint32_t stage[kSortTile];            // values in load order: cp.async landing zone (no registers while ranking)
    uint32_t warpHist[kSortWarps][kRadix]; // per warp digit counts, then first tile-local slot of (warp, digit)
    uint32_t globalBase[kRadix];           // global position of local slot 0 of each digit, minus its tile-local start
    uint32_t warpSums[kSortWarps];
    uint32_t tile;
};

__device__ inline void cpAsync4(void* smemDst, const void* src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(uint32_t(__cvta_generic_to_shared(smemDst))), "l"(src)
                 : "memory");
}

__device__ inline void cpAsync16(void* smemDst, const void* src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(uint32_t(__cvta_generic_to_shared(smemDst))), "l"(src)
                 : "memory");
}

/* One round of the look-back: consume W status words (tiles t, t-1, ...) in order up to the first tile that is not
 * ready (flag 0) or carries an inclusive prefix (flag 2).  Returns true when the inclusive prefix was reached. */
template<int W>
__device__ inline bool consumeWindow(const uint32_t (&st)[W], uint32_t& excl, int& t)
{
    bool open = true, done = false;
#pragma unroll
    for (int k = 0; k < W; ++k)
    {
        const uint32_t flag = st[k] >> kStFlagShift;
        open                = open && flag != 0u;
        if (open)
        {
            excl += st[k] & kStValueMask;
            --t;
            if (flag == 2u) done = true, open = false;
        }
    }
    return done;
}

template<class K, bool Iota>
__global__ void __launch_bounds__(kSortThreads, RS_MINBLOCKS(K)) onesweepKernel(const K* __restrict__ keysIn,
                                                                   const uint32_t* __restrict__ valsIn,
                                                                   K* __restrict__ keysOut,
                                                                   uint32_t* __restrict__ valsOut,
                                                                   uint32_t n,
                                                                   int shift,
                                                                   const uint32_t* __restrict__ binOffsets,
                                                                   uint32_t* __restrict__ ticket,
                                                                   uint32_t* __restrict__ status)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    SortSmem<K>& sm = *reinterpret_cast<SortSmem<K>*>(smemRaw);

    const unsigned tid  = threadIdx.x;
    const unsigned lane = tid & 31u;
    const unsigned warp = tid >> 5;

    /* uniform digit (see binOffsetsKernel): `ticket` is tickets + 32 * pass, the flag of the partner pass of the pair
     * sits 32 words before (odd pass) or after (even pass) */
    if (ticket[1])
    {
        const int pass = shift / kRadixBits;
        if (pass >= 2 && ((pass & 1) ? ticket[1 - 32] : ticket[1 + 32])) return; // both passes of the pair: nothing moves
        const uint32_t base = blockIdx.x * uint32_t(kSortTile);
        for (uint32_t i = base + tid; i < min(base + uint32_t(kSortTile), n); i += kSortThreads)
        {
            keysOut[i] = keysIn[i];
            valsOut[i] = Iota ? i : valsIn[i];
        }
        return;
    }

    if (tid == 0) sm.tile = atomicAdd(ticket, 1u);
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w)
        sm.warpHist[w][tid] = 0;
    __syncthreads();
    const uint32_t tile     = sm.tile;
    const uint32_t tileBase = tile * uint32_t(kSortTile);
    const uint32_t warpOff  = warp * (32u * kSortItems) + lane; // tile-local index of item 0 of this thread
    const bool fullTile     = tileBase + uint32_t(kSortTile) <= n;

    // ---- load (warp-striped: item i of lane l is element tileBase + warpOff + 32 i); order = (warp, i, lane).
    //      Values travel straight to shared memory (cp.async) and are picked up by the same thread at the reorder.
    K key[kSortItems];
    if (fullTile)
    {
        const K* src = keysIn + tileBase + warpOff;
#pragma unroll
        for (int i = 0; i < kSortItems; ++i)
            key[i] = src[32 * i];
        if (!Iota)
        {
            if ((reinterpret_cast<uintptr_t>(valsIn) & 15u) == 0)
            {
                // 16 bytes per copy, tile order; the reorder reads them after cp.async.wait_all + the block barrier
#pragma unroll
                for (int r = 0; r < kSortItems / 4; ++r)
                    cpAsync16(&sm.stage[4u * (tid + r * kSortThreads)], valsIn + tileBase + 4u * (tid + r * kSortThreads));
            }
            else
            {
                const uint32_t* vsrc = valsIn + tileBase + warpOff;
#pragma unroll
                for (int i = 0; i < kSortItems; ++i)
                    cpAsync4(&sm.stage[warpOff + 32u * i], vsrc + 32 * i);
            }
        }
    }
    else
    {
#pragma unroll
        for (int i = 0; i < kSortItems; ++i)
        {
            const uint32_t idx = tileBase + warpOff + 32u * i;
            key[i]             = (idx < n) ? keysIn[idx] : K(~K(0)); // padding sorts behind everything, never written
            if (!Iota && idx < n) cpAsync4(&sm.stage[warpOff + 32u * i], valsIn + idx);
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");

    // ---- stable ranking inside the warp
    uint16_t rank[kSortItems];
    uint32_t* myHist       = sm.warpHist[warp];
    const unsigned lanesLt = laneMaskLt();
#pragma unroll
    for (int i = 0; i < kSortItems; ++i)
    {
        unsigned d      = unsigned(key[i] >> shift) & (kRadix - 1);
        unsigned peers  = matchDigit(d);
        unsigned leader = __ffs(int(peers)) - 1;
        uint32_t before = 0;
        if (lane == leader)
        {
            before    = myHist[d];
            myHist[d] = before + __popc(peers);
        }
        before  = __shfl_sync(0xffffffffu, before, leader);
        rank[i] = uint16_t(before + __popc(peers & lanesLt));
        __syncwarp();
    }
    __syncthreads();

    // ---- per digit (thread d = digit d): tile count, aggregate published, tile-local starts of every (warp, digit)
    uint32_t st[kLookBack];
    uint32_t tileCount, dStart;
    uint32_t* const myStatus = status + size_t(tile) * kRadix + tid;
    {
        uint32_t c[kSortWarps];
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w)
        {
            c[w] = sm.warpHist[w][tid];
            run += c[w];
        }
        tileCount = run;
        if (tile == 0) stRelaxed(myStatus, kStInclusive | tileCount);
        else stRelaxed(myStatus, kStAggregate | tileCount);

        // block-wide exclusive scan of tileCount over the 256 digits
        uint32_t incl = tileCount;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (int(lane) >= o) incl += t;
        }
        if (lane == 31) sm.warpSums[warp] = incl;
        __syncthreads();
        uint32_t off = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w)
            if (w < int(warp)) off += sm.warpSums[w];
        dStart = off + incl - tileCount;
        run    = dStart;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w)
        {
            sm.warpHist[w][tid] = run;
            run += c[w];
        }
        /* first window of the decoupled look-back: the status words of the kLookBack preceding tiles are requested
         * now and consumed after the reorder, so their L2 round trip (and part of the wait for the predecessors)
         * overlaps the shared-memory traffic.  Indices below tile 0 are clamped; tile 0 is always inclusive, so a
         * clamped copy is never consumed. */
        if (tile > 0)
        {
#pragma unroll
            for (int k = 0; k < kLookBack; ++k)
                st[k] = ldRelaxed(status + size_t(max(int(tile) - 1 - k, 0)) * kRadix + tid);
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();

    // ---- reorder through shared memory
#pragma unroll
    for (int i = 0; i < kSortItems; ++i)
    {
        unsigned d   = unsigned(key[i] >> shift) & (kRadix - 1);
        uint32_t pos = sm.warpHist[warp][d] + rank[i];
        sm.keys[pos] = key[i];
        sm.vals[pos] = Iota ? tileBase + warpOff + 32u * i : sm.stage[warpOff + 32u * i];
    }

    // ---- look-back: the pre-fetched window first, then wider windows (the key registers are free by now)
    {
        uint32_t excl = 0;
        if (tile > 0)
        {
            int t = int(tile) - 1;
            if (!consumeWindow<kLookBack>(st, excl, t))
            {
                /* wait on ONE word (the first tile that was not ready), then fetch the window behind it: polling with
                 * the whole window floods the L2 when many tiles reach the look-back together (small or digit-uniform
                 * sorts: 5x slower passes) */
                uint32_t wide[kLookBackWide];
                do
                {
                    const uint32_t* sp = status + size_t(t) * kRadix + tid;
                    do
                    {
                        wide[0] = ldRelaxed(sp);
                    } while ((wide[0] >> kStFlagShift) == 0u);
#pragma unroll
                    for (int k = 1; k < kLookBackWide; ++k)
                        wide[k] = ldRelaxed(status + size_t(max(t - k, 0)) * kRadix + tid);
                } while (!consumeWindow<kLookBackWide>(wide, excl, t));
            }
            stRelaxed(myStatus, kStInclusive | (excl + tileCount));
        }
        sm.globalBase[tid] = binOffsets[tid] + excl - dStart;
    }
    __syncthreads();

    // ---- write out: consecutive slots of one digit go to consecutive global addresses
#pragma unroll
    for (int i = 0; i < kSortItems; ++i)
    {
        uint32_t slot = tid + i * kSortThreads;
        K k           = sm.keys[slot];
        unsigned d    = unsigned(k >> shift) & (kRadix - 1);
        uint32_t pos  = sm.globalBase[d] + slot;
        if (fullTile || pos < n)
        {
            keysOut[pos] = k;
            valsOut[pos] = sm.vals[slot];
        }
    }
}

} // namespace csb

namespace csb
{

template<class K>
int sortByKey(K* keys, uint32_t* values, K* keyBuf, uint32_t* valueBuf, size_t n, void* temp, uint64_t tempBytes,
              int valuesAreIota, cudaStream_t s)
{
    constexpr int P = numPasses<K>();
    static_assert(P % 2 == 0, "even number of passes: the result lands in the caller's primary buffers");
    if (n == 0) return 0;
    if (!keys || !values || !keyBuf || !valueBuf || !temp) return CSTONE_E_INVALID_ARGUMENT;
    if (n >= (size_t(1) << 30)) return CSTONE_E_TOO_LARGE;
    if (tempBytes < sortTempBytes<K>(n)) return CSTONE_E_TEMP_TOO_SMALL;

    SortTemp t = carveSortTemp<K>(temp, n);
    CSB_CUDA(cudaMemsetAsync(temp, 0, t.bytes, s));

    histogramKernel<K><<<gridFor(n, 512 * 8, 2), 512, 0, s>>>(keys, n, t.hist);
    CSB_LAUNCHED();
    binOffsetsKernel<<<P, kRadix, 0, s>>>(t.hist, t.tickets, uint32_t(n));
    CSB_LAUNCHED();

    static bool attrSet[2] = {false, false};
    const size_t smemBytes = sizeof(SortSmem<K>);
    if (!attrSet[0])
    {
        CSB_CUDA(cudaFuncSetAttribute(onesweepKernel<K, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      int(smemBytes)));
        CSB_CUDA(cudaFuncSetAttribute(onesweepKernel<K, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      int(smemBytes)));
        // per-device attribute: do not cache across devices (cheap call), keep the flag only as documentation
    }

    const unsigned numTiles = unsigned(sortNumTiles(n));
    K* kSrc                 = keys;
    K* kDst                 = keyBuf;
    uint32_t* vSrc          = values;
    uint32_t* vDst          = valueBuf;
    for (int p = 0; p < P; ++p)
    {
        uint32_t* status = t.status + size_t(p) * numTiles * kRadix;
        if (p == 0 && valuesAreIota)
            onesweepKernel<K, true><<<numTiles, kSortThreads, smemBytes, s>>>(
                kSrc, vSrc, kDst, vDst, uint32_t(n), p * kRadixBits, t.hist + p * kRadix, t.tickets + p * 32, status);
        else
            onesweepKernel<K, false><<<numTiles, kSortThreads, smemBytes, s>>>(
                kSrc, vSrc, kDst, vDst, uint32_t(n), p * kRadixBits, t.hist + p * kRadix, t.tickets + p * 32, status);
        CSB_LAUNCHED();
        K* tk        = kSrc;
        kSrc         = kDst;
        kDst         = tk;
        uint32_t* tv = vSrc;
        vSrc         = vDst;
        vDst         = tv;
    }
    return 0;
}

__global__ void __launch_bounds__(256) sequenceKernel(uint32_t* __restrict__ out, size_t n, uint32_t init)
{
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = init + uint32_t(i);
}

} // namespace csb

extern "C" {

uint64_t cstone_b200_sort_by_key_temp_bytes_k32(uint64_t n) { return csb::sortTempBytes<uint32_t>(n); }
uint64_t cstone_b200_sort_by_key_temp_bytes_k64(uint64_t n) { return csb::sortTempBytes<uint64_t>(n); }

int cstone_b200_sort_by_key_k32(uint32_t* keys, uint32_t* values, uint32_t* keyBuf, uint32_t* valueBuf, size_t n,
                                void* temp, uint64_t tempBytes, int iota, void* stream)
{
    return csb::sortByKey<uint32_t>(keys, values, keyBuf, valueBuf, n, temp, tempBytes, iota, csb::asStream(stream));
}
int cstone_b200_sort_by_key_k64(uint64_t* keys, uint32_t* values, uint64_t* keyBuf, uint32_t* valueBuf, size_t n,
                                void* temp, uint64_t tempBytes, int iota, void* stream)
{
    return csb::sortByKey<uint64_t>(keys, values, keyBuf, valueBuf, n, temp, tempBytes, iota, csb::asStream(stream));
}

int cstone_b200_sequence_u32(uint32_t* out, size_t n, uint32_t init, void* stream)
{
    if (n == 0) return 0;
    csb::sequenceKernel<<<csb::gridFor(n, 256), 256, 0, csb::asStream(stream)>>>(out, n, init);
    CSB_LAUNCHED();
    return 0;
}

} // extern "C"
