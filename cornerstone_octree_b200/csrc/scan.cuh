/*
 * scan.cuh — single-pass prefix sum with decoupled look-back (one read + one write of the data).
 *
 * Stands in for the thrust::exclusive_scan / inclusive_scan calls of the reference
 * [ref: primitives/primitives_gpu.cu:394-448, tree/csarray_gpu.cu:121,200].  The element loader is a functor, so
 * producers (rebalance decision, halo-masked counts) are fused into the scan instead of materialising their output.
 *
 * Tile = 256 threads x 8 items.  Tile ids are handed out by an atomic ticket so that every tile a block waits for
 * has already been scheduled (forward progress of the look-back does not depend on block scheduling order).
 * Tile status is one 64-bit word {flag:32 | value:32} written with a single store.
 */
#pragma once
#include "common.cuh"

namespace csb
{

constexpr int kScanThreads = 256;
constexpr int kScanItems   = 8;
constexpr int kScanTile    = kScanThreads * kScanItems;

constexpr uint32_t kFlagInvalid   = 0;
constexpr uint32_t kFlagAggregate = 1;
constexpr uint32_t kFlagInclusive = 2;

struct ScanTemp
{
    unsigned* ticket;   // 1 counter
    uint64_t* status;   // numTiles words
    size_t bytes;       // bytes to zero before the launch
};

__host__ __device__ inline size_t scanNumTiles(size_t n) { return (n + kScanTile - 1) / kScanTile; }
inline uint64_t scanTempBytes(uint64_t n) { return 128 + 8 * scanNumTiles(n) + 8; }

inline ScanTemp carveScanTemp(void* temp, size_t n)
{
    ScanTemp t;
    t.ticket = reinterpret_cast<unsigned*>(temp);
    t.status = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(temp) + 128);
    t.bytes  = size_t(scanTempBytes(n));
    return t;
}

__device__ inline uint64_t packStatus(uint32_t flag, uint32_t value) { return (uint64_t(flag) << 32) | value; }

__device__ inline uint64_t loadStatus(const uint64_t* p)
{
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ inline void storeStatus(uint64_t* p, uint64_t v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

/*! look back over the predecessors of `tile` with one warp; returns the exclusive prefix of the tile
 *  (sum of the totals of tiles 0..tile-1).  Must be called by all 32 lanes of one warp. */
__device__ inline uint32_t warpLookBack(const uint64_t* status, int tile)
{
    uint32_t prefix = 0;
    int base        = tile - 1; // most recent predecessor examined by lane 0
    while (base >= 0)
    {
        int idx      = base - int(laneId());
        uint64_t st  = packStatus(kFlagInclusive, 0); // virtual tiles before 0 contribute nothing
        if (idx >= 0)
        {
            do
            {
                st = loadStatus(status + idx);
            } while (uint32_t(st >> 32) == kFlagInvalid);
        }
        uint32_t flag     = uint32_t(st >> 32);
        unsigned inclMask = __ballot_sync(0xffffffffu, flag == kFlagInclusive);
        // lanes at or before the first inclusive entry contribute
        int firstIncl     = inclMask ? (__ffs(int(inclMask)) - 1) : 32;
        uint32_t contrib  = (int(laneId()) <= firstIncl) ? uint32_t(st) : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
        prefix += contrib;
        if (inclMask) break;
        base -= 32;
    }
    return prefix;
}

/*! generic scan kernel.
 *  Load:  uint32_t operator()(size_t i) const   — value of element i (i < n)
 *  out[i] = init + sum_{j<i} v_j (exclusive) or init + sum_{j<=i} v_j (inclusive)
 *  total (optional device pointer) receives init + sum of all elements. */
template<bool Inclusive, class Load>
__global__ void __launch_bounds__(kScanThreads)
scanKernel(Load load, uint32_t* __restrict__ out, size_t n, uint32_t init, ScanTemp tmp, uint32_t* __restrict__ total)
{
    __shared__ unsigned sTile;
    __shared__ uint32_t sWarp[kScanThreads / 32];
    __shared__ uint32_t sPrefix;

    if (threadIdx.x == 0) sTile = atomicAdd(tmp.ticket, 1u);
    __syncthreads();
    const int tile    = int(sTile);
    const size_t base = size_t(tile) * kScanTile + size_t(threadIdx.x) * kScanItems;

    uint32_t v[kScanItems];
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i)
    {
        v[i] = (base + i < n) ? load(base + i) : 0u;
        sum += v[i];
    }
    // block-wide exclusive scan of the per-thread sums
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (int(laneId()) >= o) incl += t;
    }
    const int warp = threadIdx.x >> 5;
    if (laneId() == 31) sWarp[warp] = incl;
    __syncthreads();
    uint32_t warpOffset = 0, blockTotal = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; ++w)
    {
        uint32_t t = sWarp[w];
        if (w < warp) warpOffset += t;
        blockTotal += t;
    }

    if (warp == 0)
    {
        if (tile == 0)
        {
            if (laneId() == 0)
            {
                storeStatus(tmp.status, packStatus(kFlagInclusive, blockTotal));
                sPrefix = 0;
            }
        }
        else
        {
            if (laneId() == 0) storeStatus(tmp.status + tile, packStatus(kFlagAggregate, blockTotal));
            uint32_t p = warpLookBack(tmp.status, tile);
            if (laneId() == 0)
            {
                storeStatus(tmp.status + tile, packStatus(kFlagInclusive, p + blockTotal));
                sPrefix = p;
            }
        }
    }
    __syncthreads();
    uint32_t run = init + sPrefix + warpOffset + (incl - sum);
#pragma unroll
    for (int i = 0; i < kScanItems; ++i)
    {
        if (Inclusive) run += v[i];
        if (base + i < n) out[base + i] = run;
        if (!Inclusive) run += v[i];
    }
    if (total && size_t(tile) == scanNumTiles(n) - 1 && threadIdx.x == kScanThreads - 1)
    {
        *total = init + sPrefix + blockTotal;
    }
}

struct LoadPlain
{
    const uint32_t* in;
    __device__ uint32_t operator()(size_t i) const { return in[i]; }
};

//! host launcher; temp must hold scanTempBytes(n) bytes
template<bool Inclusive, class Load>
int launchScan(Load load, uint32_t* out, size_t n, uint32_t init, void* temp, uint32_t* totalDev, cudaStream_t s)
{
    if (n == 0)
    {
        if (totalDev)
        {
            CSB_CUDA(cudaStreamSynchronize(s));
            CSB_CUDA(cudaMemcpy(totalDev, &init, sizeof(uint32_t), cudaMemcpyHostToDevice));
        }
        return 0;
    }
    ScanTemp t = carveScanTemp(temp, n);
    CSB_CUDA(cudaMemsetAsync(temp, 0, t.bytes, s));
    scanKernel<Inclusive, Load><<<unsigned(scanNumTiles(n)), kScanThreads, 0, s>>>(load, out, n, init, t, totalDev);
    CSB_LAUNCHED();
    return 0;
}

} // namespace csb
