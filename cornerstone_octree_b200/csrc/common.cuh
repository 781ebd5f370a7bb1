/*
 * common.cuh — shared device/host helpers of libcstone_b200 (sm_100a only).
 *
 * Key algebra follows the reference definitions (cited per function, paths relative to the
 * reference's include/cstone/); the code is our own.
 */
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include <cstdio>
#include <cstdlib>

#include "../../include/cstone_b200.h"

namespace csb
{

constexpr int kNumSMs = 148; // B200: 2 dies x 74 SMs

// ------------------------------------------------------------------------------------------------
// host side: status handling + launch accounting
// ------------------------------------------------------------------------------------------------
extern std::atomic<uint64_t> g_launchCount;

inline void countLaunch(int n = 1) { g_launchCount.fetch_add(uint64_t(n), std::memory_order_relaxed); }

//! CSB_TRACE_ERRORS=1 in the environment: the first failing step of a library call is named on stderr
inline int traceStatus(int status, const char* file, int line)
{
    static const bool on = std::getenv("CSB_TRACE_ERRORS") != nullptr;
    if (on && status != 0) std::fprintf(stderr, "[cstone_b200] status %d at %s:%d\n", status, file, line);
    return status;
}

#define CSB_CUDA(expr)                                                                                                 \
    do                                                                                                                 \
    {                                                                                                                  \
        cudaError_t e_ = (expr);                                                                                       \
        if (e_ != cudaSuccess) return csb::traceStatus(int(e_), __FILE__, __LINE__);                                   \
    } while (0)

#define CSB_TRY(expr)                                                                                                  \
    do                                                                                                                 \
    {                                                                                                                  \
        int tryStatus_ = (expr);                                                                                       \
        if (tryStatus_ != 0) return csb::traceStatus(tryStatus_, __FILE__, __LINE__);                                  \
    } while (0)

//! to be placed after every kernel launch
#define CSB_LAUNCHED()                                                                                                 \
    do                                                                                                                 \
    {                                                                                                                  \
        csb::countLaunch();                                                                                            \
        cudaError_t e_ = cudaGetLastError();                                                                           \
        if (e_ != cudaSuccess) return int(e_);                                                                         \
    } while (0)

inline cudaStream_t asStream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

inline unsigned divUp(size_t a, size_t b) { return unsigned((a + b - 1) / b); }

//! grid size for grid-stride kernels: a multiple of the SM count, capped by the work
inline unsigned gridFor(size_t work, unsigned perBlock, unsigned blocksPerSM = 8)
{
    size_t need = (work + perBlock - 1) / perBlock;
    size_t cap  = size_t(kNumSMs) * blocksPerSM;
    if (need >= cap) return unsigned(cap);
    return unsigned(need ? need : 1);
}

// ------------------------------------------------------------------------------------------------
// key traits  [ref: tree/definitions.h:45-91]
// ------------------------------------------------------------------------------------------------
template<class K>
struct KeyTraits;

template<>
struct KeyTraits<uint32_t>
{
    static constexpr unsigned maxLevel = 10;
    static constexpr unsigned unused   = 2;
    static constexpr unsigned keyBits  = 30;
};
template<>
struct KeyTraits<uint64_t>
{
    static constexpr unsigned maxLevel = 21;
    static constexpr unsigned unused   = 1;
    static constexpr unsigned keyBits  = 63;
};

__host__ __device__ inline int clzKey(uint32_t v)
{
#ifdef __CUDA_ARCH__
    return __clz(int(v));
#else
    return v ? __builtin_clz(v) : 32;
#endif
}
__host__ __device__ inline int clzKey(uint64_t v)
{
#ifdef __CUDA_ARCH__
    return __clzll((long long)v);
#else
    return v ? __builtin_clzll(v) : 64;
#endif
}

//! key span of a node at subdivision `level`  [ref: sfc/common.hpp:130-137]
template<class K>
__host__ __device__ inline K nodeRange(unsigned level)
{
    return K(1) << (3u * (KeyTraits<K>::maxLevel - level));
}

//! level of a power-of-8 key range  [ref: sfc/common.hpp:178-183]
template<class K>
__host__ __device__ inline unsigned treeLevel(K range)
{
    return unsigned(clzKey(K(range - 1)) - int(KeyTraits<K>::unused)) / 3u;
}

//! [ref: sfc/common.hpp:164-168]
template<class K>
__host__ __device__ inline int commonPrefix(K a, K b)
{
    return clzKey(K(a ^ b)) - int(KeyTraits<K>::unused);
}

//! octal digit at level `pos`, 1 = most significant  [ref: sfc/common.hpp:275-279]
template<class K>
__host__ __device__ inline unsigned octalDigit(K key, unsigned pos)
{
    return unsigned(key >> (3u * (KeyTraits<K>::maxLevel - pos))) & 7u;
}

//! Warren-Salmon placeholder-bit keys  [ref: sfc/common.hpp:196-235]
template<class K>
__host__ __device__ inline K encodePlaceholderBit(K key, int prefixLength)
{
    return (K(1) << prefixLength) | (key >> (3 * int(KeyTraits<K>::maxLevel) - prefixLength));
}
template<class K>
__host__ __device__ inline unsigned decodePrefixLength(K prefix)
{
    return 8u * sizeof(K) - 1u - unsigned(clzKey(prefix));
}
template<class K>
__host__ __device__ inline K decodePlaceholderBit(K prefix)
{
    int len = int(decodePrefixLength(prefix));
    return K(prefix ^ (K(1) << len)) << (3 * int(KeyTraits<K>::maxLevel) - len);
}

// ------------------------------------------------------------------------------------------------
// bounding box passed BY VALUE into kernels (same as the reference does with Box<T>)
// ------------------------------------------------------------------------------------------------
template<class T>
struct BoxT;
template<>
struct BoxT<float>
{
    typedef cstone_box_f32 type;
};
template<>
struct BoxT<double>
{
    typedef cstone_box_f64 type;
};

template<class T>
struct DBox
{
    T lo[3], hi[3], len[3], ilen[3];
    int pbc[3];
};

template<class T>
inline DBox<T> makeDBox(const typename BoxT<T>::type* b)
{
    DBox<T> d;
    for (int k = 0; k < 3; ++k)
    {
        d.lo[k]   = b->limits[2 * k];
        d.hi[k]   = b->limits[2 * k + 1];
        d.len[k]  = b->lengths[k];
        d.ilen[k] = b->inv_lengths[k];
        d.pbc[k]  = (b->boundaries[k] == CSTONE_BOUNDARY_PERIODIC);
    }
    return d;
}

// ------------------------------------------------------------------------------------------------
// Hilbert state-machine tables (tables.cu).  The reference evaluates the Hilbert curve level by level
// with bit flips and coordinate rotations [ref: sfc/hilbert.hpp:58-109].  The accumulated flip/rotation is
// one of 24 signed axis permutations; we tabulate three levels at a time:
//   enc[state*512 + (x3<<6 | y3<<3 | z3)] = 9 key bits | nextState << 9
//   dec[state*512 + 9 key bits]           = (x3<<6 | y3<<3 | z3) | nextState << 9
// ------------------------------------------------------------------------------------------------
constexpr int kHilbertStates   = 24;
constexpr int kHilbertTabSize  = kHilbertStates * 512;
constexpr int kHilbertTabBytes = kHilbertTabSize * 2;

struct SfcTables
{
    const uint16_t* enc; // device pointers
    const uint16_t* dec;
};

//! device tables for the CURRENT device (uploaded on first use); nullptr members on failure
int getSfcTables(SfcTables* out);
//! host copies (for unit checks of the table generator)
const uint16_t* hostEncTable();
const uint16_t* hostDecTable();

//! cooperative copy of one 24 KB table into shared memory
__device__ inline void loadTable(uint16_t* smemTab, const uint16_t* gTab)
{
    const uint4* src = reinterpret_cast<const uint4*>(gTab);
    uint4* dst       = reinterpret_cast<uint4*>(smemTab);
    for (int i = threadIdx.x; i < kHilbertTabBytes / 16; i += blockDim.x)
        dst[i] = src[i];
}

// ------------------------------------------------------------------------------------------------
// Morton  [ref: sfc/morton.hpp:52-125]
// ------------------------------------------------------------------------------------------------
__host__ __device__ inline uint32_t spread3(uint32_t v, uint32_t)
{
    v &= 0x3ffu;
    v = (v | (v << 16)) & 0x030000ffu;
    v = (v | (v << 8)) & 0x0300f00fu;
    v = (v | (v << 4)) & 0x030c30c3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__host__ __device__ inline uint64_t spread3(uint32_t v, uint64_t)
{
    uint64_t x = v & 0x1fffffu;
    x          = (x | (x << 32)) & 0x001f00000000ffffull;
    x          = (x | (x << 16)) & 0x001f0000ff0000ffull;
    x          = (x | (x << 8)) & 0x100f00f00f00f00full;
    x          = (x | (x << 4)) & 0x10c30c30c30c30c3ull;
    x          = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}
__host__ __device__ inline uint32_t compact3(uint32_t v)
{
    v &= 0x09249249u;
    v = (v | (v >> 2)) & 0x030c30c3u;
    v = (v | (v >> 4)) & 0x0300f00fu;
    v = (v | (v >> 8)) & 0x030000ffu;
    v = (v | (v >> 16)) & 0x3ffu;
    return v;
}
__host__ __device__ inline uint32_t compact3(uint64_t v)
{
    v &= 0x1249249249249249ull;
    v = (v | (v >> 2)) & 0x10c30c30c30c30c3ull;
    v = (v | (v >> 4)) & 0x100f00f00f00f00full;
    v = (v | (v >> 8)) & 0x001f0000ff0000ffull;
    v = (v | (v >> 16)) & 0x001f00000000ffffull;
    v = (v | (v >> 32)) & 0x1fffffull;
    return uint32_t(v);
}

template<class K>
__host__ __device__ inline K mortonEncode(uint32_t ix, uint32_t iy, uint32_t iz)
{
    return (spread3(ix, K(0)) << 2) | (spread3(iy, K(0)) << 1) | spread3(iz, K(0));
}
template<class K>
__host__ __device__ inline void mortonDecode(K key, uint32_t& ix, uint32_t& iy, uint32_t& iz)
{
    ix = compact3(K(key >> 2));
    iy = compact3(K(key >> 1));
    iz = compact3(key);
}

// ------------------------------------------------------------------------------------------------
// Hilbert through the tables (tab points to shared or global memory)
// ------------------------------------------------------------------------------------------------
template<class K>
__host__ __device__ inline K hilbertEncode(const uint16_t* enc, uint32_t ix, uint32_t iy, uint32_t iz)
{
    // number of 3-level steps; 32-bit keys (10 levels) are padded to 12 levels and shifted back
    constexpr int steps = (KeyTraits<K>::maxLevel + 2) / 3;
    constexpr int pad   = steps * 3 - int(KeyTraits<K>::maxLevel);
    ix <<= pad, iy <<= pad, iz <<= pad;
    uint64_t key   = 0;
    unsigned state = 0;
#pragma unroll
    for (int s = steps - 1; s >= 0; --s)
    {
        unsigned idx = (((ix >> (3 * s)) & 7u) << 6) | (((iy >> (3 * s)) & 7u) << 3) | ((iz >> (3 * s)) & 7u);
        unsigned e   = enc[state * 512u + idx];
        key          = (key << 9) | (e & 511u);
        state        = e >> 9;
    }
    return K(key >> (3 * pad));
}

/*! decode the first `levels` octal digits of `key`; the lower coordinate bits are returned as zero, which is
 *  exactly what the reference's hilbertIBox produces by masking a full decode [ref: sfc/hilbert.hpp:274-290] */
template<class K>
__host__ __device__ inline void
hilbertDecode(const uint16_t* dec, K key, unsigned levels, uint32_t& ix, uint32_t& iy, uint32_t& iz)
{
    constexpr int steps = (KeyTraits<K>::maxLevel + 2) / 3;
    constexpr int pad   = steps * 3 - int(KeyTraits<K>::maxLevel);
    uint64_t k          = uint64_t(key) << (3 * pad);
    unsigned state      = 0;
    uint32_t x = 0, y = 0, z = 0;
    int need = int(levels + 2) / 3; // steps that carry requested digits
#pragma unroll
    for (int s = steps - 1; s >= 0; --s)
    {
        if (steps - 1 - s < need)
        {
            unsigned d = unsigned(k >> (9 * s)) & 511u;
            unsigned e = dec[state * 512u + d];
            state      = e >> 9;
            x |= ((e >> 6) & 7u) << (3 * s);
            y |= ((e >> 3) & 7u) << (3 * s);
            z |= (e & 7u) << (3 * s);
        }
    }
    uint32_t mask = ~((1u << (KeyTraits<K>::maxLevel - levels)) - 1u);
    ix            = (x >> pad) & mask;
    iy            = (y >> pad) & mask;
    iz            = (z >> pad) & mask;
}

template<class K>
__host__ __device__ inline K sfcEncode(int kind, const uint16_t* enc, uint32_t ix, uint32_t iy, uint32_t iz)
{
    return kind == CSTONE_SFC_MORTON ? mortonEncode<K>(ix, iy, iz) : hilbertEncode<K>(enc, ix, iy, iz);
}

//! integer box of the node starting at `key` on `level`  [ref: sfc/hilbert.hpp:274-290, sfc/morton.hpp:175-182]
struct IBox
{
    int lo[3], hi[3];
};
template<class K>
__host__ __device__ inline IBox sfcIBox(int kind, const uint16_t* dec, K key, unsigned level)
{
    uint32_t c[3];
    if (kind == CSTONE_SFC_MORTON) mortonDecode<K>(key, c[0], c[1], c[2]);
    else hilbertDecode<K>(dec, key, level, c[0], c[1], c[2]);
    int cube = int((1u << KeyTraits<K>::maxLevel) >> level);
    IBox b;
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        b.lo[d] = int(c[d]);
        b.hi[d] = int(c[d]) + cube;
    }
    return b;
}

// ------------------------------------------------------------------------------------------------
// warp helpers
// ------------------------------------------------------------------------------------------------
__device__ inline unsigned laneId() { return threadIdx.x & 31u; }
__device__ inline unsigned laneMaskLt()
{
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

} // namespace csb
