/*
 * primitives.cu — the small device primitives the callers of the hot path use between its stages.
 *
 * Each entry point replaces one `extern template` of the reference's primitives / csarray / halo-gather GPU
 * interface with the same element-wise meaning:
 *   fillGpu, scaleGpu, incrementGpu, maxNormSquareGpu, lowerBoundGpu (scalar and vector form), segmentMax,
 *   reduceGpu, countGpu, sortGpu        [ref: primitives/primitives_gpu.h:39-125, primitives_gpu.cu:51-448]
 *   countSfcGapsGpu, fillSfcGapsGpu      [ref: tree/csarray_gpu.h:86-89, csarray_gpu.cu:231-265, sfc/common.hpp:392-438]
 *   gatherRanges                         [ref: halos/gather_halos_gpu.h:21-27, gather_halos_gpu.cu:25-54]
 * Used by Halos::discover (fill, segmentMax, scale: halos/halos.hpp:158-164), createSendRangesGpu (vector lower
 * bound: domain/domaindecomp_gpu.cuh:65), GpuSfcSorter::extendMap (increment: primitives/gather.cuh:136),
 * haloExchangeGpu (gatherRanges: halos/exchange_halos_gpu.cuh:71), injectKeysGpu (sortGpu: focus/inject.hpp:97).
 *
 * All of them are streaming, HBM-bound passes: grid-stride loops over a multiple of 148 blocks, 16-byte accesses
 * where the element type allows, reductions through warp shuffles + one partial per block + a one-block final pass.
 */
#include "common.cuh"

namespace csb
{

constexpr int kPrimThreads   = 256;
constexpr int kReduceBlocks  = kNumSMs * 4;

// ------------------------------------------------------------------------------------------------ element-wise
template<class T>
__global__ void __launch_bounds__(kPrimThreads) fillKernel(T* __restrict__ out, size_t n, T value)
{
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = value;
}

template<class T>
__global__ void __launch_bounds__(kPrimThreads) scaleKernel(T* __restrict__ v, size_t n, T value)
{
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        v[i] *= value;
}

template<class T>
__global__ void __launch_bounds__(kPrimThreads) incrementKernel(const T* __restrict__ in, T* __restrict__ out, size_t n,
                                                                T value)
{
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = in[i] + value;
}

// ------------------------------------------------------------------------------------------------ searches
//! first index in [0,n) with !(a[idx] < v), i.e. std::lower_bound
template<class T>
__device__ inline size_t lowerBoundDev(const T* __restrict__ a, size_t n, T v)
{
    size_t lo = 0, len = n;
    while (len > 0)
    {
        const size_t half = len >> 1;
        const bool less   = a[lo + half] < v;
        lo                = less ? lo + half + 1 : lo;
        len               = less ? len - half - 1 : half;
    }
    return lo;
}

template<class T, class I>
__global__ void __launch_bounds__(kPrimThreads)
    lowerBoundKernel(const T* __restrict__ a, size_t n, const T* __restrict__ values, size_t numValues,
                     I* __restrict__ result)
{
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < numValues; i += stride)
        result[i] = I(lowerBoundDev(a, n, values[i]));
}

template<class T>
__global__ void lowerBoundOneKernel(const T* __restrict__ a, size_t n, T v, unsigned long long* __restrict__ out)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) *out = (unsigned long long)lowerBoundDev(a, n, v);
}

// ------------------------------------------------------------------------------------------------ segmentMax
/* one thread per segment like the reference for short segments (leaves hold <= bucketSize particles); a warp per
 * segment when the mean segment is long, so that the loads of a segment coalesce */
template<class Tin, class Tout, class I>
__global__ void __launch_bounds__(kPrimThreads)
    segmentMaxKernel(const Tin* __restrict__ in, const I* __restrict__ seg, size_t numSegments, Tout* __restrict__ out)
{
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t s = size_t(blockIdx.x) * blockDim.x + threadIdx.x; s < numSegments; s += stride)
    {
        Tin m = 0;
        const I b = seg[s], e = seg[s + 1];
        for (I i = b; i < e; ++i)
            m = max(m, in[i]);
        out[s] = Tout(m);
    }
}

template<class Tin, class Tout, class I>
__global__ void __launch_bounds__(kPrimThreads)
    segmentMaxWarpKernel(const Tin* __restrict__ in, const I* __restrict__ seg, size_t numSegments, Tout* __restrict__ out)
{
    const unsigned lane  = threadIdx.x & 31u;
    const size_t warps   = (size_t(gridDim.x) * blockDim.x) >> 5;
    for (size_t s = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; s < numSegments; s += warps)
    {
        Tin m = 0;
        const I b = seg[s], e = seg[s + 1];
        for (I i = b + lane; i < e; i += 32)
            m = max(m, in[i]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) out[s] = Tout(m);
    }
}

// ------------------------------------------------------------------------------------------------ reductions
struct OpSum
{
    __device__ static unsigned long long apply(unsigned long long a, unsigned long long b) { return a + b; }
};
template<class T>
struct OpMax
{
    __device__ static T apply(T a, T b) { return max(a, b); }
};

template<class R, class Op, class Load>
__global__ void __launch_bounds__(kPrimThreads) reducePartialKernel(Load load, size_t n, R init, R* __restrict__ partial)
{
    __shared__ R sRed[kPrimThreads / 32];
    R acc               = init;
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        acc = Op::apply(acc, load(i));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        acc = Op::apply(acc, __shfl_xor_sync(0xffffffffu, acc, o));
    if ((threadIdx.x & 31) == 0) sRed[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        R v = sRed[0];
        for (int w = 1; w < kPrimThreads / 32; ++w)
            v = Op::apply(v, sRed[w]);
        partial[blockIdx.x] = v;
    }
}

template<class R, class Op>
__global__ void __launch_bounds__(kPrimThreads) reduceFinalKernel(R* __restrict__ partial, int numPartials, R init, R first)
{
    __shared__ R sRed[kPrimThreads / 32];
    R acc = init;
    for (int i = threadIdx.x; i < numPartials; i += kPrimThreads)
        acc = Op::apply(acc, partial[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        acc = Op::apply(acc, __shfl_xor_sync(0xffffffffu, acc, o));
    if ((threadIdx.x & 31) == 0) sRed[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        R v = first;
        for (int w = 0; w < kPrimThreads / 32; ++w)
            v = Op::apply(v, sRed[w]);
        partial[numPartials] = v;
    }
}

/* neutral = identity of Op for the per-thread accumulators, first = the caller's init, applied once */
template<class R, class Op, class Load>
int reduceToHost(Load load, size_t n, R neutral, R first, R* outHost, void* temp, cudaStream_t s)
{
    if (!outHost) return CSTONE_E_INVALID_ARGUMENT;
    if (n == 0)
    {
        *outHost = first;
        return 0;
    }
    if (!temp) return CSTONE_E_INVALID_ARGUMENT;
    R* partial       = reinterpret_cast<R*>(temp);
    const int blocks = int(gridFor(n, kPrimThreads * 4, 4));
    reducePartialKernel<R, Op, Load><<<blocks, kPrimThreads, 0, s>>>(load, n, neutral, partial);
    CSB_LAUNCHED();
    reduceFinalKernel<R, Op><<<1, kPrimThreads, 0, s>>>(partial, blocks, neutral, first);
    CSB_LAUNCHED();
    CSB_CUDA(cudaMemcpyAsync(outHost, partial + blocks, sizeof(R), cudaMemcpyDeviceToHost, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    return 0;
}

template<class T>
struct LoadNormSq
{
    const T *x, *y, *z;
    // [ref: primitives_gpu.cu:196-201] x*x + y*y + z*z, left to right, no contraction
    __device__ T operator()(size_t i) const { return x[i] * x[i] + y[i] * y[i] + z[i] * z[i]; }
};
struct LoadU32
{
    const uint32_t* v;
    __device__ unsigned long long operator()(size_t i) const { return v[i]; }
};
template<class T>
struct LoadEquals
{
    const T* v;
    T value;
    __device__ unsigned long long operator()(size_t i) const { return v[i] == value ? 1ull : 0ull; }
};

// ------------------------------------------------------------------------------------------------ SFC gaps
//! last non-zero octal digit place of a key, maxLevel for key 0  [ref: sfc/common.hpp:339-346]
template<class K>
__device__ inline int lastNzPlace(K key)
{
    if (key == 0) return int(KeyTraits<K>::maxLevel);
    const int tz = (sizeof(K) == 8) ? (__ffsll((long long)key) - 1) : (__ffs(int(key)) - 1);
    return int(KeyTraits<K>::maxLevel) - tz / 3;
}

/* the cornerstone keys that span [a, b) with the fewest nodes  [ref: sfc/common.hpp:392-430]; Store = false only
 * counts them */
template<class K, bool Store>
__device__ inline int spanSfcRange(K a, K b, K* out)
{
    int numValues           = 0;
    const int firstDiffPos  = (clzKey(K(a ^ b)) + 3 - int(KeyTraits<K>::unused)) / 3;
    const int aLast         = lastNzPlace(a);
    const int bLast         = lastNzPlace(b);
    for (int pos = aLast; pos > firstDiffPos; --pos)
    {
        int numDigits = (8 - int(octalDigit(a, unsigned(pos)))) % 8;
        numValues += numDigits;
        while (numDigits--)
        {
            if (Store) *out++ = a;
            a += nodeRange<K>(unsigned(pos));
        }
    }
    for (int pos = firstDiffPos; pos <= bLast; ++pos)
    {
        int numDigits = int(octalDigit(b, unsigned(pos))) - int(octalDigit(a, unsigned(pos)));
        numValues += numDigits;
        while (numDigits--)
        {
            if (Store) *out++ = a;
            a += nodeRange<K>(unsigned(pos));
        }
    }
    return numValues;
}

template<class K>
__global__ void __launch_bounds__(kPrimThreads) countSfcGapsKernel(const K* __restrict__ tree, int numNodes,
                                                                   int* __restrict__ nodeOps)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < numNodes; i += stride)
        nodeOps[i] = spanSfcRange<K, false>(tree[i], tree[i + 1], nullptr);
}

template<class K>
__global__ void __launch_bounds__(kPrimThreads) fillSfcGapsKernel(const K* __restrict__ tree, int numNodes,
                                                                  const int* __restrict__ nodeOps, K* __restrict__ newTree)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= numNodes; i += stride)
    {
        if (i < numNodes) spanSfcRange<K, true>(tree[i], tree[i + 1], newTree + nodeOps[i]);
        else newTree[nodeOps[i]] = tree[numNodes];
    }
}

// ------------------------------------------------------------------------------------------------ gatherRanges
/* buffer[t] = src[rangeOffsets[r] + t - rangeScan[r]] with r = the range that holds packed index t.  One warp
 * handles 32 consecutive packed indices: the search runs once per warp when they fall into one range (the common
 * case: halo ranges are whole leaves), per lane otherwise. */
template<class E, class I>
__global__ void __launch_bounds__(kPrimThreads)
    gatherRangesKernel(const I* __restrict__ rangeScan, const I* __restrict__ rangeOffsets, int numRanges,
                       const E* __restrict__ src, E* __restrict__ buffer, size_t bufferSize)
{
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t t0 = size_t(blockIdx.x) * blockDim.x + (threadIdx.x & ~31u); t0 < bufferSize; t0 += stride)
    {
        const size_t t = t0 + (threadIdx.x & 31u);
        // upper_bound(rangeScan, rangeScan + numRanges, t0) - 1
        int lo = 0, len = numRanges;
        while (len > 0)
        {
            const int half = len >> 1;
            const bool le  = size_t(rangeScan[lo + half]) <= t0;
            lo             = le ? lo + half + 1 : lo;
            len            = le ? len - half - 1 : half;
        }
        int r = lo - 1;
        if (t < bufferSize)
        {
            while (r + 1 < numRanges && size_t(rangeScan[r + 1]) <= t)
                ++r;
            buffer[t] = src[size_t(rangeOffsets[r]) + t - size_t(rangeScan[r])];
        }
    }
}

struct alignas(4) E4 { uint32_t v; };
struct alignas(8) E8 { uint2 v; };
struct alignas(4) E12 { uint32_t v[3]; };
struct alignas(16) E16 { uint4 v; };

template<class I>
int gatherRangesDispatch(const void* scan, const void* offsets, int numRanges, const void* src, void* buffer,
                         size_t bufferSize, int elemBytes, cudaStream_t s)
{
    const unsigned grid = gridFor(bufferSize, kPrimThreads, 8);
    const I* sc         = static_cast<const I*>(scan);
    const I* of         = static_cast<const I*>(offsets);
    switch (elemBytes)
    {
        case 4:
            gatherRangesKernel<E4, I><<<grid, kPrimThreads, 0, s>>>(sc, of, numRanges, static_cast<const E4*>(src),
                                                                     static_cast<E4*>(buffer), bufferSize);
            break;
        case 8:
            gatherRangesKernel<E8, I><<<grid, kPrimThreads, 0, s>>>(sc, of, numRanges, static_cast<const E8*>(src),
                                                                     static_cast<E8*>(buffer), bufferSize);
            break;
        case 12:
            gatherRangesKernel<E12, I><<<grid, kPrimThreads, 0, s>>>(sc, of, numRanges, static_cast<const E12*>(src),
                                                                      static_cast<E12*>(buffer), bufferSize);
            break;
        case 16:
            gatherRangesKernel<E16, I><<<grid, kPrimThreads, 0, s>>>(sc, of, numRanges, static_cast<const E16*>(src),
                                                                      static_cast<E16*>(buffer), bufferSize);
            break;
        default: return CSTONE_E_INVALID_ARGUMENT;
    }
    CSB_LAUNCHED();
    return 0;
}

template<class T>
int fillT(void* first, size_t n, const void* valueHost, cudaStream_t s)
{
    T v;
    memcpy(&v, valueHost, sizeof(T));
    fillKernel<T><<<gridFor(n, kPrimThreads, 8), kPrimThreads, 0, s>>>(static_cast<T*>(first), n, v);
    CSB_LAUNCHED();
    return 0;
}

template<class T>
int lowerBoundOne(const T* first, size_t n, T value, uint64_t* outHost, void* temp, cudaStream_t s)
{
    if (!outHost) return CSTONE_E_INVALID_ARGUMENT;
    if (n == 0)
    {
        *outHost = 0;
        return 0;
    }
    if (!first || !temp) return CSTONE_E_INVALID_ARGUMENT;
    lowerBoundOneKernel<T><<<1, 32, 0, s>>>(first, n, value, static_cast<unsigned long long*>(temp));
    CSB_LAUNCHED();
    CSB_CUDA(cudaMemcpyAsync(outHost, temp, 8, cudaMemcpyDeviceToHost, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    return 0;
}

template<class T, class I>
int lowerBoundMany(const T* first, size_t n, const T* values, size_t numValues, I* result, cudaStream_t s)
{
    if (numValues == 0) return 0;
    if ((n && !first) || !values || !result) return CSTONE_E_INVALID_ARGUMENT;
    lowerBoundKernel<T, I><<<gridFor(numValues, kPrimThreads, 8), kPrimThreads, 0, s>>>(first, n, values, numValues, result);
    CSB_LAUNCHED();
    return 0;
}

template<class Tin, class Tout>
int segmentMaxT(const Tin* in, const void* segments, int indexBytes, size_t numSegments, size_t hintTotal, Tout* out,
                cudaStream_t s)
{
    if (numSegments == 0) return 0;
    if (!in || !segments || !out || (indexBytes != 4 && indexBytes != 8)) return CSTONE_E_INVALID_ARGUMENT;
    const bool wide = hintTotal / numSegments >= 128; // long segments: a warp each
    if (indexBytes == 4)
    {
        const uint32_t* sg = static_cast<const uint32_t*>(segments);
        if (wide)
            segmentMaxWarpKernel<<<gridFor(numSegments * 32, kPrimThreads, 8), kPrimThreads, 0, s>>>(in, sg, numSegments, out);
        else segmentMaxKernel<<<gridFor(numSegments, kPrimThreads, 8), kPrimThreads, 0, s>>>(in, sg, numSegments, out);
    }
    else
    {
        const uint64_t* sg = static_cast<const uint64_t*>(segments);
        if (wide)
            segmentMaxWarpKernel<<<gridFor(numSegments * 32, kPrimThreads, 8), kPrimThreads, 0, s>>>(in, sg, numSegments, out);
        else segmentMaxKernel<<<gridFor(numSegments, kPrimThreads, 8), kPrimThreads, 0, s>>>(in, sg, numSegments, out);
    }
    CSB_LAUNCHED();
    return 0;
}

inline uint64_t alignUp256(uint64_t v) { return (v + 255u) & ~uint64_t(255); }

template<class K>
int sortKeys(K* keys, K* keyBuf, size_t n, void* temp, uint64_t tempBytes, uint64_t sortBytes, cudaStream_t s)
{
    if (n == 0) return 0;
    if (!keys || !keyBuf || !temp) return CSTONE_E_INVALID_ARGUMENT;
    const uint64_t valBytes = alignUp256(uint64_t(n) * 4);
    if (tempBytes < alignUp256(sortBytes) + 2 * valBytes) return CSTONE_E_TEMP_TOO_SMALL;
    char* base     = static_cast<char*>(temp);
    uint32_t* vals = reinterpret_cast<uint32_t*>(base + alignUp256(sortBytes));
    uint32_t* vbuf = reinterpret_cast<uint32_t*>(base + alignUp256(sortBytes) + valBytes);
    if constexpr (sizeof(K) == 4) return cstone_b200_sort_by_key_k32(keys, vals, keyBuf, vbuf, n, temp, sortBytes, 1, s);
    else return cstone_b200_sort_by_key_k64(keys, vals, keyBuf, vbuf, n, temp, sortBytes, 1, s);
}

} // namespace csb

extern "C" {

using namespace csb;

uint64_t cstone_b200_reduce_temp_bytes(void) { return uint64_t(kReduceBlocks + 1) * 8; }

int cstone_b200_fill(void* first, size_t n, const void* valueHost, int elemBytes, void* stream)
{
    if (n == 0) return 0;
    if (!first || !valueHost) return CSTONE_E_INVALID_ARGUMENT;
    switch (elemBytes)
    {
        case 1: return fillT<uint8_t>(first, n, valueHost, asStream(stream));
        case 4: return fillT<uint32_t>(first, n, valueHost, asStream(stream));
        case 8: return fillT<uint64_t>(first, n, valueHost, asStream(stream));
        default: return CSTONE_E_INVALID_ARGUMENT;
    }
}

int cstone_b200_scale_f32(float* first, size_t n, float value, void* stream)
{
    if (n == 0) return 0;
    if (!first) return CSTONE_E_INVALID_ARGUMENT;
    scaleKernel<float><<<gridFor(n, kPrimThreads, 8), kPrimThreads, 0, asStream(stream)>>>(first, n, value);
    CSB_LAUNCHED();
    return 0;
}
int cstone_b200_scale_f64(double* first, size_t n, double value, void* stream)
{
    if (n == 0) return 0;
    if (!first) return CSTONE_E_INVALID_ARGUMENT;
    scaleKernel<double><<<gridFor(n, kPrimThreads, 8), kPrimThreads, 0, asStream(stream)>>>(first, n, value);
    CSB_LAUNCHED();
    return 0;
}

int cstone_b200_increment_u32(const uint32_t* in, uint32_t* out, size_t n, uint32_t value, void* stream)
{
    if (n == 0) return 0;
    if (!in || !out) return CSTONE_E_INVALID_ARGUMENT;
    incrementKernel<uint32_t><<<gridFor(n, kPrimThreads, 8), kPrimThreads, 0, asStream(stream)>>>(in, out, n, value);
    CSB_LAUNCHED();
    return 0;
}
int cstone_b200_increment_u64(const uint64_t* in, uint64_t* out, size_t n, uint64_t value, void* stream)
{
    if (n == 0) return 0;
    if (!in || !out) return CSTONE_E_INVALID_ARGUMENT;
    incrementKernel<uint64_t><<<gridFor(n, kPrimThreads, 8), kPrimThreads, 0, asStream(stream)>>>(in, out, n, value);
    CSB_LAUNCHED();
    return 0;
}

int cstone_b200_max_norm_square_f32(const float* x, const float* y, const float* z, size_t n, float* outHost, void* temp,
                                    void* stream)
{
    if (n && (!x || !y || !z)) return CSTONE_E_INVALID_ARGUMENT;
    return reduceToHost<float, OpMax<float>>(LoadNormSq<float>{x, y, z}, n, 0.0f, 0.0f, outHost, temp, asStream(stream));
}
int cstone_b200_max_norm_square_f64(const double* x, const double* y, const double* z, size_t n, double* outHost,
                                    void* temp, void* stream)
{
    if (n && (!x || !y || !z)) return CSTONE_E_INVALID_ARGUMENT;
    return reduceToHost<double, OpMax<double>>(LoadNormSq<double>{x, y, z}, n, 0.0, 0.0, outHost, temp, asStream(stream));
}

int cstone_b200_reduce_sum_u32(const uint32_t* in, size_t n, uint64_t init, uint64_t* outHost, void* temp, void* stream)
{
    if (n && !in) return CSTONE_E_INVALID_ARGUMENT;
    unsigned long long r = 0;
    int st = reduceToHost<unsigned long long, OpSum>(LoadU32{in}, n, 0ull, (unsigned long long)init, &r, temp,
                                                     asStream(stream));
    if (st == 0 && outHost) *outHost = r;
    return outHost ? st : CSTONE_E_INVALID_ARGUMENT;
}

int cstone_b200_count_u32(const uint32_t* first, size_t n, uint32_t value, uint64_t* outHost, void* temp, void* stream)
{
    if (n && !first) return CSTONE_E_INVALID_ARGUMENT;
    unsigned long long r = 0;
    int st = reduceToHost<unsigned long long, OpSum>(LoadEquals<uint32_t>{first, value}, n, 0ull, 0ull, &r, temp,
                                                     asStream(stream));
    if (st == 0 && outHost) *outHost = r;
    return outHost ? st : CSTONE_E_INVALID_ARGUMENT;
}
int cstone_b200_count_u64(const uint64_t* first, size_t n, uint64_t value, uint64_t* outHost, void* temp, void* stream)
{
    if (n && !first) return CSTONE_E_INVALID_ARGUMENT;
    unsigned long long r = 0;
    int st = reduceToHost<unsigned long long, OpSum>(LoadEquals<uint64_t>{first, value}, n, 0ull, 0ull, &r, temp,
                                                     asStream(stream));
    if (st == 0 && outHost) *outHost = r;
    return outHost ? st : CSTONE_E_INVALID_ARGUMENT;
}

#define CSB_LOWER_BOUND_ONE(SFX, T)                                                                                    \
    int cstone_b200_lower_bound_##SFX(const T* first, size_t n, T value, uint64_t* outHost, void* temp, void* stream)  \
    {                                                                                                                  \
        return lowerBoundOne<T>(first, n, value, outHost, temp, asStream(stream));                                     \
    }
CSB_LOWER_BOUND_ONE(u32, uint32_t)
CSB_LOWER_BOUND_ONE(u64, uint64_t)
CSB_LOWER_BOUND_ONE(i32, int32_t)
CSB_LOWER_BOUND_ONE(i64, int64_t)
CSB_LOWER_BOUND_ONE(f32, float)

#define CSB_LOWER_BOUND_MANY(KS, K, IS, I)                                                                             \
    int cstone_b200_lower_bound_many_##KS##_##IS(const K* first, size_t n, const K* values, size_t numValues,          \
                                                 I* result, void* stream)                                              \
    {                                                                                                                  \
        return lowerBoundMany<K, I>(first, n, values, numValues, result, asStream(stream));                            \
    }
CSB_LOWER_BOUND_MANY(k32, uint32_t, u32, uint32_t)
CSB_LOWER_BOUND_MANY(k64, uint64_t, u32, uint32_t)
CSB_LOWER_BOUND_MANY(k32, uint32_t, u64, uint64_t)
CSB_LOWER_BOUND_MANY(k64, uint64_t, u64, uint64_t)

int cstone_b200_segment_max_f32_f32(const float* in, const void* segments, int indexBytes, size_t numSegments,
                                    size_t numElements, float* out, void* stream)
{
    return segmentMaxT<float, float>(in, segments, indexBytes, numSegments, numElements, out, asStream(stream));
}
int cstone_b200_segment_max_f64_f32(const double* in, const void* segments, int indexBytes, size_t numSegments,
                                    size_t numElements, float* out, void* stream)
{
    return segmentMaxT<double, float>(in, segments, indexBytes, numSegments, numElements, out, asStream(stream));
}
int cstone_b200_segment_max_f64_f64(const double* in, const void* segments, int indexBytes, size_t numSegments,
                                    size_t numElements, double* out, void* stream)
{
    return segmentMaxT<double, double>(in, segments, indexBytes, numSegments, numElements, out, asStream(stream));
}

uint64_t cstone_b200_sort_keys_temp_bytes_k32(uint64_t n)
{
    return alignUp256(cstone_b200_sort_by_key_temp_bytes_k32(n)) + 2 * alignUp256(n * 4);
}
uint64_t cstone_b200_sort_keys_temp_bytes_k64(uint64_t n)
{
    return alignUp256(cstone_b200_sort_by_key_temp_bytes_k64(n)) + 2 * alignUp256(n * 4);
}
int cstone_b200_sort_keys_k32(uint32_t* keys, uint32_t* keyBuf, size_t n, void* temp, uint64_t tempBytes, void* stream)
{
    return sortKeys<uint32_t>(keys, keyBuf, n, temp, tempBytes, cstone_b200_sort_by_key_temp_bytes_k32(n),
                              asStream(stream));
}
int cstone_b200_sort_keys_k64(uint64_t* keys, uint64_t* keyBuf, size_t n, void* temp, uint64_t tempBytes, void* stream)
{
    return sortKeys<uint64_t>(keys, keyBuf, n, temp, tempBytes, cstone_b200_sort_by_key_temp_bytes_k64(n),
                              asStream(stream));
}

int cstone_b200_count_sfc_gaps_k32(const uint32_t* tree, int32_t numNodes, int32_t* nodeOps, void* stream)
{
    if (numNodes <= 0) return 0;
    if (!tree || !nodeOps) return CSTONE_E_INVALID_ARGUMENT;
    countSfcGapsKernel<uint32_t><<<gridFor(size_t(numNodes), kPrimThreads, 8), kPrimThreads, 0, asStream(stream)>>>(
        tree, numNodes, nodeOps);
    CSB_LAUNCHED();
    return 0;
}
int cstone_b200_count_sfc_gaps_k64(const uint64_t* tree, int32_t numNodes, int32_t* nodeOps, void* stream)
{
    if (numNodes <= 0) return 0;
    if (!tree || !nodeOps) return CSTONE_E_INVALID_ARGUMENT;
    countSfcGapsKernel<uint64_t><<<gridFor(size_t(numNodes), kPrimThreads, 8), kPrimThreads, 0, asStream(stream)>>>(
        tree, numNodes, nodeOps);
    CSB_LAUNCHED();
    return 0;
}
int cstone_b200_fill_sfc_gaps_k32(const uint32_t* tree, int32_t numNodes, const int32_t* nodeOps, uint32_t* newTree,
                                  void* stream)
{
    if (numNodes < 0 || !tree || !nodeOps || !newTree) return CSTONE_E_INVALID_ARGUMENT;
    fillSfcGapsKernel<uint32_t><<<gridFor(size_t(numNodes) + 1, kPrimThreads, 8), kPrimThreads, 0, asStream(stream)>>>(
        tree, numNodes, nodeOps, newTree);
    CSB_LAUNCHED();
    return 0;
}
int cstone_b200_fill_sfc_gaps_k64(const uint64_t* tree, int32_t numNodes, const int32_t* nodeOps, uint64_t* newTree,
                                  void* stream)
{
    if (numNodes < 0 || !tree || !nodeOps || !newTree) return CSTONE_E_INVALID_ARGUMENT;
    fillSfcGapsKernel<uint64_t><<<gridFor(size_t(numNodes) + 1, kPrimThreads, 8), kPrimThreads, 0, asStream(stream)>>>(
        tree, numNodes, nodeOps, newTree);
    CSB_LAUNCHED();
    return 0;
}

int cstone_b200_gather_ranges(const void* rangeScan, const void* rangeOffsets, int indexBytes, int32_t numRanges,
                              const void* src, void* buffer, size_t bufferSize, int elemBytes, void* stream)
{
    if (bufferSize == 0) return 0;
    if (!rangeScan || !rangeOffsets || !src || !buffer || numRanges <= 0) return CSTONE_E_INVALID_ARGUMENT;
    if (indexBytes == 4)
        return gatherRangesDispatch<uint32_t>(rangeScan, rangeOffsets, numRanges, src, buffer, bufferSize, elemBytes,
                                              asStream(stream));
    if (indexBytes == 8)
        return gatherRangesDispatch<uint64_t>(rangeScan, rangeOffsets, numRanges, src, buffer, bufferSize, elemBytes,
                                              asStream(stream));
    return CSTONE_E_INVALID_ARGUMENT;
}

} // extern "C"
