/*
 * gather.cu — a5: gather / scatter through the SFC permutation, a3: fused coordinate extrema.
 *
 * Replaces gatherGpu / scatterGpu [ref: primitives/primitives_gpu.h:48-56, primitives_gpu.cu:109-165] and
 * MinMaxGpu [ref: primitives/primitives_gpu.h:58-62, primitives_gpu.cu:177-187].
 * gather_multi moves up to 16 arrays with one read of the map: 4 + F*2e bytes per particle instead of the
 * reference's F*(4 + 2e) (one kernel per array, gatherArrays domain/layout.hpp:202-239).
 */
#include "common.cuh"

namespace csb
{

template<int BYTES>
struct Blob
{
    typedef uint32_t word;
    word w[BYTES / 4];
};
template<>
struct alignas(8) Blob<8>
{
    uint64_t w[1];
};
template<>
struct alignas(16) Blob<16>
{
    uint64_t w[2];
};
template<>
struct alignas(8) Blob<24>
{
    uint64_t w[3];
};
template<>
struct alignas(16) Blob<32>
{
    uint64_t w[4];
};

template<class E, bool Scatter>
__global__ void __launch_bounds__(256)
    permuteKernel(const uint32_t* __restrict__ map, size_t n, const E* __restrict__ src, E* __restrict__ dst)
{
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        if (Scatter) dst[map[i]] = src[i];
        else dst[i] = src[map[i]];
    }
}

template<bool Scatter>
int permute(const uint32_t* map, size_t n, const void* src, void* dst, int elemBytes, cudaStream_t s)
{
    if (n == 0) return 0;
    if (!map || !src || !dst) return CSTONE_E_INVALID_ARGUMENT;
    unsigned grid = gridFor(n, 256, 8);
#define CSB_PERMUTE(B)                                                                                                 \
    case B:                                                                                                            \
        permuteKernel<Blob<B>, Scatter><<<grid, 256, 0, s>>>(map, n, (const Blob<B>*)src, (Blob<B>*)dst);              \
        break;
    switch (elemBytes)
    {
        CSB_PERMUTE(4)
        CSB_PERMUTE(8)
        CSB_PERMUTE(12)
        CSB_PERMUTE(16)
        CSB_PERMUTE(24)
        CSB_PERMUTE(32)
        default: return CSTONE_E_INVALID_ARGUMENT;
    }
#undef CSB_PERMUTE
    CSB_LAUNCHED();
    return 0;
}

constexpr int kMaxGatherArrays = 16;
struct GatherArgs
{
    const void* src[kMaxGatherArrays];
    void* dst[kMaxGatherArrays];
    int n4;      // number of arrays with 4-byte elements (listed first)
    int n8;      // number of arrays with 8-byte elements (listed after the 4-byte ones)
};

__global__ void __launch_bounds__(256) gatherMultiKernel(const uint32_t* __restrict__ map, size_t n, GatherArgs a)
{
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        uint32_t j = map[i];
        uint32_t v4[kMaxGatherArrays];
        uint64_t v8[kMaxGatherArrays];
        // all loads first (independent, in flight together), then all stores
#pragma unroll
        for (int k = 0; k < kMaxGatherArrays; ++k)
            if (k < a.n4) v4[k] = reinterpret_cast<const uint32_t*>(a.src[k])[j];
#pragma unroll
        for (int k = 0; k < kMaxGatherArrays; ++k)
            if (k < a.n8) v8[k] = reinterpret_cast<const uint64_t*>(a.src[a.n4 + k])[j];
#pragma unroll
        for (int k = 0; k < kMaxGatherArrays; ++k)
            if (k < a.n4) reinterpret_cast<uint32_t*>(a.dst[k])[i] = v4[k];
#pragma unroll
        for (int k = 0; k < kMaxGatherArrays; ++k)
            if (k < a.n8) reinterpret_cast<uint64_t*>(a.dst[a.n4 + k])[i] = v8[k];
    }
}

/* Common case (coordinates + properties in double): N8 arrays of 8-byte elements, four consecutive elements per
 * thread.  One 16-byte read of the map, 4*N8 independent loads in flight per thread (the single-element kernel
 * above reaches 3.2 TB/s on a sorted map because its loads wait for the map), two 16-byte stores per array. */
template<int N8>
__global__ void __launch_bounds__(256) gather8x4Kernel(const uint32_t* __restrict__ map, size_t n, GatherArgs a)
{
    const size_t nq     = n / 4;
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    const size_t t0     = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    for (size_t q = t0; q < nq; q += stride)
    {
        const uint4 j = reinterpret_cast<const uint4*>(map)[q];
        uint64_t v[N8][4];
#pragma unroll
        for (int k = 0; k < N8; ++k)
        {
            const uint64_t* src = reinterpret_cast<const uint64_t*>(a.src[k]);
            v[k][0] = src[j.x], v[k][1] = src[j.y], v[k][2] = src[j.z], v[k][3] = src[j.w];
        }
#pragma unroll
        for (int k = 0; k < N8; ++k)
        {
            ulonglong2* dst = reinterpret_cast<ulonglong2*>(a.dst[k]) + 2 * q;
            dst[0]          = make_ulonglong2(v[k][0], v[k][1]);
            dst[1]          = make_ulonglong2(v[k][2], v[k][3]);
        }
    }
    const size_t i = 4 * nq + t0; // tail
    if (i < n)
    {
        const uint32_t j = map[i];
#pragma unroll
        for (int k = 0; k < N8; ++k)
            reinterpret_cast<uint64_t*>(a.dst[k])[i] = reinterpret_cast<const uint64_t*>(a.src[k])[j];
    }
}

// ------------------------------------------------------------------------------------------------
// a3 extrema of up to three arrays in one pass; comparisons only, so the result is exact
// ------------------------------------------------------------------------------------------------
template<class T>
__device__ inline T warpMin(T v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template<class T>
__device__ inline T warpMax(T v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

constexpr int kMinMaxBlocks = kNumSMs * 4;

template<class T>
__global__ void __launch_bounds__(256)
    minMaxPartialKernel(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z, size_t n,
                        T* __restrict__ partial /* [gridDim.x][6] */)
{
    const T* arr[3] = {x, y, z};
    __shared__ T sRed[8][6];
    T lo[3], hi[3];
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        lo[d] = arr[d] ? arr[d][0] : T(0);
        hi[d] = lo[d];
    }
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    size_t i            = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    // four strided elements of each array per trip: 12 independent loads in flight per thread (a streaming reduction
    // with one load per trip reached 55 % of the HBM peak)
    for (; i + 3 * stride < n; i += 4 * stride)
    {
        T v[3][4];
#pragma unroll
        for (int d = 0; d < 3; ++d)
            if (arr[d])
            {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    v[d][u] = arr[d][i + u * stride];
            }
#pragma unroll
        for (int d = 0; d < 3; ++d)
            if (arr[d])
            {
                lo[d] = min(min(lo[d], min(v[d][0], v[d][1])), min(v[d][2], v[d][3]));
                hi[d] = max(max(hi[d], max(v[d][0], v[d][1])), max(v[d][2], v[d][3]));
            }
    }
    for (; i < n; i += stride)
    {
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            if (arr[d])
            {
                T v   = arr[d][i];
                lo[d] = min(lo[d], v);
                hi[d] = max(hi[d], v);
            }
        }
    }
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        lo[d] = warpMin(lo[d]);
        hi[d] = warpMax(hi[d]);
    }
    if (laneId() == 0)
    {
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            sRed[threadIdx.x >> 5][2 * d]     = lo[d];
            sRed[threadIdx.x >> 5][2 * d + 1] = hi[d];
        }
    }
    __syncthreads();
    if (threadIdx.x < 6)
    {
        T v = sRed[0][threadIdx.x];
        for (int w = 1; w < 8; ++w)
            v = (threadIdx.x & 1) ? max(v, sRed[w][threadIdx.x]) : min(v, sRed[w][threadIdx.x]);
        partial[blockIdx.x * 6 + threadIdx.x] = v;
    }
}

template<class T>
__global__ void minMaxFinalKernel(T* __restrict__ partial, int numPartials)
{
    // 6 warps, one per output value
    int q = threadIdx.x >> 5;
    T v   = partial[q];
    for (int i = int(laneId()); i < numPartials; i += 32)
    {
        T p = partial[i * 6 + q];
        v   = (q & 1) ? max(v, p) : min(v, p);
    }
    v = (q & 1) ? warpMax(v) : warpMin(v);
    __syncthreads();
    if (laneId() == 0) partial[q] = v;
}

template<class T>
int minMax3(const T* x, const T* y, const T* z, size_t n, T* out, void* temp, cudaStream_t s)
{
    if (n == 0 || !out || !temp) return CSTONE_E_INVALID_ARGUMENT;
    T* partial  = reinterpret_cast<T*>(temp);
    int blocks  = int(gridFor(n, 256 * 4, 4));
    minMaxPartialKernel<T><<<blocks, 256, 0, s>>>(x, y, z, n, partial);
    CSB_LAUNCHED();
    minMaxFinalKernel<T><<<1, 192, 0, s>>>(partial, blocks);
    CSB_LAUNCHED();
    CSB_CUDA(cudaMemcpyAsync(out, partial, 6 * sizeof(T), cudaMemcpyDeviceToHost, s));
    CSB_CUDA(cudaStreamSynchronize(s));
    return 0;
}

} // namespace csb

extern "C" {

int cstone_b200_gather(const uint32_t* map, size_t n, const void* src, void* dst, int elemBytes, void* stream)
{
    return csb::permute<false>(map, n, src, dst, elemBytes, csb::asStream(stream));
}
int cstone_b200_scatter(const uint32_t* map, size_t n, const void* src, void* dst, int elemBytes, void* stream)
{
    return csb::permute<true>(map, n, src, dst, elemBytes, csb::asStream(stream));
}

int cstone_b200_gather_multi(const uint32_t* map, size_t n, int numArrays, const void* const* src, void* const* dst,
                             const int* elemBytes, void* stream)
{
    if (n == 0 || numArrays == 0) return 0;
    if (numArrays < 0 || numArrays > csb::kMaxGatherArrays || !map || !src || !dst || !elemBytes)
        return CSTONE_E_INVALID_ARGUMENT;
    csb::GatherArgs a{};
    int pos = 0;
    for (int pass = 0; pass < 2; ++pass)
    {
        int want = pass == 0 ? 4 : 8;
        for (int k = 0; k < numArrays; ++k)
        {
            if (elemBytes[k] != 4 && elemBytes[k] != 8)
            {
                // wider rows (12 - 32 bytes, util::array properties): one single-array gather each, in the first pass
                if (pass == 0)
                {
                    if (!src[k] || !dst[k]) return CSTONE_E_INVALID_ARGUMENT;
                    int st = csb::permute<false>(map, n, src[k], dst[k], elemBytes[k], csb::asStream(stream));
                    if (st) return st;
                }
                continue;
            }
            if (elemBytes[k] != want) continue;
            if (!src[k] || !dst[k]) return CSTONE_E_INVALID_ARGUMENT;
            a.src[pos] = src[k];
            a.dst[pos] = dst[k];
            ++pos;
            (pass == 0 ? a.n4 : a.n8)++;
        }
    }
    if (pos == 0) return 0;
    bool aligned = (reinterpret_cast<uintptr_t>(map) & 15u) == 0;
    for (int k = 0; k < a.n8; ++k)
        aligned = aligned && (reinterpret_cast<uintptr_t>(a.dst[k]) & 15u) == 0;
    cudaStream_t s = csb::asStream(stream);
    if (aligned && a.n4 == 0 && a.n8 <= 8 && n >= 4096)
    {
        const unsigned grid = csb::gridFor(n / 4, 256, 4);
        switch (a.n8)
        {
            case 1: csb::gather8x4Kernel<1><<<grid, 256, 0, s>>>(map, n, a); break;
            case 2: csb::gather8x4Kernel<2><<<grid, 256, 0, s>>>(map, n, a); break;
            case 3: csb::gather8x4Kernel<3><<<grid, 256, 0, s>>>(map, n, a); break;
            case 4: csb::gather8x4Kernel<4><<<grid, 256, 0, s>>>(map, n, a); break;
            case 5: csb::gather8x4Kernel<5><<<grid, 256, 0, s>>>(map, n, a); break;
            case 6: csb::gather8x4Kernel<6><<<grid, 256, 0, s>>>(map, n, a); break;
            case 7: csb::gather8x4Kernel<7><<<grid, 256, 0, s>>>(map, n, a); break;
            default: csb::gather8x4Kernel<8><<<grid, 256, 0, s>>>(map, n, a); break;
        }
    }
    else { csb::gatherMultiKernel<<<csb::gridFor(n, 256, 8), 256, 0, s>>>(map, n, a); }
    CSB_LAUNCHED();
    return 0;
}

uint64_t cstone_b200_minmax_temp_bytes(void) { return uint64_t(csb::kMinMaxBlocks + 1) * 6 * sizeof(double); }

int cstone_b200_minmax3_f32(const float* x, const float* y, const float* z, size_t n, float* out, void* temp,
                            void* stream)
{
    return csb::minMax3<float>(x, y, z, n, out, temp, csb::asStream(stream));
}
int cstone_b200_minmax3_f64(const double* x, const double* y, const double* z, size_t n, double* out, void* temp,
                            void* stream)
{
    return csb::minMax3<double>(x, y, z, n, out, temp, csb::asStream(stream));
}

} // extern "C"
