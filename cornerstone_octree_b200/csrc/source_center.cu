/*
 * source_center.cu — n3 (first part): expansion (mass) centres of the tree nodes.
 *
 * Replaces computeLeafSourceCenterGpu / upsweepCentersGpu [ref: focus/source_center_gpu.h:51-74,
 * focus/source_center_gpu.cu:38-111]; semantics follow the CPU functions
 *   addBody / normalizeMass / massCenter     [ref: focus/source_center.hpp:44-78]
 *   CombineSourceCenter                      [ref: focus/source_center.hpp:81-96]
 *   computeLeafMassCenter                    [ref: focus/source_center.hpp:110-126]
 *
 * Parity: the accumulation order IS the result (floating point sums), so a leaf is summed by one thread, particle
 * after particle, and an internal node adds its eight children in index order; no contraction (-fmad=false), the
 * normalisation multiplies by 1/sum like the reference.  The particle streams are staged through shared memory with
 * coalesced loads (algorithmic bytes: 3 sizeof(Tc) + sizeof(Tm) per particle + 4 sizeof(Tf) per leaf).
 */
#include "common.cuh"

namespace csb
{

constexpr int kCenterThreads = 256;

template<class T>
__device__ __forceinline__ void addBody(T (&c)[4], T sx, T sy, T sz, T sm)
{
    T w = fabs(sm);
    c[0] += w * sx;
    c[1] += w * sy;
    c[2] += w * sz;
    c[3] += w;
}

template<class T>
__device__ __forceinline__ void normalizeMass(T (&c)[4])
{
    T invM = (c[3] != T(0.0)) ? T(1.0) / c[3] : T(1.0);
    c[0] *= invM;
    c[1] *= invM;
    c[2] *= invM;
}

template<class T>
struct alignas(16) Quad
{
    T v[4];
};

constexpr int kLeafThreads = 256;  // leaves per block
constexpr int kCenterChunk = 1024; // particles staged per round: 4 x 1024 x sizeof(Tf) = 32 KB for double

/* One thread per leaf, 256 consecutive leaves per block.  The particles of those leaves are one contiguous run: the
 * block streams it through shared memory in chunks (coalesced loads; the products |m| x, |m| y, |m| z are formed
 * there, by all threads), and every thread adds the part of its leaf that lies in the chunk, in particle order — the
 * running sums live in registers across chunks (and the next chunk is already in flight to registers while this one
 * is summed), so the summation order is exactly the sequential one of massCenter()
 * and every product / sum is the same IEEE operation (no contraction). */
template<class Tc, class Tm, class Tf>
__global__ void __launch_bounds__(kLeafThreads)
    leafSourceCenterKernel(const Tc* __restrict__ x, const Tc* __restrict__ y, const Tc* __restrict__ z,
                           const Tm* __restrict__ m, const int* __restrict__ leafToInternal, int numLeaves,
                           const uint32_t* __restrict__ layout, Quad<Tf>* __restrict__ centers)
{
    extern __shared__ __align__(16) unsigned char sRaw[];
    Tf* sx = reinterpret_cast<Tf*>(sRaw);
    Tf* sy = sx + kCenterChunk;
    Tf* sz = sy + kCenterChunk;
    Tf* sw = sz + kCenterChunk;
    for (int leafBase = blockIdx.x * kLeafThreads; leafBase < numLeaves; leafBase += gridDim.x * kLeafThreads)
    {
        const int nb         = min(kLeafThreads, numLeaves - leafBase);
        const int leaf       = leafBase + int(threadIdx.x);
        const bool mine      = int(threadIdx.x) < nb;
        const uint32_t p0    = layout[leafBase], p1 = layout[leafBase + nb];
        const uint32_t first = mine ? layout[leaf] : p1, last = mine ? layout[leaf + 1] : p1;
        Tf c0 = 0, c1 = 0, c2 = 0, c3 = 0;
        // register-staged pipeline: the loads of chunk k+1 are in flight while the leaves of chunk k are summed
        constexpr int kPer = kCenterChunk / kLeafThreads;
        Tc rx[kPer], ry[kPer], rz[kPer];
        Tm rm[kPer];
        auto fetch = [&](uint32_t start)
        {
#pragma unroll
            for (int j = 0; j < kPer; ++j)
            {
                const uint32_t k = start + threadIdx.x + uint32_t(j) * kLeafThreads;
                if (k < p1) { rx[j] = x[k], ry[j] = y[k], rz[j] = z[k], rm[j] = m[k]; }
            }
        };
        if (p0 < p1) fetch(p0);
        for (uint32_t start = p0; start < p1; start += kCenterChunk)
        {
            const uint32_t len = min(uint32_t(kCenterChunk), p1 - start);
            __syncthreads();
#pragma unroll
            for (int j = 0; j < kPer; ++j)
            {
                const uint32_t k = threadIdx.x + uint32_t(j) * kLeafThreads;
                if (k < len)
                {
                    const Tf w = fabs(Tf(rm[j]));
                    sx[k] = w * Tf(rx[j]), sy[k] = w * Tf(ry[j]), sz[k] = w * Tf(rz[j]), sw[k] = w;
                }
            }
            __syncthreads();
            if (start + kCenterChunk < p1) fetch(start + kCenterChunk);
            const uint32_t a = max(first, start), b = min(last, start + len);
            for (uint32_t i = a; i < b; ++i)
            {
                c0 += sx[i - start], c1 += sy[i - start], c2 += sz[i - start], c3 += sw[i - start];
            }
        }
        if (mine)
        {
            Tf c[4] = {c0, c1, c2, c3};
            normalizeMass(c);
            Quad<Tf> q;
            q.v[0] = c[0], q.v[1] = c[1], q.v[2] = c[2], q.v[3] = c[3];
            centers[leafToInternal[leaf]] = q;
        }
    }
}

template<class T>
__global__ void __launch_bounds__(kCenterThreads)
    upsweepCentersKernel(int first, int last, const int* __restrict__ childOffsets, Quad<T>* centers)
{
    for (int cell = first + blockIdx.x * blockDim.x + threadIdx.x; cell < last; cell += gridDim.x * blockDim.x)
    {
        int child = childOffsets[cell];
        if (!child) continue;
        T c[4] = {0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < 8; ++k)
        {
            Quad<T> s = centers[child + k];
            addBody<T>(c, s.v[0], s.v[1], s.v[2], s.v[3]);
        }
        normalizeMass(c);
        Quad<T> q;
        q.v[0] = c[0], q.v[1] = c[1], q.v[2] = c[2], q.v[3] = c[3];
        centers[cell] = q;
    }
}

template<class Tc, class Tm, class Tf>
int leafSourceCenter(const Tc* x, const Tc* y, const Tc* z, const Tm* m, const int* l2i, int numLeaves,
                     const uint32_t* layout, Tf* centers, void* stream)
{
    if (numLeaves <= 0) return 0;
    if (!x || !y || !z || !m || !l2i || !layout || !centers) return CSTONE_E_INVALID_ARGUMENT;
    constexpr int smem = 4 * kCenterChunk * int(sizeof(Tf));
    static bool attr   = [] {
        return cudaFuncSetAttribute(leafSourceCenterKernel<Tc, Tm, Tf>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    smem) == cudaSuccess;
    }();
    if (!attr) return int(cudaErrorInvalidValue);
    leafSourceCenterKernel<Tc, Tm, Tf><<<gridFor(size_t(numLeaves), kLeafThreads, 6), kLeafThreads, smem,
                                         asStream(stream)>>>(x, y, z, m, l2i, numLeaves, layout,
                                                             reinterpret_cast<Quad<Tf>*>(centers));
    CSB_LAUNCHED();
    return 0;
}

template<class T>
int upsweepCenters(int numLevels, const int* levelRangeHost, const int* childOffsets, T* centers, void* stream)
{
    if (numLevels <= 0) return 0;
    if (!levelRangeHost || !childOffsets || !centers) return CSTONE_E_INVALID_ARGUMENT;
    for (int level = numLevels - 1; level >= 0; --level)
    {
        int first = levelRangeHost[level], last = levelRangeHost[level + 1];
        if (last <= first) continue;
        upsweepCentersKernel<T><<<gridFor(size_t(last - first), kCenterThreads, 8), kCenterThreads, 0,
                                  asStream(stream)>>>(first, last, childOffsets, reinterpret_cast<Quad<T>*>(centers));
        CSB_LAUNCHED();
    }
    return 0;
}

} // namespace csb

extern "C" {

int cstone_b200_compute_leaf_source_center_f64_f64_f64(const double* x, const double* y, const double* z,
                                                       const double* m, const int32_t* l2i, int32_t numLeaves,
                                                       const uint32_t* layout, double* centers4, void* stream)
{
    return csb::leafSourceCenter<double, double, double>(x, y, z, m, l2i, numLeaves, layout, centers4, stream);
}
int cstone_b200_compute_leaf_source_center_f64_f32_f64(const double* x, const double* y, const double* z,
                                                       const float* m, const int32_t* l2i, int32_t numLeaves,
                                                       const uint32_t* layout, double* centers4, void* stream)
{
    return csb::leafSourceCenter<double, float, double>(x, y, z, m, l2i, numLeaves, layout, centers4, stream);
}
int cstone_b200_compute_leaf_source_center_f32_f32_f32(const float* x, const float* y, const float* z, const float* m,
                                                       const int32_t* l2i, int32_t numLeaves, const uint32_t* layout,
                                                       float* centers4, void* stream)
{
    return csb::leafSourceCenter<float, float, float>(x, y, z, m, l2i, numLeaves, layout, centers4, stream);
}
int cstone_b200_upsweep_centers_f32(int32_t numLevels, const int32_t* levelRangeHost, const int32_t* childOffsets,
                                    float* centers4, void* stream)
{
    return csb::upsweepCenters<float>(numLevels, levelRangeHost, childOffsets, centers4, stream);
}
int cstone_b200_upsweep_centers_f64(int32_t numLevels, const int32_t* levelRangeHost, const int32_t* childOffsets,
                                    double* centers4, void* stream)
{
    return csb::upsweepCenters<double>(numLevels, levelRangeHost, childOffsets, centers4, stream);
}

} // extern "C"
