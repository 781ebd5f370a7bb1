/*
 * sfc_keys.cu — a1/a2: SFC keys from coordinates, integer encode/decode.
 *
 * Replaces computeSfcKeysGpu [ref: sfc/sfc_gpu.h:38-39, sfc/sfc_gpu.cu:38-57] and follows the arithmetic of
 * sfc3D [ref: sfc/sfc.hpp:157-194]:  ix = int(floor(x*mx) - xmin*mx), mx = 2^maxLevel * (1/lx), in type T with
 * NO fused multiply-add (the library is compiled with -fmad=false; the intrinsics below make it explicit),
 * truncation, clamp to 2^maxLevel-1.  Keys equal to removeKey = 2^(3*maxLevel) are left untouched.
 *
 * B200 design: pure streaming kernel, 3*sizeof(T) + 2*sizeof(K) algorithmic bytes per particle (40 B for
 * u64/double).  Each thread handles VEC consecutive particles with 16-byte vector loads/stores; the Hilbert
 * curve is evaluated with a 3-levels-per-step state-machine table staged in shared memory (7 look-ups for a
 * 63-bit key) so that integer work stays well below the HBM time.
 */
#include "common.cuh"

namespace csb
{

template<class T, int N>
struct alignas(sizeof(T) * N) Pack
{
    T v[N];
};

__device__ inline float mulNoFma(float a, float b) { return __fmul_rn(a, b); }
__device__ inline double mulNoFma(double a, double b) { return __dmul_rn(a, b); }
__device__ inline float subNoFma(float a, float b) { return __fsub_rn(a, b); }
__device__ inline double subNoFma(double a, double b) { return __dsub_rn(a, b); }

template<class K, class T>
__device__ inline K keyOf(int kind, const uint16_t* enc, T x, T y, T z, const T* lo, const T* m)
{
    constexpr int mcoord = (1 << KeyTraits<K>::maxLevel) - 1;
    int ix               = int(subNoFma(floor(mulNoFma(x, m[0])), mulNoFma(lo[0], m[0])));
    int iy               = int(subNoFma(floor(mulNoFma(y, m[1])), mulNoFma(lo[1], m[1])));
    int iz               = int(subNoFma(floor(mulNoFma(z, m[2])), mulNoFma(lo[2], m[2])));
    ix                   = min(ix, mcoord);
    iy                   = min(iy, mcoord);
    iz                   = min(iz, mcoord);
    return sfcEncode<K>(kind, enc, uint32_t(ix), uint32_t(iy), uint32_t(iz));
}

template<class T>
struct KeyParams
{
    T lo[3];
    T m[3];
};

template<class K, class T, int VEC>
__global__ void __launch_bounds__(256) sfcKeysKernel(int kind,
                                                      const uint16_t* __restrict__ encTab,
                                                      const T* __restrict__ x,
                                                      const T* __restrict__ y,
                                                      const T* __restrict__ z,
                                                      K* __restrict__ keys,
                                                      size_t n,
                                                      KeyParams<T> p)
{
    __shared__ __align__(16) uint16_t tab[kHilbertTabSize];
    if (kind == CSTONE_SFC_HILBERT)
    {
        loadTable(tab, encTab);
        __syncthreads();
    }
    constexpr K removeKey = K(1) << (3 * KeyTraits<K>::maxLevel);
    typedef Pack<T, VEC> TV;
    typedef Pack<K, VEC> KV;

    size_t numVec = n / VEC;
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t v = size_t(blockIdx.x) * blockDim.x + threadIdx.x; v < numVec; v += stride)
    {
        TV xv = reinterpret_cast<const TV*>(x)[v];
        TV yv = reinterpret_cast<const TV*>(y)[v];
        TV zv = reinterpret_cast<const TV*>(z)[v];
        KV kv = reinterpret_cast<const KV*>(keys)[v];
#pragma unroll
        for (int i = 0; i < VEC; ++i)
        {
            K k = keyOf<K, T>(kind, tab, xv.v[i], yv.v[i], zv.v[i], p.lo, p.m);
            if (kv.v[i] != removeKey) kv.v[i] = k;
        }
        reinterpret_cast<KV*>(keys)[v] = kv;
    }
    // tail (n % VEC elements) by the first threads of block 0
    if (blockIdx.x == 0 && threadIdx.x < n - numVec * VEC)
    {
        size_t i = numVec * VEC + threadIdx.x;
        if (keys[i] != removeKey) keys[i] = keyOf<K, T>(kind, tab, x[i], y[i], z[i], p.lo, p.m);
    }
}

template<class K>
__global__ void __launch_bounds__(256) encodeKernel(int kind,
                                                     const uint16_t* __restrict__ encTab,
                                                     const uint32_t* __restrict__ ix,
                                                     const uint32_t* __restrict__ iy,
                                                     const uint32_t* __restrict__ iz,
                                                     K* __restrict__ keys,
                                                     size_t n)
{
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        keys[i] = sfcEncode<K>(kind, encTab, ix[i], iy[i], iz[i]);
}

template<class K>
__global__ void __launch_bounds__(256) decodeKernel(int kind,
                                                     const uint16_t* __restrict__ decTab,
                                                     const K* __restrict__ keys,
                                                     uint32_t* __restrict__ ix,
                                                     uint32_t* __restrict__ iy,
                                                     uint32_t* __restrict__ iz,
                                                     size_t n)
{
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        uint32_t a, b, c;
        if (kind == CSTONE_SFC_MORTON) mortonDecode<K>(keys[i], a, b, c);
        else hilbertDecode<K>(decTab, keys[i], KeyTraits<K>::maxLevel, a, b, c);
        ix[i] = a, iy[i] = b, iz[i] = c;
    }
}

template<class A>
inline bool aligned(const A* p, size_t bytes)
{
    return (reinterpret_cast<uintptr_t>(p) % bytes) == 0;
}

template<class K, class T>
int computeSfcKeys(int kind, const T* x, const T* y, const T* z, K* keys, size_t n,
                   const typename BoxT<T>::type* box, void* stream)
{
    if (kind != CSTONE_SFC_MORTON && kind != CSTONE_SFC_HILBERT) return CSTONE_E_INVALID_ARGUMENT;
    if (n == 0) return 0;
    if (!x || !y || !z || !keys || !box) return CSTONE_E_INVALID_ARGUMENT;
    SfcTables tabs;
    int st = getSfcTables(&tabs);
    if (st) return st;

    KeyParams<T> p;
    const unsigned cube = 1u << KeyTraits<K>::maxLevel;
    for (int d = 0; d < 3; ++d)
    {
        p.lo[d] = box->limits[2 * d];
        // [ref: sfc/sfc.hpp:190-193] cubeLength * box.ilx(): unsigned converted to T, one rounding
        volatile T m = T(cube) * box->inv_lengths[d];
        p.m[d]       = m;
    }

    constexpr int VEC = 16 / sizeof(T);
    cudaStream_t s    = asStream(stream);
    bool vecOk = aligned(x, sizeof(T) * VEC) && aligned(y, sizeof(T) * VEC) && aligned(z, sizeof(T) * VEC) &&
                 aligned(keys, sizeof(K) * VEC);
    if (vecOk)
    {
        unsigned grid = gridFor(n / VEC + 1, 256, 4);
        sfcKeysKernel<K, T, VEC><<<grid, 256, 0, s>>>(kind, tabs.enc, x, y, z, keys, n, p);
    }
    else
    {
        unsigned grid = gridFor(n, 256, 4);
        sfcKeysKernel<K, T, 1><<<grid, 256, 0, s>>>(kind, tabs.enc, x, y, z, keys, n, p);
    }
    CSB_LAUNCHED();
    return 0;
}

template<class K>
int encodeSfc(int kind, const uint32_t* ix, const uint32_t* iy, const uint32_t* iz, K* keys, size_t n, void* stream)
{
    if (n == 0) return 0;
    SfcTables tabs;
    int st = getSfcTables(&tabs);
    if (st) return st;
    encodeKernel<K><<<gridFor(n, 256), 256, 0, asStream(stream)>>>(kind, tabs.enc, ix, iy, iz, keys, n);
    CSB_LAUNCHED();
    return 0;
}

template<class K>
int decodeSfc(int kind, const K* keys, uint32_t* ix, uint32_t* iy, uint32_t* iz, size_t n, void* stream)
{
    if (n == 0) return 0;
    SfcTables tabs;
    int st = getSfcTables(&tabs);
    if (st) return st;
    decodeKernel<K><<<gridFor(n, 256), 256, 0, asStream(stream)>>>(kind, tabs.dec, keys, ix, iy, iz, n);
    CSB_LAUNCHED();
    return 0;
}

} // namespace csb

extern "C" {

int cstone_b200_compute_sfc_keys_k32_f32(int kind, const float* x, const float* y, const float* z, uint32_t* keys,
                                         size_t n, const cstone_box_f32* box, void* stream)
{
    return csb::computeSfcKeys<uint32_t, float>(kind, x, y, z, keys, n, box, stream);
}
int cstone_b200_compute_sfc_keys_k32_f64(int kind, const double* x, const double* y, const double* z, uint32_t* keys,
                                         size_t n, const cstone_box_f64* box, void* stream)
{
    return csb::computeSfcKeys<uint32_t, double>(kind, x, y, z, keys, n, box, stream);
}
int cstone_b200_compute_sfc_keys_k64_f32(int kind, const float* x, const float* y, const float* z, uint64_t* keys,
                                         size_t n, const cstone_box_f32* box, void* stream)
{
    return csb::computeSfcKeys<uint64_t, float>(kind, x, y, z, keys, n, box, stream);
}
int cstone_b200_compute_sfc_keys_k64_f64(int kind, const double* x, const double* y, const double* z, uint64_t* keys,
                                         size_t n, const cstone_box_f64* box, void* stream)
{
    return csb::computeSfcKeys<uint64_t, double>(kind, x, y, z, keys, n, box, stream);
}

int cstone_b200_encode_sfc_k32(int kind, const uint32_t* ix, const uint32_t* iy, const uint32_t* iz, uint32_t* keys,
                               size_t n, void* stream)
{
    return csb::encodeSfc<uint32_t>(kind, ix, iy, iz, keys, n, stream);
}
int cstone_b200_encode_sfc_k64(int kind, const uint32_t* ix, const uint32_t* iy, const uint32_t* iz, uint64_t* keys,
                               size_t n, void* stream)
{
    return csb::encodeSfc<uint64_t>(kind, ix, iy, iz, keys, n, stream);
}
int cstone_b200_decode_sfc_k32(int kind, const uint32_t* keys, uint32_t* ix, uint32_t* iy, uint32_t* iz, size_t n,
                               void* stream)
{
    return csb::decodeSfc<uint32_t>(kind, keys, ix, iy, iz, n, stream);
}
int cstone_b200_decode_sfc_k64(int kind, const uint64_t* keys, uint32_t* ix, uint32_t* iy, uint32_t* iz, size_t n,
                               void* stream)
{
    return csb::decodeSfc<uint64_t>(kind, keys, ix, iy, iz, n, stream);
}

} // extern "C"
