/*
 * cstone_b200_shim.cpp — the reference-side binding: definitions of the reference's `extern` GPU function templates
 * (the ones its static library `cstone_gpu` instantiates in *.cu files) as thin forwards into the C ABI of
 * libcstone_b200.so.  A maintainer drops this ONE translation unit in place of
 *     sfc/sfc_gpu.cu  primitives/primitives_gpu.cu  tree/csarray_gpu.cu  tree/octree_gpu.cu
 *     traversal/collisions_gpu.cu  halos/gather_halos_gpu.cu
 * and links -lcstone_b200; everything above (Domain, Halos, GlobalAssignment, FocusedOctree, ...) keeps calling the
 * same names with the same signatures.
 *
 * Build (host compiler only, no nvcc, no CUDA headers):
 *     g++ -std=c++20 -DUSE_CUDA -I<reference>/include -I<repo>/include -c integration/cstone_b200_shim.cpp
 * tests/test_integration_shim.py does exactly that against /root/reference and checks the mangled symbols;
 * integration/shim_kat.cpp calls the pipeline through these names on a GPU.
 *
 * Every explicit instantiation below mirrors one of the reference's (file:line cited).  Instantiations the reference
 * makes for types that are not on the octree / domain path are listed at the end with the reason they are absent.
 */
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <tuple>

#include "cstone/sfc/sfc_gpu.h"
#include "cstone/primitives/primitives_gpu.h"
#include "cstone/tree/csarray_gpu.h"
#include "cstone/tree/octree_gpu.h"
#include "cstone/traversal/collisions_gpu.h"
#include "cstone/halos/gather_halos_gpu.h"

#include "cstone_b200.h"

namespace cstone
{

namespace b200
{

//! the reference's error behaviour: CUDA errors print and exit (cuda/errorcheck.cuh:30-40), logic errors throw
static void check(int st, const char* what)
{
    if (st > 0)
    {
        std::fprintf(stderr, "%s: %s\n", what, cstone_b200_status_string(st));
        std::exit(EXIT_FAILURE);
    }
    if (st < 0) throw std::runtime_error(std::string(what) + ": " + cstone_b200_status_string(st));
}

//! Box<T> has the member layout of cstone_box_f32 / cstone_box_f64 (limits[6], lengths[3], inverse lengths[3], boundaries[3])
static const cstone_box_f32* cbox(const Box<float>& b) { return reinterpret_cast<const cstone_box_f32*>(&b); }
static const cstone_box_f64* cbox(const Box<double>& b) { return reinterpret_cast<const cstone_box_f64*>(&b); }
static_assert(sizeof(Box<float>) == sizeof(cstone_box_f32) && sizeof(Box<double>) == sizeof(cstone_box_f64),
              "cstone::Box<T> and the C ABI box differ");

//! grow-only device scratch owned by the shim (the reference's .cu files allocate thrust temporaries in the same places)
class Scratch
{
public:
    void* get(std::size_t bytes)
    {
        if (bytes > cap_)
        {
            if (p_) cstone_b200_device_free(p_);
            cap_ = bytes + bytes / 8 + 256;
            check(cstone_b200_device_malloc(&p_, cap_), "shim scratch");
        }
        return p_;
    }
    std::size_t capacity() const { return cap_; }

private:
    void* p_         = nullptr;
    std::size_t cap_ = 0;
};
static Scratch& scratch(int which)
{
    static thread_local Scratch s[4];
    return s[which];
}

template<class K>
inline constexpr bool is64 = sizeof(K) == 8;

//! `converged` of the last computeNodeOpsGpu, returned by the rebalanceTreeGpu that follows it
static thread_local int32_t lastConverged = 0;

} // namespace b200

// ------------------------------------------------------------------------------------------------------------------
// sfc/sfc_gpu.h:38-39  (instantiations: sfc/sfc_gpu.cu:59-75)
// ------------------------------------------------------------------------------------------------------------------
template<class KeyType, class T>
void computeSfcKeysGpu(const T* x, const T* y, const T* z, KeyType* keys, size_t numKeys, const Box<T>& box)
{
    using I            = typename KeyType::ValueType;
    constexpr int kind = IsHilbert<KeyType>{} ? CSTONE_SFC_HILBERT : CSTONE_SFC_MORTON;
    int st;
    if constexpr (sizeof(I) == 8 && sizeof(T) == 8)
        st = cstone_b200_compute_sfc_keys_k64_f64(kind, x, y, z, reinterpret_cast<uint64_t*>(keys), numKeys, b200::cbox(box), nullptr);
    else if constexpr (sizeof(I) == 8)
        st = cstone_b200_compute_sfc_keys_k64_f32(kind, x, y, z, reinterpret_cast<uint64_t*>(keys), numKeys, b200::cbox(box), nullptr);
    else if constexpr (sizeof(T) == 8)
        st = cstone_b200_compute_sfc_keys_k32_f64(kind, x, y, z, reinterpret_cast<uint32_t*>(keys), numKeys, b200::cbox(box), nullptr);
    else
        st = cstone_b200_compute_sfc_keys_k32_f32(kind, x, y, z, reinterpret_cast<uint32_t*>(keys), numKeys, b200::cbox(box), nullptr);
    b200::check(st, "computeSfcKeysGpu");
}

template void computeSfcKeysGpu(const float*, const float*, const float*, MortonKey<unsigned>*, size_t, const Box<float>&);
template void computeSfcKeysGpu(const double*, const double*, const double*, MortonKey<unsigned>*, size_t, const Box<double>&);
template void computeSfcKeysGpu(const float*, const float*, const float*, MortonKey<uint64_t>*, size_t, const Box<float>&);
template void computeSfcKeysGpu(const double*, const double*, const double*, MortonKey<uint64_t>*, size_t, const Box<double>&);
template void computeSfcKeysGpu(const float*, const float*, const float*, HilbertKey<unsigned>*, size_t, const Box<float>&);
template void computeSfcKeysGpu(const double*, const double*, const double*, HilbertKey<unsigned>*, size_t, const Box<double>&);
template void computeSfcKeysGpu(const float*, const float*, const float*, HilbertKey<uint64_t>*, size_t, const Box<float>&);
template void computeSfcKeysGpu(const double*, const double*, const double*, HilbertKey<uint64_t>*, size_t, const Box<double>&);

// ------------------------------------------------------------------------------------------------------------------
// primitives/primitives_gpu.h:39-125  (instantiations: primitives/primitives_gpu.cu)
// ------------------------------------------------------------------------------------------------------------------
template<class T>
void fillGpu(T* first, T* last, T value)
{
    b200::check(cstone_b200_fill(first, size_t(last - first), &value, int(sizeof(T)), nullptr), "fillGpu");
}
template void fillGpu(double*, double*, double);       // primitives_gpu.cu:57-62
template void fillGpu(float*, float*, float);
template void fillGpu(int*, int*, int);
template void fillGpu(char*, char*, char);
template void fillGpu(unsigned*, unsigned*, unsigned);
template void fillGpu(uint64_t*, uint64_t*, uint64_t);

template<class T>
void scaleGpu(T* first, T* last, T value)
{
    if constexpr (sizeof(T) == 8) b200::check(cstone_b200_scale_f64(first, size_t(last - first), value, nullptr), "scaleGpu");
    else b200::check(cstone_b200_scale_f32(first, size_t(last - first), value, nullptr), "scaleGpu");
}
template void scaleGpu(double*, double*, double);      // :83-84
template void scaleGpu(float*, float*, float);

template<class T>
void incrementGpu(const T* first, const T* last, T* d_first, T value)
{
    if constexpr (sizeof(T) == 8)
        b200::check(cstone_b200_increment_u64(reinterpret_cast<const uint64_t*>(first), reinterpret_cast<uint64_t*>(d_first),
                                              size_t(last - first), uint64_t(value), nullptr), "incrementGpu");
    else
        b200::check(cstone_b200_increment_u32(first, d_first, size_t(last - first), value, nullptr), "incrementGpu");
}
template void incrementGpu(const unsigned*, const unsigned*, unsigned*, unsigned);   // :105-107
template void incrementGpu(const uint64_t*, const uint64_t*, uint64_t*, uint64_t);

//! index maps are non-negative: an `int` map has the bit patterns of a LocalIndex map
template<class T, class IndexType>
void gatherGpu(const IndexType* ordering, size_t numElements, const T* src, T* buffer)
{
    static_assert(sizeof(IndexType) == 4);
    b200::check(cstone_b200_gather(reinterpret_cast<const uint32_t*>(ordering), numElements, src, buffer, int(sizeof(T)), nullptr),
                "gatherGpu");
}
template void gatherGpu(const int*, size_t, const int*, int*);                       // :126-134
template void gatherGpu(const int*, size_t, const uint32_t*, uint32_t*);
template void gatherGpu(const int*, size_t, const uint64_t*, uint64_t*);
template void gatherGpu(const int*, size_t, const util::array<float, 4>*, util::array<float, 4>*);
template void gatherGpu(const int*, size_t, const util::array<float, 8>*, util::array<float, 8>*);
template void gatherGpu(const int*, size_t, const util::array<double, 4>*, util::array<double, 4>*);
template void gatherGpu(const unsigned*, size_t, const double*, double*);            // :136-148
template void gatherGpu(const unsigned*, size_t, const float*, float*);
template void gatherGpu(const unsigned*, size_t, const int*, int*);
template void gatherGpu(const unsigned*, size_t, const long*, long*);
template void gatherGpu(const unsigned*, size_t, const unsigned*, unsigned*);
template void gatherGpu(const unsigned*, size_t, const unsigned long*, unsigned long*);
template void gatherGpu(const unsigned*, size_t, const unsigned long long*, unsigned long long*);
template void gatherGpu(const unsigned*, size_t, const util::array<float, 1>*, util::array<float, 1>*);
template void gatherGpu(const unsigned*, size_t, const util::array<float, 2>*, util::array<float, 2>*);
template void gatherGpu(const unsigned*, size_t, const util::array<float, 3>*, util::array<float, 3>*);
template void gatherGpu(const unsigned*, size_t, const util::array<float, 4>*, util::array<float, 4>*);

template<class T, class IndexType>
void scatterGpu(const IndexType* ordering, size_t numElements, const T* src, T* buffer)
{
    static_assert(sizeof(IndexType) == 4);
    b200::check(cstone_b200_scatter(reinterpret_cast<const uint32_t*>(ordering), numElements, src, buffer, int(sizeof(T)), nullptr),
                "scatterGpu");
}
template void scatterGpu(const int*, size_t, const int*, int*);                      // :167-175
template void scatterGpu(const int*, size_t, const uint32_t*, uint32_t*);
template void scatterGpu(const int*, size_t, const uint64_t*, uint64_t*);
template void scatterGpu(const int*, size_t, const util::array<float, 4>*, util::array<float, 4>*);
template void scatterGpu(const int*, size_t, const util::array<float, 8>*, util::array<float, 8>*);
template void scatterGpu(const int*, size_t, const util::array<double, 4>*, util::array<double, 4>*);

template<class T>
std::tuple<T, T> MinMaxGpu<T>::operator()(const T* first, const T* last)
{
    T mm[6];
    void* temp = b200::scratch(0).get(cstone_b200_minmax_temp_bytes());
    if constexpr (sizeof(T) == 8)
        b200::check(cstone_b200_minmax3_f64(first, nullptr, nullptr, size_t(last - first), mm, temp, nullptr), "MinMaxGpu");
    else
        b200::check(cstone_b200_minmax3_f32(first, nullptr, nullptr, size_t(last - first), mm, temp, nullptr), "MinMaxGpu");
    return std::make_tuple(mm[0], mm[1]);
}
template class MinMaxGpu<double>;                                                     // :189-190
template class MinMaxGpu<float>;

template<class T>
T maxNormSquareGpu(const T* x, const T* y, const T* z, size_t numElements)
{
    T out;
    void* temp = b200::scratch(0).get(cstone_b200_reduce_temp_bytes());
    if constexpr (sizeof(T) == 8)
        b200::check(cstone_b200_max_norm_square_f64(x, y, z, numElements, &out, temp, nullptr), "maxNormSquareGpu");
    else
        b200::check(cstone_b200_max_norm_square_f32(x, y, z, numElements, &out, temp, nullptr), "maxNormSquareGpu");
    return out;
}
template float maxNormSquareGpu(const float*, const float*, const float*, size_t);    // :214-215
template double maxNormSquareGpu(const double*, const double*, const double*, size_t);

template<class T>
size_t lowerBoundGpu(const T* first, const T* last, T value)
{
    uint64_t out;
    void* temp     = b200::scratch(0).get(cstone_b200_reduce_temp_bytes());
    const size_t n = size_t(last - first);
    int st;
    if constexpr (std::is_same_v<T, unsigned>) st = cstone_b200_lower_bound_u32(first, n, value, &out, temp, nullptr);
    else if constexpr (std::is_same_v<T, uint64_t>) st = cstone_b200_lower_bound_u64(first, n, value, &out, temp, nullptr);
    else if constexpr (std::is_same_v<T, int>) st = cstone_b200_lower_bound_i32(first, n, value, &out, temp, nullptr);
    else if constexpr (std::is_same_v<T, int64_t>) st = cstone_b200_lower_bound_i64(first, n, value, &out, temp, nullptr);
    else st = cstone_b200_lower_bound_f32(first, n, value, &out, temp, nullptr);
    b200::check(st, "lowerBoundGpu");
    return size_t(out);
}
template size_t lowerBoundGpu(const unsigned*, const unsigned*, unsigned);           // :223-227
template size_t lowerBoundGpu(const uint64_t*, const uint64_t*, uint64_t);
template size_t lowerBoundGpu(const int*, const int*, int);
template size_t lowerBoundGpu(const int64_t*, const int64_t*, int64_t);
template size_t lowerBoundGpu(const float*, const float*, float);

template<class T, class IndexType>
void lowerBoundGpu(const T* first, const T* last, const T* valueFirst, const T* valueLast, IndexType* result)
{
    const size_t n = size_t(last - first), nv = size_t(valueLast - valueFirst);
    int st;
    if constexpr (sizeof(T) == 8 && sizeof(IndexType) == 8)
        st = cstone_b200_lower_bound_many_k64_u64(first, n, valueFirst, nv, result, nullptr);
    else if constexpr (sizeof(T) == 8)
        st = cstone_b200_lower_bound_many_k64_u32(first, n, valueFirst, nv, result, nullptr);
    else if constexpr (sizeof(IndexType) == 8)
        st = cstone_b200_lower_bound_many_k32_u64(first, n, valueFirst, nv, result, nullptr);
    else
        st = cstone_b200_lower_bound_many_k32_u32(first, n, valueFirst, nv, result, nullptr);
    b200::check(st, "lowerBoundGpu");
}
template void lowerBoundGpu(const unsigned*, const unsigned*, const unsigned*, const unsigned*, unsigned*);   // :235-238
template void lowerBoundGpu(const uint64_t*, const uint64_t*, const uint64_t*, const uint64_t*, unsigned*);
template void lowerBoundGpu(const unsigned*, const unsigned*, const unsigned*, const unsigned*, uint64_t*);
template void lowerBoundGpu(const uint64_t*, const uint64_t*, const uint64_t*, const uint64_t*, uint64_t*);

/* segmentMax [primitives_gpu.h:73-85]: the C ABI wants the number of input elements as a work-size hint (it is
 * segments[numSegments], which lives on the device); a per-leaf average of 64 bounds it well enough for the grid. */
template<class Tin, class Tout, class IndexType>
void segmentMax(const Tin* input, const IndexType* segments, size_t numSegments, Tout* output)
{
    const size_t hint = 64 * numSegments;
    int st;
    if constexpr (sizeof(Tin) == 4) st = cstone_b200_segment_max_f32_f32(input, segments, int(sizeof(IndexType)), numSegments, hint, output, nullptr);
    else if constexpr (sizeof(Tout) == 4) st = cstone_b200_segment_max_f64_f32(input, segments, int(sizeof(IndexType)), numSegments, hint, output, nullptr);
    else st = cstone_b200_segment_max_f64_f64(input, segments, int(sizeof(IndexType)), numSegments, hint, output, nullptr);
    b200::check(st, "segmentMax");
}
template void segmentMax(const float*, const unsigned*, size_t, float*);             // :270-275
template void segmentMax(const double*, const unsigned*, size_t, float*);
template void segmentMax(const double*, const unsigned*, size_t, double*);
template void segmentMax(const float*, const uint64_t*, size_t, float*);
template void segmentMax(const double*, const uint64_t*, size_t, float*);
template void segmentMax(const double*, const uint64_t*, size_t, double*);

template<class Tin, class Tout>
Tout reduceGpu(const Tin* input, size_t numElements, Tout init)
{
    uint64_t out;
    void* temp = b200::scratch(0).get(cstone_b200_reduce_temp_bytes());
    b200::check(cstone_b200_reduce_sum_u32(input, numElements, uint64_t(init), &out, temp, nullptr), "reduceGpu");
    return Tout(out);
}
template size_t reduceGpu(const unsigned*, size_t, size_t);                          // :283

template<class IndexType>
void sequenceGpu(IndexType* input, size_t numElements, IndexType init)
{
    static_assert(sizeof(IndexType) == 4);
    b200::check(cstone_b200_sequence_u32(reinterpret_cast<uint32_t*>(input), numElements, uint32_t(init), nullptr), "sequenceGpu");
}
template void sequenceGpu(int*, size_t, int);                                        // :291-292
template void sequenceGpu(unsigned*, size_t, unsigned);

template<class KeyType>
void sortGpu(KeyType* first, KeyType* last, KeyType* keyBuf)
{
    const size_t n = size_t(last - first);
    if constexpr (sizeof(KeyType) == 8)
    {
        const uint64_t tb = cstone_b200_sort_keys_temp_bytes_k64(n);
        b200::check(cstone_b200_sort_keys_k64(first, keyBuf, n, b200::scratch(1).get(tb), tb, nullptr), "sortGpu");
    }
    else
    {
        const uint64_t tb = cstone_b200_sort_keys_temp_bytes_k32(n);
        b200::check(cstone_b200_sort_keys_k32(first, keyBuf, n, b200::scratch(1).get(tb), tb, nullptr), "sortGpu");
    }
}
template void sortGpu(uint32_t*, uint32_t*, uint32_t*);                              // :322-323
template void sortGpu(uint64_t*, uint64_t*, uint64_t*);

template<class KeyType, class ValueType>
uint64_t sortByKeyTempStorage(uint64_t numElements)
{
    return sizeof(KeyType) == 8 ? cstone_b200_sort_by_key_temp_bytes_k64(numElements)
                                : cstone_b200_sort_by_key_temp_bytes_k32(numElements);
}

//! a too small temp buffer throws std::runtime_error, as primitives_gpu.cu:353
template<class KeyType, class ValueType>
void sortByKeyGpu(KeyType* first, KeyType* last, ValueType* values, KeyType* keyBuf, ValueType* valueBuf, void* temp,
                  uint64_t tempBytes)
{
    static_assert(sizeof(ValueType) == 4);
    const size_t n = size_t(last - first);
    int st;
    if constexpr (sizeof(KeyType) == 8)
        st = cstone_b200_sort_by_key_k64(first, reinterpret_cast<uint32_t*>(values), keyBuf, reinterpret_cast<uint32_t*>(valueBuf), n,
                                         temp, tempBytes, 0, nullptr);
    else
        st = cstone_b200_sort_by_key_k32(first, reinterpret_cast<uint32_t*>(values), keyBuf, reinterpret_cast<uint32_t*>(valueBuf), n,
                                         temp, tempBytes, 0, nullptr);
    b200::check(st, "sortByKeyGpu");
}
#define SORT_BY_KEY_GPU_DB(KeyType, ValueType)                                                                         \
    template void sortByKeyGpu(KeyType*, KeyType*, ValueType*, KeyType*, ValueType*, void*, uint64_t);                 \
    template uint64_t sortByKeyTempStorage<KeyType, ValueType>(uint64_t)
SORT_BY_KEY_GPU_DB(unsigned, unsigned);                                              // :375-378
SORT_BY_KEY_GPU_DB(unsigned, int);
SORT_BY_KEY_GPU_DB(uint64_t, unsigned);
SORT_BY_KEY_GPU_DB(uint64_t, int);
#undef SORT_BY_KEY_GPU_DB

//! the allocating form (thrust::sort_by_key in the reference): alternate buffers come from the shim's scratch
template<class KeyType, class ValueType>
void sortByKeyGpu(KeyType* first, KeyType* last, ValueType* values)
{
    const size_t n    = size_t(last - first);
    const uint64_t tb = sortByKeyTempStorage<KeyType, ValueType>(n);
    auto* keyBuf      = static_cast<KeyType*>(b200::scratch(1).get(n * sizeof(KeyType)));
    auto* valueBuf    = static_cast<ValueType*>(b200::scratch(2).get(n * sizeof(ValueType)));
    sortByKeyGpu(first, last, values, keyBuf, valueBuf, b200::scratch(3).get(tb), tb);
}
template void sortByKeyGpu(unsigned*, unsigned*, unsigned*);                         // :388-391
template void sortByKeyGpu(unsigned*, unsigned*, int*);
template void sortByKeyGpu(uint64_t*, uint64_t*, unsigned*);
template void sortByKeyGpu(uint64_t*, uint64_t*, int*);

//! 32-bit wrapping sums: int and unsigned share their bit patterns
template<class IndexType, class SumType>
void exclusiveScanGpu(const IndexType* first, const IndexType* last, SumType* output, SumType init)
{
    static_assert(sizeof(IndexType) == 4 && sizeof(SumType) == 4);
    const size_t n = size_t(last - first);
    void* temp     = b200::scratch(0).get(cstone_b200_scan_temp_bytes(n));
    b200::check(cstone_b200_exclusive_scan_u32(reinterpret_cast<const uint32_t*>(first), reinterpret_cast<uint32_t*>(output), n,
                                               uint32_t(init), temp, nullptr), "exclusiveScanGpu");
}
template void exclusiveScanGpu(const int*, const int*, int*, int);                   // :400-403
template void exclusiveScanGpu(const int*, const int*, unsigned*, unsigned);
template void exclusiveScanGpu(const unsigned*, const unsigned*, unsigned*, unsigned);

template<class IndexType, class SumType>
void inclusiveScanGpu(const IndexType* first, const IndexType* last, SumType* output)
{
    static_assert(sizeof(IndexType) == 4 && sizeof(SumType) == 4);
    const size_t n = size_t(last - first);
    void* temp     = b200::scratch(0).get(cstone_b200_scan_temp_bytes(n));
    b200::check(cstone_b200_inclusive_scan_u32(reinterpret_cast<const uint32_t*>(first), reinterpret_cast<uint32_t*>(output), n, temp,
                                               nullptr), "inclusiveScanGpu");
}
template void inclusiveScanGpu(const int*, const int*, int*);                        // :434-437
template void inclusiveScanGpu(const int*, const int*, unsigned*);
template void inclusiveScanGpu(const unsigned*, const unsigned*, unsigned*);

template<class ValueType>
size_t countGpu(const ValueType* first, const ValueType* last, ValueType v)
{
    uint64_t out;
    void* temp     = b200::scratch(0).get(cstone_b200_reduce_temp_bytes());
    const size_t n = size_t(last - first);
    if constexpr (sizeof(ValueType) == 8)
        b200::check(cstone_b200_count_u64(reinterpret_cast<const uint64_t*>(first), n, uint64_t(v), &out, temp, nullptr), "countGpu");
    else
        b200::check(cstone_b200_count_u32(reinterpret_cast<const uint32_t*>(first), n, uint32_t(v), &out, temp, nullptr), "countGpu");
    return size_t(out);
}
template size_t countGpu(const int* first, const int* last, int v);                  // :446-448
template size_t countGpu(const unsigned* first, const unsigned* last, unsigned v);
template size_t countGpu(const uint64_t* first, const uint64_t* last, uint64_t v);

// ------------------------------------------------------------------------------------------------------------------
// tree/csarray_gpu.h:55-89  (instantiations: tree/csarray_gpu.cu:132-135, 208-209, 228-229, 245-246, 264-265)
// ------------------------------------------------------------------------------------------------------------------
template<class KeyType>
void computeNodeCountsGpu(const KeyType* tree, unsigned* counts, TreeNodeIndex numNodes, const KeyType* firstKey,
                          const KeyType* lastKey, unsigned maxCount, bool useCountsAsGuess)
{
    if constexpr (b200::is64<KeyType>)
        b200::check(cstone_b200_compute_node_counts_k64(tree, counts, numNodes, firstKey, lastKey, maxCount, useCountsAsGuess, nullptr),
                    "computeNodeCountsGpu");
    else
        b200::check(cstone_b200_compute_node_counts_k32(tree, counts, numNodes, firstKey, lastKey, maxCount, useCountsAsGuess, nullptr),
                    "computeNodeCountsGpu");
}
template void computeNodeCountsGpu(const unsigned*, unsigned*, TreeNodeIndex, const unsigned*, const unsigned*, unsigned, bool);
template void computeNodeCountsGpu(const uint64_t*, unsigned*, TreeNodeIndex, const uint64_t*, const uint64_t*, unsigned, bool);

//! returns the node count of the rebalanced tree; nodeOps holds numNodes + 1 scan entries [tree/csarray_gpu.h:64-77]
template<class KeyType>
TreeNodeIndex computeNodeOpsGpu(const KeyType* tree, TreeNodeIndex numNodes, const unsigned* counts, unsigned bucketSize,
                                TreeNodeIndex* nodeOps)
{
    int32_t newNumNodes = 0;
    int32_t& converged  = b200::lastConverged;
    void* temp = b200::scratch(0).get(cstone_b200_scan_temp_bytes(size_t(numNodes) + 1));
    if constexpr (b200::is64<KeyType>)
        b200::check(cstone_b200_compute_node_ops_k64(tree, numNodes, counts, bucketSize, nodeOps, &newNumNodes, &converged, temp, nullptr),
                    "computeNodeOpsGpu");
    else
        b200::check(cstone_b200_compute_node_ops_k32(tree, numNodes, counts, bucketSize, nodeOps, &newNumNodes, &converged, temp, nullptr),
                    "computeNodeOpsGpu");
    return newNumNodes;
}
template TreeNodeIndex computeNodeOpsGpu(const unsigned*, TreeNodeIndex, const unsigned*, unsigned, TreeNodeIndex*);
template TreeNodeIndex computeNodeOpsGpu(const uint64_t*, TreeNodeIndex, const unsigned*, unsigned, TreeNodeIndex*);

/* returns `converged` = every decision of the preceding computeNodeOpsGpu was "keep".  The reference carries that
 * flag from one call to the other in a __device__ counter (tree/csarray_gpu.cu:40,160-169,224-227); here it comes back
 * from the decision kernel's scan and waits in b200::lastConverged. */
template<class KeyType>
bool rebalanceTreeGpu(const KeyType* tree, TreeNodeIndex numNodes, TreeNodeIndex newNumNodes, const TreeNodeIndex* nodeOps,
                      KeyType* newTree)
{
    if constexpr (b200::is64<KeyType>)
        b200::check(cstone_b200_rebalance_tree_k64(tree, numNodes, newNumNodes, nodeOps, newTree, nullptr), "rebalanceTreeGpu");
    else
        b200::check(cstone_b200_rebalance_tree_k32(tree, numNodes, newNumNodes, nodeOps, newTree, nullptr), "rebalanceTreeGpu");
    return b200::lastConverged != 0;
}
template bool rebalanceTreeGpu(const unsigned*, TreeNodeIndex, TreeNodeIndex, const TreeNodeIndex*, unsigned*);
template bool rebalanceTreeGpu(const uint64_t*, TreeNodeIndex, TreeNodeIndex, const TreeNodeIndex*, uint64_t*);

template<class KeyType>
void countSfcGapsGpu(const KeyType* tree, TreeNodeIndex numNodes, TreeNodeIndex* nodeOps)
{
    if constexpr (b200::is64<KeyType>) b200::check(cstone_b200_count_sfc_gaps_k64(tree, numNodes, nodeOps, nullptr), "countSfcGapsGpu");
    else b200::check(cstone_b200_count_sfc_gaps_k32(tree, numNodes, nodeOps, nullptr), "countSfcGapsGpu");
}
template void countSfcGapsGpu(const uint32_t*, TreeNodeIndex, TreeNodeIndex*);
template void countSfcGapsGpu(const uint64_t*, TreeNodeIndex, TreeNodeIndex*);

template<class KeyType>
void fillSfcGapsGpu(const KeyType* tree, TreeNodeIndex numNodes, const TreeNodeIndex* nodeOps, KeyType* newTree)
{
    if constexpr (b200::is64<KeyType>) b200::check(cstone_b200_fill_sfc_gaps_k64(tree, numNodes, nodeOps, newTree, nullptr), "fillSfcGapsGpu");
    else b200::check(cstone_b200_fill_sfc_gaps_k32(tree, numNodes, nodeOps, newTree, nullptr), "fillSfcGapsGpu");
}
template void fillSfcGapsGpu(const uint32_t*, TreeNodeIndex, const TreeNodeIndex*, uint32_t*);
template void fillSfcGapsGpu(const uint64_t*, TreeNodeIndex, const TreeNodeIndex*, uint64_t*);

// ------------------------------------------------------------------------------------------------------------------
// tree/octree_gpu.h:49-53  (instantiations: tree/octree_gpu.cu:175-176)
// ------------------------------------------------------------------------------------------------------------------
template<class KeyType>
void buildOctreeGpu(const KeyType* cstoneTree, OctreeView<KeyType> d)
{
    if constexpr (b200::is64<KeyType>)
    {
        const uint64_t tb = cstone_b200_build_octree_temp_bytes_k64(d.numLeafNodes);
        b200::check(cstone_b200_build_octree_k64(cstoneTree, d.numLeafNodes, d.prefixes, d.childOffsets, d.parents, d.levelRange,
                                                 d.internalToLeaf, d.leafToInternal, b200::scratch(1).get(tb), tb, nullptr),
                    "buildOctreeGpu");
    }
    else
    {
        const uint64_t tb = cstone_b200_build_octree_temp_bytes_k32(d.numLeafNodes);
        b200::check(cstone_b200_build_octree_k32(cstoneTree, d.numLeafNodes, d.prefixes, d.childOffsets, d.parents, d.levelRange,
                                                 d.internalToLeaf, d.leafToInternal, b200::scratch(1).get(tb), tb, nullptr),
                    "buildOctreeGpu");
    }
}
template void buildOctreeGpu(const uint32_t*, OctreeView<uint32_t>);
template void buildOctreeGpu(const uint64_t*, OctreeView<uint64_t>);

//! lvlRange is a HOST array in the reference's call sites (focus/octree_focus_mpi.hpp: octree().levelRange of the CPU tree)
void upsweepSumGpu(int numLvl, const TreeNodeIndex* lvlRange, const TreeNodeIndex* childOffsets, LocalIndex* counts)
{
    b200::check(cstone_b200_upsweep_sum(numLvl, lvlRange, childOffsets, counts, nullptr), "upsweepSumGpu");
}

// ------------------------------------------------------------------------------------------------------------------
// traversal/collisions_gpu.h:58-77  (instantiations: traversal/collisions_gpu.cu:112-115, 171-174)
// plain integer key types are walked on the Hilbert curve there (SfcKind<KeyType> = HilbertKey, sfc/sfc.hpp)
// ------------------------------------------------------------------------------------------------------------------
template<class KeyType, class RadiusType, class T>
void findHalosGpu(const KeyType* prefixes, const TreeNodeIndex* childOffsets, const TreeNodeIndex* internalToLeaf,
                  const KeyType* leaves, const RadiusType* interactionRadii, const Box<T>& box, TreeNodeIndex firstNode,
                  TreeNodeIndex lastNode, int* collisionFlags)
{
    static_assert(std::is_same_v<RadiusType, float>);
    int st;
    if constexpr (b200::is64<KeyType> && sizeof(T) == 8)
        st = cstone_b200_find_halos_k64_f64(CSTONE_SFC_HILBERT, prefixes, childOffsets, internalToLeaf, leaves, interactionRadii,
                                            b200::cbox(box), firstNode, lastNode, collisionFlags, nullptr);
    else if constexpr (b200::is64<KeyType>)
        st = cstone_b200_find_halos_k64_f32(CSTONE_SFC_HILBERT, prefixes, childOffsets, internalToLeaf, leaves, interactionRadii,
                                            b200::cbox(box), firstNode, lastNode, collisionFlags, nullptr);
    else if constexpr (sizeof(T) == 8)
        st = cstone_b200_find_halos_k32_f64(CSTONE_SFC_HILBERT, prefixes, childOffsets, internalToLeaf, leaves, interactionRadii,
                                            b200::cbox(box), firstNode, lastNode, collisionFlags, nullptr);
    else
        st = cstone_b200_find_halos_k32_f32(CSTONE_SFC_HILBERT, prefixes, childOffsets, internalToLeaf, leaves, interactionRadii,
                                            b200::cbox(box), firstNode, lastNode, collisionFlags, nullptr);
    b200::check(st, "findHalosGpu");
}
#define FIND_HALOS_GPU(KeyType, RadiusType, T)                                                                         \
    template void findHalosGpu(const KeyType* prefixes, const TreeNodeIndex* childOffsets,                             \
                               const TreeNodeIndex* internalToLeaf, const KeyType* leaves,                             \
                               const RadiusType* interactionRadii, const Box<T>& box, TreeNodeIndex firstNode,         \
                               TreeNodeIndex lastNode, int* collisionFlags)
FIND_HALOS_GPU(uint32_t, float, float);
FIND_HALOS_GPU(uint32_t, float, double);
FIND_HALOS_GPU(uint64_t, float, float);
FIND_HALOS_GPU(uint64_t, float, double);
#undef FIND_HALOS_GPU

template<class T, class KeyType>
void markMacsGpu(const KeyType* prefixes, const TreeNodeIndex* childOffsets, const Vec4<T>* centers, const Box<T>& box,
                 const KeyType* focusNodes, TreeNodeIndex numFocusNodes, bool limitSource, char* markings)
{
    int st;
    if constexpr (b200::is64<KeyType> && sizeof(T) == 8)
        st = cstone_b200_mark_macs_k64_f64(CSTONE_SFC_HILBERT, prefixes, childOffsets, reinterpret_cast<const double*>(centers),
                                           b200::cbox(box), focusNodes, numFocusNodes, limitSource, markings, nullptr);
    else if constexpr (b200::is64<KeyType>)
        st = cstone_b200_mark_macs_k64_f32(CSTONE_SFC_HILBERT, prefixes, childOffsets, reinterpret_cast<const float*>(centers),
                                           b200::cbox(box), focusNodes, numFocusNodes, limitSource, markings, nullptr);
    else if constexpr (sizeof(T) == 8)
        st = cstone_b200_mark_macs_k32_f64(CSTONE_SFC_HILBERT, prefixes, childOffsets, reinterpret_cast<const double*>(centers),
                                           b200::cbox(box), focusNodes, numFocusNodes, limitSource, markings, nullptr);
    else
        st = cstone_b200_mark_macs_k32_f32(CSTONE_SFC_HILBERT, prefixes, childOffsets, reinterpret_cast<const float*>(centers),
                                           b200::cbox(box), focusNodes, numFocusNodes, limitSource, markings, nullptr);
    b200::check(st, "markMacsGpu");
}
#define MARK_MACS_GPU(KeyType, T)                                                                                      \
    template void markMacsGpu(const KeyType* prefixes, const TreeNodeIndex* childOffsets, const Vec4<T>* centers,      \
                              const Box<T>& box, const KeyType* focusNodes, TreeNodeIndex numFocusNodes,               \
                              bool limitSource, char* markings)
MARK_MACS_GPU(uint64_t, double);
MARK_MACS_GPU(uint64_t, float);
MARK_MACS_GPU(unsigned, double);
MARK_MACS_GPU(unsigned, float);
#undef MARK_MACS_GPU

// ------------------------------------------------------------------------------------------------------------------
// halos/gather_halos_gpu.h:21-27  (instantiations: halos/gather_halos_gpu.cu:56-75)
// ------------------------------------------------------------------------------------------------------------------
template<class T, class IndexType>
void gatherRanges(const IndexType* rangeScan, const IndexType* rangeOffsets, int numRanges, const T* src, T* buffer,
                  size_t bufferSize)
{
    b200::check(cstone_b200_gather_ranges(rangeScan, rangeOffsets, int(sizeof(IndexType)), numRanges, src, buffer, bufferSize,
                                          int(sizeof(T)), nullptr),
                "gatherRanges");
}
template void gatherRanges(const unsigned*, const unsigned*, int, const int*, int*, size_t);
template void gatherRanges(const uint64_t*, const uint64_t*, int, const int*, int*, size_t);
template void gatherRanges(const unsigned*, const unsigned*, int, const util::array<float, 1>*, util::array<float, 1>*, size_t);
template void gatherRanges(const unsigned*, const unsigned*, int, const util::array<float, 2>*, util::array<float, 2>*, size_t);
template void gatherRanges(const unsigned*, const unsigned*, int, const util::array<float, 3>*, util::array<float, 3>*, size_t);
template void gatherRanges(const unsigned*, const unsigned*, int, const util::array<float, 4>*, util::array<float, 4>*, size_t);
template void gatherRanges(const uint64_t*, const uint64_t*, int, const util::array<float, 1>*, util::array<float, 1>*, size_t);
template void gatherRanges(const uint64_t*, const uint64_t*, int, const util::array<float, 2>*, util::array<float, 2>*, size_t);
template void gatherRanges(const uint64_t*, const uint64_t*, int, const util::array<float, 3>*, util::array<float, 3>*, size_t);
template void gatherRanges(const uint64_t*, const uint64_t*, int, const util::array<float, 4>*, util::array<float, 4>*, size_t);

/* Instantiations of the reference that this shim does NOT provide (none of them is reached from Domain::sync,
 * exchangeHalos or the tree / halo / neighbour functions of the hot path):
 *   gatherGpu / scatterGpu for util::array<float,12>, util::array<double,8|12> (48 - 96 byte multipole rows; the C ABI
 *       moves 4 - 32 byte elements), gatherGpu for 1-byte elements (uint8_t, char);
 *   sequenceGpu<uint64_t>, sortGpu<float>, sortByKeyGpu<uint64_t,uint64_t>, sortByKeyGpu<float,unsigned>;
 *   exclusiveScanGpu with uint64_t sums.
 * A program that needs one of them gets an undefined-symbol error at link time, not a wrong answer. */

} // namespace cstone
