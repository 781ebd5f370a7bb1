/*
 * neighbors.cu — a15: radius neighbour search with the reference's CPU semantics and output layout.
 *
 * Mirrors findNeighbors [ref: findneighbors.hpp:95-188]: for particle i all j != i in the visited leaves with
 * |x_j - x_i|^2 < 4 h_i^2 (strict), PBC fold with rint when the 2h cube of i leaves the box, at most ngmax
 * indices stored at neighbors[(i-first)*ngmax + k] in DEPTH-FIRST visiting order, full count returned.
 * Arithmetic (no FMA, same association) follows distanceSq [ref: findneighbors.hpp:51-78], minDistance /
 * insideBox [ref: traversal/boxoverlap.hpp:185-217], applyPbc [ref: sfc/box.hpp:194-206] and norm2, which is a
 * right fold x*x + (y*y + z*z) [ref: util/array.hpp:253-256,333-336].
 *
 * B200 design: one warp walks the tree for 32 consecutive (SFC-adjacent) targets.  Every stack entry carries
 * the mask of lanes whose OWN depth-first walk reaches that node, so each lane sees exactly the node sequence
 * (and therefore the neighbour order) of the reference's per-particle walk, while node data and leaf
 * particles are fetched once per warp (uniform loads) instead of once per particle.
 */
#include "common.cuh"

namespace csb
{

constexpr int kNbWarpsPerBlock = 8;
constexpr int kNbStack         = 160; // >= 7 * 21 + 1 entries can be pending at any time

template<class T>
__device__ inline T absT(T v)
{
    return fabs(v);
}
__device__ inline float absT(float v) { return fabsf(v); }

template<class T>
struct NbParams
{
    const T* x;
    const T* y;
    const T* z;
    const T* h;
    const int* childOffsets;
    const int* internalToLeaf;
    const uint32_t* layout;
    const T* centers;
    const T* sizes;
    uint32_t* neighbors;
    uint32_t* neighborsCount;
    uint32_t firstId, lastId, ngmax;
    float ext;
    DBox<T> box;
};

template<class T, bool AnyPbc>
__global__ void __launch_bounds__(kNbWarpsPerBlock * 32) findNeighborsKernel(NbParams<T> p)
{
    __shared__ int sNode[kNbWarpsPerBlock][kNbStack];
    __shared__ unsigned sMask[kNbWarpsPerBlock][kNbStack];

    const unsigned lane   = laneId();
    const unsigned warp   = threadIdx.x >> 5;
    const unsigned full   = 0xffffffffu;
    int* stackNode        = sNode[warp];
    unsigned* stackMask   = sMask[warp];
    const uint32_t numTargets = p.lastId - p.firstId;
    const uint32_t numGroups  = (numTargets + 31u) / 32u;
    const uint32_t warpsTotal = gridDim.x * kNbWarpsPerBlock;

    for (uint32_t group = blockIdx.x * kNbWarpsPerBlock + warp; group < numGroups; group += warpsTotal)
    {
        const uint32_t w  = group * 32u + lane;
        const bool valid  = w < numTargets;
        const uint32_t i  = p.firstId + (valid ? w : numTargets - 1);
        const T xi = p.x[i], yi = p.y[i], zi = p.z[i], hi = p.h[i];
        const T radiusSq     = T(4.0) * hi * hi;
        const T cellRadiusSq = radiusSq * p.ext * p.ext;
        uint32_t* out        = p.neighbors + size_t(valid ? w : 0) * p.ngmax;
        uint32_t count       = 0;

        bool usePbc = false;
        if (AnyPbc)
        {
            T r2h       = T(2) * hi;
            bool inside = (xi - r2h >= p.box.lo[0]) && (yi - r2h >= p.box.lo[1]) && (zi - r2h >= p.box.lo[2]) &&
                          (xi + r2h <= p.box.hi[0]) && (yi + r2h <= p.box.hi[1]) && (zi + r2h <= p.box.hi[2]);
            usePbc = !inside;
        }

        auto overlaps = [&](int idx) -> bool
        {
            T dx = p.centers[3 * idx] - xi;
            T dy = p.centers[3 * idx + 1] - yi;
            T dz = p.centers[3 * idx + 2] - zi;
            if (AnyPbc && usePbc)
            {
                dx -= T(p.box.pbc[0]) * p.box.len[0] * rint(dx * p.box.ilen[0]);
                dy -= T(p.box.pbc[1]) * p.box.len[1] * rint(dy * p.box.ilen[1]);
                dz -= T(p.box.pbc[2]) * p.box.len[2] * rint(dz * p.box.ilen[2]);
            }
            dx = absT(dx) - p.sizes[3 * idx];
            dy = absT(dy) - p.sizes[3 * idx + 1];
            dz = absT(dz) - p.sizes[3 * idx + 2];
            dx += absT(dx);
            dy += absT(dy);
            dz += absT(dz);
            dx *= T(0.5);
            dy *= T(0.5);
            dz *= T(0.5);
            return dx * dx + (dy * dy + dz * dz) < cellRadiusSq;
        };

        auto searchLeaf = [&](int idx, unsigned mask)
        {
            int leaf       = p.internalToLeaf[idx];
            uint32_t jBeg  = p.layout[leaf];
            uint32_t jEnd  = p.layout[leaf + 1];
            if (!((mask >> lane) & 1u)) return;
            for (uint32_t j = jBeg; j < jEnd; ++j)
            {
                T dx = p.x[j] - xi;
                T dy = p.y[j] - yi;
                T dz = p.z[j] - zi;
                if (AnyPbc && usePbc)
                {
                    dx -= T(p.box.pbc[0]) * p.box.len[0] * rint(dx * p.box.ilen[0]);
                    dy -= T(p.box.pbc[1]) * p.box.len[1] * rint(dy * p.box.ilen[1]);
                    dz -= T(p.box.pbc[2]) * p.box.len[2] * rint(dz * p.box.ilen[2]);
                }
                T d2 = dx * dx + dy * dy + dz * dz;
                if (d2 < radiusSq && j != i)
                {
                    if (count < p.ngmax) out[count] = j;
                    ++count;
                }
            }
        };

        unsigned rootMask = __ballot_sync(full, valid && overlaps(0));
        if (rootMask)
        {
            if (p.childOffsets[0] == 0) { searchLeaf(0, rootMask); }
            else
            {
                int sp = 0;
                if (lane == 0)
                {
                    stackNode[0] = 0;
                    stackMask[0] = 0;
                }
                sp            = 1;
                int node      = 0;
                unsigned mask = rootMask;
                do
                {
                    const int firstChild = p.childOffsets[node];
                    const bool mine      = (mask >> lane) & 1u;
#pragma unroll 1
                    for (int o = 0; o < 8; ++o)
                    {
                        int child       = firstChild + o;
                        unsigned cmask  = __ballot_sync(full, mine && overlaps(child));
                        if (!cmask) continue;
                        if (p.childOffsets[child] == 0) { searchLeaf(child, cmask); }
                        else
                        {
                            if (lane == 0)
                            {
                                stackNode[sp] = child;
                                stackMask[sp] = cmask;
                            }
                            ++sp;
                        }
                    }
                    __syncwarp();
                    --sp;
                    node = stackNode[sp];
                    mask = stackMask[sp];
                    __syncwarp();
                } while (node != 0);
            }
        }
        if (valid) p.neighborsCount[w] = count;
    }
}

template<class T>
int findNeighbors(const T* x, const T* y, const T* z, const T* h, uint32_t firstId, uint32_t lastId,
                  const typename BoxT<T>::type* box, const int* childOffsets, const int* internalToLeaf,
                  const uint32_t* layout, const T* centers, const T* sizes, float ext, uint32_t ngmax,
                  uint32_t* neighbors, uint32_t* neighborsCount, void* stream)
{
    if (lastId <= firstId) return 0;
    if (!x || !y || !z || !h || !box || !childOffsets || !internalToLeaf || !layout || !centers || !sizes ||
        !neighborsCount || (ngmax && !neighbors))
        return CSTONE_E_INVALID_ARGUMENT;
    NbParams<T> p{x,         y,         z,      h,       childOffsets, internalToLeaf, layout, centers,
                  sizes,     neighbors, neighborsCount, firstId, lastId,       ngmax,          ext,    makeDBox<T>(box)};
    bool anyPbc       = p.box.pbc[0] || p.box.pbc[1] || p.box.pbc[2];
    uint32_t groups   = (lastId - firstId + 31u) / 32u;
    unsigned blocks   = gridFor(groups, kNbWarpsPerBlock, 6);
    cudaStream_t s    = asStream(stream);
    if (anyPbc) findNeighborsKernel<T, true><<<blocks, kNbWarpsPerBlock * 32, 0, s>>>(p);
    else findNeighborsKernel<T, false><<<blocks, kNbWarpsPerBlock * 32, 0, s>>>(p);
    CSB_LAUNCHED();
    return 0;
}

} // namespace csb

extern "C" {

int cstone_b200_find_neighbors_f32(const float* x, const float* y, const float* z, const float* h, uint32_t firstId,
                                   uint32_t lastId, const cstone_box_f32* box, const int32_t* childOffsets,
                                   const int32_t* internalToLeaf, const uint32_t* layout, const float* centers,
                                   const float* sizes, float ext, uint32_t ngmax, uint32_t* neighbors,
                                   uint32_t* neighborsCount, void* stream)
{
    return csb::findNeighbors<float>(x, y, z, h, firstId, lastId, box, childOffsets, internalToLeaf, layout, centers,
                                     sizes, ext, ngmax, neighbors, neighborsCount, stream);
}
int cstone_b200_find_neighbors_f64(const double* x, const double* y, const double* z, const double* h,
                                   uint32_t firstId, uint32_t lastId, const cstone_box_f64* box,
                                   const int32_t* childOffsets, const int32_t* internalToLeaf, const uint32_t* layout,
                                   const double* centers, const double* sizes, float ext, uint32_t ngmax,
                                   uint32_t* neighbors, uint32_t* neighborsCount, void* stream)
{
    return csb::findNeighbors<double>(x, y, z, h, firstId, lastId, box, childOffsets, internalToLeaf, layout, centers,
                                      sizes, ext, ngmax, neighbors, neighborsCount, stream);
}

} // extern "C"
