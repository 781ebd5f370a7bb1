/*
 * neighbors.cu — a15: radius neighbour search with the reference's CPU semantics and output layout.
 *
 * Mirrors findNeighbors [ref: findneighbors.hpp:95-188]: for particle i all j != i in the visited leaves with
 * |x_j - x_i|^2 < 4 h_i^2 (strict), PBC fold with rint when the 2h cube of i leaves the box, at most ngmax
 * indices stored at neighbors[(i-first)*ngmax + k] in DEPTH-FIRST visiting order, full count returned.
 * Exact arithmetic (no FMA, same association) follows distanceSq [ref: findneighbors.hpp:51-78], minDistance /
 * insideBox [ref: traversal/boxoverlap.hpp:185-217], applyPbc [ref: sfc/box.hpp:194-206] and norm2, which is a
 * right fold x*x + (y*y + z*z) [ref: util/array.hpp:253-256,333-336].
 *
 * B200 design
 *  - one warp walks the tree for 32 consecutive (SFC-adjacent) targets.  Every stack entry carries the mask of
 *    lanes whose OWN depth-first walk reaches that node, so each lane sees exactly the node sequence (and
 *    therefore the neighbour order) of the reference's per-particle walk, while node data and leaf particles
 *    are fetched once per warp: the 8 children of a node and the particles of a leaf are loaded with coalesced
 *    loads, staged in shared memory and read back as broadcasts.
 *  - double precision: every decision (node overlap, pair distance) is first taken by a float32 FILTER on
 *    coordinates relative to a warp-local origin, with a rigorous error bound (see "filter bound" below).  Only
 *    decisions inside the error band are re-evaluated with the reference's exact double expression, so the
 *    result is bit-identical to the reference while the FP64 pipe is almost idle.
 *  - neighbour indices are collected in a per-warp shared-memory buffer and written to the particle-major
 *    output rows with coalesced warp-wide stores.
 *
 * filter bound (u = 2^-24).  All staged values are float(v - c) with c = target of lane 0.  With Q >= |v - c|
 * for every staged value of a node / leaf and D >= |x_i - c| for the 32 targets, each coordinate difference
 * carries an absolute error <= 3.1 u (Q + D); we use e = 8 u (Q + D).  For a squared distance d2 compared
 * against r2 the float result d2f satisfies |d2f - d2| <= E(d2) = 2 sqrt(3) e sqrt(d2) + 3 e^2 + 4u d2
 * (monotone in d2), so  d2f < r2f - m  =>  d2 < r2  and  d2f > r2f + m  =>  d2 >= r2  for
 * m = 6.93 sqrt(r2) e + 8 e^2 + 16u r2  (>= 2 E(r2) + conversion errors of r2f).  Anything else, NaNs
 * included, is "in the band" and decided exactly.
 */
#include "common.cuh"

namespace csb
{

#ifndef NB_MINBLOCKS
#define NB_MINBLOCKS 6
#endif
constexpr int kNbWarps     = 4;
constexpr int kNbThreads   = kNbWarps * 32;
constexpr int kNbStack     = 160; // >= 7 * 21 + 1 entries can be pending at any time
constexpr int kNbRing      = 32;  // chunks whose hit masks are collected before they are expanded into indices

template<class T>
__device__ inline T absT(T v)
{
    return fabs(v);
}
__device__ inline float absT(float v) { return fabsf(v); }

template<class T>
struct NbFilter
{
    static constexpr bool value = false;
};
template<>
struct NbFilter<double>
{
    static constexpr bool value = true;
};

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

struct alignas(16) NbWarpSmem
{
    float4 candA[16];                       // staged leaf particles, two per entry: x0,x1,y0,y1 ...
    float2 candB[16];                       // ... and z0,z1 (relative float coordinates, or raw floats)
    float rec[8 * 8];                       // per child: centre xyz, size xyz, e, 8 e^2
    int2 cinfo[8];                          // per child: leaf -> particle range [x,y); internal -> (node, -1)
    int stackNode[kNbStack];
    unsigned stackMask[kNbStack];
    uint32_t hitMask[kNbRing][32];          // per processed chunk and lane: bit k = candidate base+k is a neighbour
    uint32_t hitBase[kNbRing];              // first particle index of the chunk
};

// packed 2 x f32 arithmetic of sm_100 (FADD2 / FMUL2 / FFMA2): two candidates per instruction
__device__ inline uint64_t pack2(float lo, float hi)
{
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ inline void unpack2(uint64_t v, float& lo, float& hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ inline uint64_t add2(uint64_t a, uint64_t b)
{
    uint64_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ inline uint64_t fma2(uint64_t a, uint64_t b, uint64_t c)
{
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

__device__ inline float warpMaxF(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

//! the reference's pair test, exact arithmetic  [ref: findneighbors.hpp:51-78,126-140]
template<class T, bool AnyPbc>
__device__ inline bool exactPair(const NbParams<T>& p, T xj, T yj, T zj, T xi, T yi, T zi, bool usePbc, T radiusSq)
{
    T dx = xj - xi;
    T dy = yj - yi;
    T dz = zj - zi;
    if (AnyPbc && usePbc)
    {
        dx -= T(p.box.pbc[0]) * p.box.len[0] * rint(dx * p.box.ilen[0]);
        dy -= T(p.box.pbc[1]) * p.box.len[1] * rint(dy * p.box.ilen[1]);
        dz -= T(p.box.pbc[2]) * p.box.len[2] * rint(dz * p.box.ilen[2]);
    }
    T d2 = dx * dx + dy * dy + dz * dz;
    return d2 < radiusSq;
}

//! the reference's node test, exact arithmetic  [ref: traversal/boxoverlap.hpp:197-217, findneighbors.hpp:120-124]
template<class T, bool AnyPbc>
__device__ inline bool exactNode(const NbParams<T>& p, int idx, T xi, T yi, T zi, bool usePbc, T cellRadiusSq)
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
}

template<class T, bool AnyPbc>
__global__ void __launch_bounds__(kNbThreads, NB_MINBLOCKS) findNeighborsKernel(NbParams<T> p)
{
    constexpr bool F = NbFilter<T>::value;
    __shared__ NbWarpSmem smemAll[kNbWarps];
    NbWarpSmem& sm = smemAll[threadIdx.x >> 5];

    const unsigned lane       = laneId();
    const unsigned full       = 0xffffffffu;
    const uint32_t numTargets = p.lastId - p.firstId;
    const uint32_t numGroups  = (numTargets + 31u) / 32u;
    const uint32_t warpsTotal = gridDim.x * kNbWarps;
    constexpr float kU8       = 4.76837158203125e-7f;  // 8u = 2^-21
    constexpr float kViol     = 3145728.0f;            // 1.5 / 8u: staged magnitude limit as a multiple of e
    const float4* rec4        = reinterpret_cast<const float4*>(sm.rec);

    for (uint32_t group = blockIdx.x * kNbWarps + (threadIdx.x >> 5); group < numGroups; group += warpsTotal)
    {
        const uint32_t w  = group * 32u + lane;
        const bool valid  = w < numTargets;
        const uint32_t i  = p.firstId + (valid ? w : numTargets - 1);
        const T xi = p.x[i], yi = p.y[i], zi = p.z[i], hi = p.h[i];
        const T radiusSq     = T(4.0) * hi * hi;
        const T cellRadiusSq = radiusSq * p.ext * p.ext;
        uint32_t count = 0;

        bool usePbc = false;
        if (AnyPbc)
        {
            T r2h       = T(2) * hi;
            bool inside = (xi - r2h >= p.box.lo[0]) && (yi - r2h >= p.box.lo[1]) && (zi - r2h >= p.box.lo[2]) &&
                          (xi + r2h <= p.box.hi[0]) && (yi + r2h <= p.box.hi[1]) && (zi + r2h <= p.box.hi[2]);
            usePbc = !inside;
        }

        // ---- float filter constants of this lane (double instantiation only)
        T c0x = 0, c0y = 0, c0z = 0;
        float ax = 0, ay = 0, az = 0, D = 0;
        float r2f = 0, Ap = 0, Bp = 0, c2f = 0, Ac = 0, Bc = 0;
        bool exactLane = true;
        if constexpr (F)
        {
            c0x = __shfl_sync(full, xi, 0);
            c0y = __shfl_sync(full, yi, 0);
            c0z = __shfl_sync(full, zi, 0);
            ax  = float(xi - c0x);
            ay  = float(yi - c0y);
            az  = float(zi - c0z);
            D   = warpMaxF(fmaxf(fabsf(ax), fmaxf(fabsf(ay), fabsf(az))));
            r2f = float(radiusSq);
            Ap  = __double2float_ru(6.93 * sqrt(double(radiusSq)));
            Bp  = __double2float_ru(double(radiusSq) * 9.5367431640625e-7);
            c2f = float(cellRadiusSq);
            Ac  = __double2float_ru(6.93 * sqrt(double(cellRadiusSq)));
            Bc  = __double2float_ru(double(cellRadiusSq) * 9.5367431640625e-7);
            exactLane = (AnyPbc && usePbc) || !(radiusSq >= T(1e-30) && cellRadiusSq >= T(1e-30) &&
                                                 radiusSq <= T(1e30) && cellRadiusSq <= T(1e30)) ||
                        ax != ax || ay != ay || az != az;
        }
        const uint64_t nax2 = pack2(-ax, -ax), nay2 = pack2(-ay, -ay), naz2 = pack2(-az, -az), nr2 = pack2(-r2f, -r2f);

        uint32_t* const outRow = p.neighbors + size_t(valid ? w : 0) * p.ngmax;

        /* Expand the collected hit masks into neighbour indices.  Every lane walks ITS column of the ring as one
         * flat loop (next non-empty word, or next set bit), so the trip count is the per-lane total and not the sum
         * of per-chunk maxima over the warp; order = chunk order, then ascending particle index. */
        int ring       = 0; // chunks in the ring (warp-uniform)
        uint32_t ringNz = 0; // bit r: this lane has hits in ring slot r
        auto drain = [&]()
        {
            __syncwarp();
            uint32_t nz = ringNz, m = 0, b = 0;
            for (;;)
            {
                if (m == 0)
                {
                    if (nz == 0) break;
                    const int r = __ffs(int(nz)) - 1;
                    nz &= nz - 1u;
                    m = sm.hitMask[r][lane];
                    b = sm.hitBase[r];
                }
                const uint32_t k = uint32_t(__ffs(int(m))) - 1u;
                m &= m - 1u;
                if (count < p.ngmax) outRow[count] = b + k;
                ++count;
            }
            ring   = 0;
            ringNz = 0;
            __syncwarp();
        };

        // all particles of one leaf against the lanes in `mask`; e / e8 = filter error terms of the leaf node
        auto processLeaf = [&](uint32_t jBeg, uint32_t jEnd, unsigned mask, float e, float e8)
        {
            const bool mine = (mask >> lane) & 1u;
            for (uint32_t base = jBeg; base < jEnd; base += 32u)
            {
                const uint32_t cnt = min(32u, jEnd - base);
                bool viol          = false;
                {
                    // stage: pairs of candidates as (x0,x1,y0,y1) + (z0,z1); slots >= cnt hold a far-away dummy
                    float bx = 1.0e18f, by = 1.0e18f, bz = 1.0e18f;
                    if (lane < cnt)
                    {
                        const uint32_t j = base + lane;
                        const T xj = p.x[j], yj = p.y[j], zj = p.z[j];
                        if constexpr (F)
                        {
                            bx = float(xj - c0x), by = float(yj - c0y), bz = float(zj - c0z);
                            const float mx = fmaxf(fabsf(bx), fmaxf(fabsf(by), fabsf(bz)));
                            viol           = !(mx <= e * kViol) || bx != bx || by != by || bz != bz;
                        }
                        else { bx = float(xj), by = float(yj), bz = float(zj); }
                    }
                    float* cA = reinterpret_cast<float*>(sm.candA) + (lane >> 1) * 4 + (lane & 1u);
                    cA[0]     = bx;
                    cA[2]     = by;
                    reinterpret_cast<float*>(sm.candB)[lane] = bz;
                }
                const bool anyViol = F ? bool(__any_sync(full, viol)) : false;
                __syncwarp();
                uint32_t hm = 0;
                if (mine)
                {
                    const int self = int(i - base); // slot of the target itself, if it lives in this chunk
                    if constexpr (F)
                    {
                        bool exact = exactLane || anyViol;
                        if (!exact)
                        {
                            const float m    = __fmaf_rn(Ap, e, Bp) + e8;
                            const float negm = -m;
                            float minAbs     = __int_as_float(0x7f800000);
#pragma unroll
                            for (int blk = 0; blk < 4; ++blk)
                            {
                                if (uint32_t(8 * blk) >= cnt) break; // warp-uniform
#pragma unroll
                                for (int q = 0; q < 4; ++q)
                                {
                                    const int kk      = 4 * blk + q;
                                    const float4 ca   = sm.candA[kk];
                                    const float2 cb   = sm.candB[kk];
                                    const uint64_t dx = add2(pack2(ca.x, ca.y), nax2);
                                    const uint64_t dy = add2(pack2(ca.z, ca.w), nay2);
                                    const uint64_t dz = add2(pack2(cb.x, cb.y), naz2);
                                    float t0, t1; // d2 - r2 of the two candidates
                                    unpack2(fma2(dz, dz, fma2(dy, dy, fma2(dx, dx, nr2))), t0, t1);
                                    if (t0 < negm) hm |= 1u << (2 * kk);
                                    if (t1 < negm) hm |= 2u << (2 * kk);
                                    minAbs = fminf(minAbs, fminf(fabsf(t0), fabsf(t1)));
                                }
                            }
                            exact = !(minAbs > m); // some pair inside the error band (or m is NaN): decide exactly
                        }
                        if (exact)
                        {
                            hm = 0;
                            for (uint32_t k = 0; k < cnt; ++k)
                            {
                                const uint32_t j = base + k;
                                if (exactPair<T, AnyPbc>(p, p.x[j], p.y[j], p.z[j], xi, yi, zi, usePbc, radiusSq))
                                    hm |= 1u << k;
                            }
                        }
                    }
                    else
                    {
#pragma unroll 4
                        for (uint32_t k = 0; k < cnt; ++k)
                        {
                            const float* cA = reinterpret_cast<const float*>(sm.candA) + (k >> 1) * 4 + (k & 1u);
                            const float cz  = reinterpret_cast<const float*>(sm.candB)[k];
                            if (exactPair<T, AnyPbc>(p, T(cA[0]), T(cA[2]), T(cz), xi, yi, zi, usePbc, radiusSq))
                                hm |= 1u << k;
                        }
                    }
                    if (self >= 0 && self < 32) hm &= ~(1u << self);
                }
                sm.hitMask[ring][lane] = hm;
                if (lane == 0) sm.hitBase[ring] = base;
                ringNz |= (hm ? 1u : 0u) << ring;
                if (++ring == kNbRing) drain();
                else __syncwarp();
            }
        };

        // coalesced load of the 8 children of a node (24 centre + 24 size values, child offsets, leaf ranges)
        auto stageChildren = [&](int fc) -> unsigned
        {
            if (lane < 24u)
            {
                const T cd      = p.centers[3 * size_t(fc) + lane];
                const T sd      = p.sizes[3 * size_t(fc) + lane];
                const int child = int(lane) / 3, dim = int(lane) - 3 * child;
                if constexpr (F)
                {
                    const T c0                  = dim == 0 ? c0x : (dim == 1 ? c0y : c0z);
                    sm.rec[child * 8 + dim]     = float(cd - c0);
                    sm.rec[child * 8 + 3 + dim] = __double2float_ru(sd);
                }
                else
                {
                    sm.rec[child * 8 + dim]     = float(cd);
                    sm.rec[child * 8 + 3 + dim] = float(sd);
                }
            }
            int co = 1;
            if (lane < 8u)
            {
                co = p.childOffsets[fc + int(lane)];
                if (co == 0)
                {
                    const int leaf = p.internalToLeaf[fc + int(lane)];
                    sm.cinfo[lane] = make_int2(int(p.layout[leaf]), int(p.layout[leaf + 1]));
                }
                else { sm.cinfo[lane] = make_int2(fc + int(lane), -1); }
            }
            const unsigned leafBits = __ballot_sync(full, co == 0);
            __syncwarp();
            if constexpr (F)
            {
                if (lane < 8u)
                {
                    const float* r = sm.rec + lane * 8;
                    const float Q  = fmaxf(fabsf(r[0]) + r[3], fmaxf(fabsf(r[1]) + r[4], fabsf(r[2]) + r[5]));
                    const float e  = kU8 * (Q + D);
                    sm.rec[lane * 8 + 6] = e;
                    sm.rec[lane * 8 + 7] = 8.0f * e * e;
                }
                __syncwarp();
            }
            return leafBits;
        };

        // does the search sphere of this lane overlap staged child o (= node `child`)?
        auto overlapChild = [&](int o, int child) -> bool
        {
            if constexpr (F)
            {
                if (exactLane) return exactNode<T, AnyPbc>(p, child, xi, yi, zi, usePbc, cellRadiusSq);
                const float4 r0 = rec4[o * 2], r1 = rec4[o * 2 + 1];
                const float dx  = fmaxf(fabsf(r0.x - ax) - r0.w, 0.f);
                const float dy  = fmaxf(fabsf(r0.y - ay) - r1.x, 0.f);
                const float dz  = fmaxf(fabsf(r0.z - az) - r1.y, 0.f);
                const float d2  = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
                const float m   = __fmaf_rn(Ac, r1.z, Bc) + r1.w;
                if (d2 < c2f - m) return true;
                if (d2 > c2f + m) return false;
                return exactNode<T, AnyPbc>(p, child, xi, yi, zi, usePbc, cellRadiusSq);
            }
            else
            {
                const float* r = sm.rec + o * 8; // raw float centre / size: the exact expression on staged data
                T dx = T(r[0]) - xi, dy = T(r[1]) - yi, dz = T(r[2]) - zi;
                if (AnyPbc && usePbc)
                {
                    dx -= T(p.box.pbc[0]) * p.box.len[0] * rint(dx * p.box.ilen[0]);
                    dy -= T(p.box.pbc[1]) * p.box.len[1] * rint(dy * p.box.ilen[1]);
                    dz -= T(p.box.pbc[2]) * p.box.len[2] * rint(dz * p.box.ilen[2]);
                }
                dx = absT(dx) - T(r[3]);
                dy = absT(dy) - T(r[4]);
                dz = absT(dz) - T(r[5]);
                dx += absT(dx);
                dy += absT(dy);
                dz += absT(dz);
                dx *= T(0.5);
                dy *= T(0.5);
                dz *= T(0.5);
                return dx * dx + (dy * dy + dz * dz) < cellRadiusSq;
            }
        };

        const unsigned rootMask =
            __ballot_sync(full, valid && exactNode<T, AnyPbc>(p, 0, xi, yi, zi, usePbc, cellRadiusSq));
        if (rootMask)
        {
            if (p.childOffsets[0] == 0)
            {
                // single-node tree: decide everything exactly (infinite error terms put every pair in the band)
                const int leaf = p.internalToLeaf[0];
                processLeaf(p.layout[leaf], p.layout[leaf + 1], rootMask, __int_as_float(0x7f800000),
                            __int_as_float(0x7f800000));
            }
            else
            {
                if (lane == 0)
                {
                    sm.stackNode[0] = 0;
                    sm.stackMask[0] = 0;
                }
                int sp        = 1;
                int node      = 0;
                unsigned mask = rootMask;
                do
                {
                    const int firstChild    = p.childOffsets[node];
                    const unsigned leafBits = stageChildren(firstChild);
                    const bool mine         = (mask >> lane) & 1u;
#pragma unroll 1
                    for (int o = 0; o < 8; ++o)
                    {
                        const int child      = firstChild + o;
                        const unsigned cmask = __ballot_sync(full, mine && overlapChild(o, child));
                        if (!cmask) continue;
                        if ((leafBits >> o) & 1u)
                        {
                            const int2 r = sm.cinfo[o];
                            processLeaf(uint32_t(r.x), uint32_t(r.y), cmask, sm.rec[o * 8 + 6], sm.rec[o * 8 + 7]);
                        }
                        else
                        {
                            if (lane == 0)
                            {
                                sm.stackNode[sp] = child;
                                sm.stackMask[sp] = cmask;
                            }
                            ++sp;
                        }
                    }
                    __syncwarp();
                    --sp;
                    node = sm.stackNode[sp];
                    mask = sm.stackMask[sp];
                    __syncwarp();
                } while (node != 0);
            }
        }
        if (ring) drain();
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
    unsigned blocks   = gridFor(groups, kNbWarps, NB_MINBLOCKS);
    cudaStream_t s    = asStream(stream);
    if (anyPbc) findNeighborsKernel<T, true><<<blocks, kNbThreads, 0, s>>>(p);
    else findNeighborsKernel<T, false><<<blocks, kNbThreads, 0, s>>>(p);
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
