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
#define NB_MINBLOCKS 7
#endif
constexpr int kNbWarps     = 4;
constexpr int kNbThreads   = kNbWarps * 32;
constexpr int kNbStack     = 160; // >= 7 * 21 + 1 entries can be pending at any time
#ifndef NB_RING
#define NB_RING 32
#endif
#ifndef NB_ASYNC
#define NB_ASYNC 1 // workspace kernels: cp.async pipeline for candidate chunks and node records
#endif
#ifndef NB_MERGE
#define NB_MERGE 1 // workspace kernels: consecutive reached leaf children that fit into one chunk are tested together
#endif
#ifndef NB_PFNODE
#define NB_PFNODE 0
#endif
#ifndef NB_PFLEAF
#define NB_PFLEAF 0
#endif
constexpr int kNbRing      = NB_RING;  // chunks whose hit masks are collected before they are expanded into indices

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
    const int2* nodeInfo; // optional (workspace entry points): per node (first child, -1) or the particle range of a leaf
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
    int2 cinfo[8];                          // per child: leaf -> particle range [x,y); internal -> (its first child, -1)
    double c0[3];                           // warp-local origin of the float filter (target of lane 0)
    float D;                                // max |target - origin| over the warp
    float2 eMax;                            // largest e (and 8 e^2) among the 8 staged children
    int stackFc[kNbStack];                  // pending internal nodes: index of their first child ...
    unsigned stackMask[kNbStack];           // ... and the lanes whose walk reaches them
    uint32_t hitMask[kNbRing][32];          // per processed chunk and lane: bit k = candidate base+k is a neighbour
    uint32_t hitBase[kNbRing];              // first particle index of the chunk
};

/* Landing zones of the asynchronous copies (workspace kernels only): the raw coordinates of the NEXT chunk of
 * candidates and the raw records of the NEXT node arrive here (cp.async) while the current chunk is tested. */
template<class T>
struct alignas(16) NbWarpRaw
{
    T cand[3][32];
    T centers[24];
    T sizes[24];
    int2 info[8];
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
__device__ inline uint64_t mul2(uint64_t a, uint64_t b)
{
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ inline uint64_t fma2(uint64_t a, uint64_t b, uint64_t c)
{
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

//! predicated store without a branch: row[count] = v if count < limit
__device__ inline void storeIfBelow(uint32_t* row, uint32_t v, uint32_t count, uint32_t limit)
{
    asm volatile("{\n\t.reg .pred p;\n\t.reg .u64 a;\n\tsetp.lt.u32 p, %2, %3;\n\tmad.wide.u32 a, %2, 4, %0;\n\t"
                 "@p st.global.u32 [a], %1;\n\t}"
                 :
                 : "l"(row), "r"(v), "r"(count), "r"(limit)
                 : "memory");
}
__device__ inline void prefetchL1(const void* ptr) { asm volatile("prefetch.global.L1 [%0];" ::"l"(ptr)); }

template<int Bytes>
__device__ inline void cpAsync(void* smemDst, const void* src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(uint32_t(__cvta_generic_to_shared(smemDst))), "l"(src),
                 "n"(Bytes)
                 : "memory");
}
__device__ inline void cpAsyncCommit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template<int N>
__device__ inline void cpAsyncWait()
{
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__device__ inline float warpMaxF(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

//! per-target quantities of the reference's search [ref: findneighbors.hpp:104-118]
template<class T>
struct NbTarget
{
    T x, y, z, radiusSq, cellRadiusSq;
    bool usePbc;
};

template<class T, bool AnyPbc>
__device__ inline NbTarget<T> loadTarget(const NbParams<T>& p, uint32_t i)
{
    NbTarget<T> t;
    t.x = p.x[i], t.y = p.y[i], t.z = p.z[i];
    const T hi     = p.h[i];
    t.radiusSq     = T(4.0) * hi * hi;
    t.cellRadiusSq = t.radiusSq * p.ext * p.ext;
    t.usePbc       = false;
    if (AnyPbc)
    {
        T r2h       = T(2) * hi;
        bool inside = (t.x - r2h >= p.box.lo[0]) && (t.y - r2h >= p.box.lo[1]) && (t.z - r2h >= p.box.lo[2]) &&
                      (t.x + r2h <= p.box.hi[0]) && (t.y + r2h <= p.box.hi[1]) && (t.z + r2h <= p.box.hi[2]);
        t.usePbc = !inside;
    }
    return t;
}

//! the reference's pair test, exact arithmetic  [ref: findneighbors.hpp:51-78,126-140]
template<class T, bool AnyPbc>
__device__ inline bool exactPair(const NbParams<T>& p, T xj, T yj, T zj, const NbTarget<T>& t)
{
    T dx = xj - t.x;
    T dy = yj - t.y;
    T dz = zj - t.z;
    if (AnyPbc && t.usePbc)
    {
        dx -= T(p.box.pbc[0]) * p.box.len[0] * rint(dx * p.box.ilen[0]);
        dy -= T(p.box.pbc[1]) * p.box.len[1] * rint(dy * p.box.ilen[1]);
        dz -= T(p.box.pbc[2]) * p.box.len[2] * rint(dz * p.box.ilen[2]);
    }
    T d2 = dx * dx + dy * dy + dz * dz;
    return d2 < t.radiusSq;
}

//! the reference's node test, exact arithmetic  [ref: traversal/boxoverlap.hpp:197-217, findneighbors.hpp:120-124]
template<class T, bool AnyPbc>
__device__ inline bool exactNode(const NbParams<T>& p, int idx, const NbTarget<T>& t)
{
    T dx = p.centers[3 * idx] - t.x;
    T dy = p.centers[3 * idx + 1] - t.y;
    T dz = p.centers[3 * idx + 2] - t.z;
    if (AnyPbc && t.usePbc)
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
    return dx * dx + (dy * dy + dz * dz) < t.cellRadiusSq;
}

/* Out-of-line exact decisions of the filtered (double) kernel.  They re-read the target from global memory, so that
 * the main path keeps no double-precision target state in registers (64 registers / 8 blocks per SM). */
template<class T, bool AnyPbc>
__device__ __noinline__ bool exactNodeSlow(const NbParams<T>& p, int idx, uint32_t i)
{
    const NbTarget<T> t = loadTarget<T, AnyPbc>(p, i);
    return exactNode<T, AnyPbc>(p, idx, t);
}
template<class T, bool AnyPbc>
__device__ __noinline__ uint32_t exactChunkSlow(const NbParams<T>& p, uint32_t base, uint32_t cnt, uint32_t i)
{
    const NbTarget<T> t = loadTarget<T, AnyPbc>(p, i);
    uint32_t hm         = 0;
    for (uint32_t k = 0; k < cnt; ++k)
    {
        const uint32_t j = base + k;
        if (exactPair<T, AnyPbc>(p, p.x[j], p.y[j], p.z[j], t)) hm |= 1u << k;
    }
    return hm;
}

/* Workspace variant: one record per node that answers what the walk needs next, so that staging a node costs ONE
 * round of independent loads instead of the chain childOffsets -> internalToLeaf -> layout. */
__global__ void nodeInfoKernel(const int* childOffsets, const int* internalToLeaf, const uint32_t* layout, int numNodes,
                               int2* nodeInfo)
{
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < numNodes; idx += gridDim.x * blockDim.x)
    {
        const int co = childOffsets[idx];
        int2 r       = make_int2(co, -1);
        if (co == 0)
        {
            const int leaf = internalToLeaf[idx];
            r              = make_int2(int(layout[leaf]), int(layout[leaf + 1]));
        }
        nodeInfo[idx] = r;
    }
}

template<class T, bool AnyPbc, bool Info>
__global__ void __launch_bounds__(kNbThreads, NB_MINBLOCKS) findNeighborsKernel(const __grid_constant__ NbParams<T> p)
{
    constexpr bool F     = NbFilter<T>::value;
    constexpr bool Async = Info && NB_ASYNC;
    __shared__ NbWarpSmem smemAll[kNbWarps];
    __shared__ NbWarpRaw<T> rawAll[Async ? kNbWarps : 1];
    NbWarpSmem& sm    = smemAll[threadIdx.x >> 5];
    NbWarpRaw<T>& raw = rawAll[Async ? (threadIdx.x >> 5) : 0];

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
        uint32_t count    = 0;

        // ---- per-lane constants.  Double kernel: only floats survive this block (the exact paths reload the target);
        //      float kernel: the target itself (tg) is the working set.
        NbTarget<T> tg;
        float c2f = 0;
        float Ap = 0, Bp = 0, Ac = 0, Bc = 0;
        uint64_t nax2 = 0, nay2 = 0, naz2 = 0, nr2 = 0;
        bool exactLane = true;
        unsigned rootMask;
        {
            const NbTarget<T> t0 = loadTarget<T, AnyPbc>(p, i);
            if constexpr (F)
            {
                const T c0x = __shfl_sync(full, t0.x, 0), c0y = __shfl_sync(full, t0.y, 0), c0z = __shfl_sync(full, t0.z, 0);
                const float ax = float(t0.x - c0x), ay = float(t0.y - c0y), az = float(t0.z - c0z);
                const float D  = warpMaxF(fmaxf(fabsf(ax), fmaxf(fabsf(ay), fabsf(az))));
                __syncwarp();
                if (lane == 0)
                {
                    sm.c0[0] = c0x, sm.c0[1] = c0y, sm.c0[2] = c0z;
                    sm.D     = D;
                }
                const float r2f = float(t0.radiusSq);
                c2f = float(t0.cellRadiusSq);
                const float4 lc = make_float4(__double2float_ru(6.93 * sqrt(double(t0.radiusSq))),
                                              __double2float_ru(double(t0.radiusSq) * 9.5367431640625e-7),
                                              __double2float_ru(6.93 * sqrt(double(t0.cellRadiusSq))),
                                              __double2float_ru(double(t0.cellRadiusSq) * 9.5367431640625e-7));
                Ap = lc.x, Bp = lc.y, Ac = lc.z, Bc = lc.w;
                exactLane = (AnyPbc && t0.usePbc) ||
                            !(t0.radiusSq >= T(1e-30) && t0.cellRadiusSq >= T(1e-30) && t0.radiusSq <= T(1e30) &&
                              t0.cellRadiusSq <= T(1e30)) ||
                            ax != ax || ay != ay || az != az;
                nax2 = pack2(-ax, -ax), nay2 = pack2(-ay, -ay), naz2 = pack2(-az, -az), nr2 = pack2(-r2f, -r2f);
            }
            else { tg = t0; }
            rootMask = __ballot_sync(full, valid && exactNode<T, AnyPbc>(p, 0, t0));
        }
        __syncwarp();

        uint32_t* const outRow = p.neighbors + size_t(valid ? w : 0) * p.ngmax;

        /* Expand the collected hit masks into neighbour indices.  Every lane walks ITS column of the ring as one
         * flat loop (next non-empty word, or next set bit), so the trip count is the per-lane total and not the sum
         * of per-chunk maxima over the warp; order = chunk order, then ascending particle index. */
        int ring        = 0; // chunks in the ring (warp-uniform)
        uint32_t ringNz = 0; // bit r: this lane has hits in ring slot r
        auto drain = [&]()
        {
            __syncwarp();
            uint32_t nz = ringNz, m = 0, b = 0;
            const uint32_t ngmax = p.ngmax;
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
                storeIfBelow(outRow, b + k, count, ngmax);
                ++count;
            }
            ring   = 0;
            ringNz = 0;
            __syncwarp();
        };

        /* stage one chunk of <= 32 candidates: pairs of candidates as (x0,x1,y0,y1) + (z0,z1), relative float
         * coordinates (or raw floats); slots >= cnt hold a far-away dummy.  Returns whether a staged value violates
         * the magnitude limit of the filter bound (then the chunk is decided exactly). */
        auto stageChunk = [&](uint32_t cnt, T xj, T yj, T zj, float e) -> bool
        {
            bool viol = false;
            float bx = 1.0e18f, by = 1.0e18f, bz = 1.0e18f;
            if (lane < cnt)
            {
                if constexpr (F)
                {
                    bx = float(xj - sm.c0[0]), by = float(yj - sm.c0[1]), bz = float(zj - sm.c0[2]);
                    const float mx = fmaxf(fabsf(bx), fmaxf(fabsf(by), fabsf(bz)));
                    viol           = !(mx <= e * kViol) || bx != bx || by != by || bz != bz;
                }
                else { bx = float(xj), by = float(yj), bz = float(zj); }
            }
            float* cA = reinterpret_cast<float*>(sm.candA) + (lane >> 1) * 4 + (lane & 1u);
            cA[0]     = bx;
            cA[2]     = by;
            reinterpret_cast<float*>(sm.candB)[lane] = bz;
            const bool anyViol = F ? bool(__any_sync(full, viol)) : false;
            __syncwarp();
            return anyViol;
        };

        // the staged chunk [base, base + cnt) against this lane (if `mine`); the hit mask goes into the ring
        auto testChunk = [&](uint32_t base, uint32_t cnt, bool mine, bool anyViol, float e, float e8,
                             uint32_t vmask = 0xffffffffu)
        {
            uint32_t hm = 0;
            if (mine)
            {
                const uint32_t self = i - base; // slot of the target itself, if it lives in this chunk (< 32)
                if constexpr (F)
                {
                    bool exact = exactLane || anyViol;
                    if (!exact)
                    {
                        const float m = __fmaf_rn(Ap, e, Bp) + e8;
                        /* s = d2 - r2 + m per candidate: s < 0 is a sure hit (the sign bit is shifted into hm),
                         * 0 <= s <= 2m is the error band (found as the smallest non-negative s, compared as
                         * unsigned bit patterns), s > 2m a sure miss.  Rounding m - r2 costs <= u r2, inside the
                         * slack of m. */
                        float r2neg, dummy;
                        unpack2(nr2, r2neg, dummy);
                        const float cm    = m + r2neg;
                        const uint64_t c2 = pack2(cm, cm);
                        uint32_t minBits  = 0xffffffffu;
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
                                float s0, s1;
                                unpack2(fma2(dz, dz, fma2(dy, dy, fma2(dx, dx, c2))), s0, s1);
                                hm      = __funnelshift_l(__float_as_uint(s0), hm, 1);
                                hm      = __funnelshift_l(__float_as_uint(s1), hm, 1);
                                minBits = min(minBits, min(__float_as_uint(s0), __float_as_uint(s1)));
                            }
                        }
                        // candidate k sits at bit n-1-k, n = slots processed: back to bit k
                        hm    = __brev(hm) >> (32u - min(32u, (cnt + 7u) & ~7u));
                        exact = !(minBits > __float_as_uint(m + m));
                    }
                    if (exact) hm = exactChunkSlow<T, AnyPbc>(p, base, cnt, i);
                }
                else if (AnyPbc && tg.usePbc)
                {
#pragma unroll 4
                    for (uint32_t k = 0; k < cnt; ++k)
                    {
                        const float* cA = reinterpret_cast<const float*>(sm.candA) + (k >> 1) * 4 + (k & 1u);
                        const float cz  = reinterpret_cast<const float*>(sm.candB)[k];
                        if (exactPair<T, AnyPbc>(p, T(cA[0]), T(cA[2]), T(cz), tg)) hm |= 1u << k;
                    }
                }
                else
                {
                    /* float coordinates: the reference's own expression ((dx*dx + dy*dy) + dz*dz < r2, round to nearest,
                     * no contraction) on two candidates per instruction; the sign bit of d2 - r2 is the hit bit (the
                     * difference of two floats is negative exactly when the first is smaller; NaNs come out of the
                     * arithmetic with a clear sign bit = no hit, like the comparison) */
                    const uint64_t ntx = pack2(-float(tg.x), -float(tg.x)), nty = pack2(-float(tg.y), -float(tg.y));
                    const uint64_t ntz = pack2(-float(tg.z), -float(tg.z));
                    const uint64_t nr  = pack2(-float(tg.radiusSq), -float(tg.radiusSq));
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
                            const uint64_t dx = add2(pack2(ca.x, ca.y), ntx);
                            const uint64_t dy = add2(pack2(ca.z, ca.w), nty);
                            const uint64_t dz = add2(pack2(cb.x, cb.y), ntz);
                            float s0, s1;
                            unpack2(add2(add2(add2(mul2(dx, dx), mul2(dy, dy)), mul2(dz, dz)), nr), s0, s1);
                            hm = __funnelshift_l(__float_as_uint(s0), hm, 1);
                            hm = __funnelshift_l(__float_as_uint(s1), hm, 1);
                        }
                    }
                    hm = __brev(hm) >> (32u - min(32u, (cnt + 7u) & ~7u));
                    if (cnt < 32u) hm &= (1u << cnt) - 1u; // padding slots hold a far-away dummy, not a non-candidate
                }
                hm &= vmask & ~((self < 32u ? 1u : 0u) << (self & 31u));
            }
            sm.hitMask[ring][lane] = hm;
            if (lane == 0) sm.hitBase[ring] = base;
            ringNz |= (hm ? 1u : 0u) << ring;
            if (++ring == kNbRing) drain();
            else __syncwarp();
        };

        // all particles of one leaf against the lanes in `mask`; e / e8 = filter error terms of the leaf node
        auto processLeaf = [&](uint32_t jBeg, uint32_t jEnd, unsigned mask, float e, float e8)
        {
            const bool mine = (mask >> lane) & 1u;
            for (uint32_t base = jBeg; base < jEnd; base += 32u)
            {
                const uint32_t cnt = min(32u, jEnd - base);
                T xj = 0, yj = 0, zj = 0;
                if (lane < cnt)
                {
                    const uint32_t j = base + lane;
                    xj = p.x[j], yj = p.y[j], zj = p.z[j];
                }
                const bool anyViol = stageChunk(cnt, xj, yj, zj, e);
                testChunk(base, cnt, mine, anyViol, e, e8);
            }
        };

        /* Workspace kernels: the reached leaf children of the staged node, chunk by chunk.  The raw coordinates of
         * chunk k+1 travel (cp.async) while chunk k is tested; `bits` = reached children of this lane. */
        /* Next chunk.  Consecutive reached leaf children whose particles are contiguous and fit into ONE chunk of 32
         * are merged into a run (`run` = their child bits): one staging + one test for the whole run, every lane
         * keeps only the candidates of the leaves its own walk reaches (same order: leaf children are visited in
         * ascending child order before any internal child). */
        auto advanceChunk = [&](unsigned& lt, int& o, unsigned& run, uint32_t& base, uint32_t& end) -> bool
        {
            if (o >= 0 && base + 32u < end)
            {
                base += 32u;
                return true;
            }
            while (lt)
            {
                o = __ffs(int(lt)) - 1;
                lt &= lt - 1u;
                const int2 r = sm.cinfo[o];
                base = uint32_t(r.x), end = uint32_t(r.y);
                run = 1u << o;
#if NB_MERGE
                for (int q = o + 1; (lt >> q) & 1u; ++q)
                {
                    const int2 rn = sm.cinfo[q];
                    if (uint32_t(rn.x) != end || uint32_t(rn.y) - base > 32u) break;
                    end = uint32_t(rn.y);
                    run |= 1u << q;
                    lt &= ~(1u << q);
                }
#endif
                if (end > base) return true;
            }
            return false;
        };
        auto issueChunk = [&](uint32_t base, uint32_t end)
        {
            const uint32_t j = base + lane;
            if (j < end && lane < 32u)
            {
                cpAsync<sizeof(T)>(&raw.cand[0][lane], p.x + j);
                cpAsync<sizeof(T)>(&raw.cand[1][lane], p.y + j);
                cpAsync<sizeof(T)>(&raw.cand[2][lane], p.z + j);
            }
        };
        // precondition: the copies of the first chunk (if any) were committed as the second-youngest group
        auto processLeaves = [&](unsigned lt, int o, unsigned run, uint32_t base, uint32_t end, bool have, uint32_t bits)
        {
            bool first = true;
            while (have)
            {
                const uint32_t cbase = base, cnt = min(32u, end - base);
                const int co = o;
                const unsigned crun = run;
                if (first) cpAsyncWait<1>();
                else cpAsyncWait<0>();
                first = false;
                __syncwarp();
                float e = sm.rec[co * 8 + 6], e8 = sm.rec[co * 8 + 7];
                bool mine      = (bits >> co) & 1u;
                uint32_t vmask = 0xffffffffu;
                if (crun & (crun - 1u)) // merged run (warp-uniform): per-lane mask of the candidates in reached leaves
                {
                    vmask = 0;
                    for (unsigned rb = crun; rb; rb &= rb - 1u)
                    {
                        const int q      = __ffs(int(rb)) - 1;
                        const int2 r     = sm.cinfo[q];
                        const uint32_t n = uint32_t(r.y - r.x);
                        e                = fmaxf(e, sm.rec[q * 8 + 6]);
                        e8               = fmaxf(e8, sm.rec[q * 8 + 7]);
                        const uint32_t m = n ? ((0xffffffffu >> (32u - n)) << (uint32_t(r.x) - cbase)) : 0u;
                        if ((bits >> q) & 1u) vmask |= m;
                    }
                    mine = vmask != 0;
                }
                T xj = 0, yj = 0, zj = 0;
                if (lane < cnt) xj = raw.cand[0][lane], yj = raw.cand[1][lane], zj = raw.cand[2][lane];
                const bool anyViol = stageChunk(cnt, xj, yj, zj, e);
                have = advanceChunk(lt, o, run, base, end);
                if (have) issueChunk(base, end);
                cpAsyncCommit();
                testChunk(cbase, cnt, mine, anyViol, e, e8, vmask);
            }
        };
        // raw records of the 8 children fc..fc+7 (centres, sizes, node records) on their way to shared memory
        auto issueNode = [&](int fc)
        {
            if (lane < 24u)
            {
                cpAsync<sizeof(T)>(&raw.centers[lane], p.centers + 3 * size_t(fc) + lane);
                cpAsync<sizeof(T)>(&raw.sizes[lane], p.sizes + 3 * size_t(fc) + lane);
            }
            if (lane < 8u) cpAsync<8>(&raw.info[lane], p.nodeInfo + fc + int(lane));
        };

        /* coalesced load of the 8 children fc..fc+7 of a node: 24 centre + 24 size values, child offsets and, for leaf
         * children, their particle ranges.  All loads that do not depend on each other are issued together. */
        auto stageChildren = [&](int fc) -> unsigned
        {
            if (lane < 24u)
            {
                T cd, sd;
                if constexpr (Async) { cd = raw.centers[lane], sd = raw.sizes[lane]; }
                else { cd = p.centers[3 * size_t(fc) + lane], sd = p.sizes[3 * size_t(fc) + lane]; }
                const int child = int(lane) / 3, dim = int(lane) - 3 * child;
                if constexpr (F)
                {
                    sm.rec[child * 8 + dim]     = float(cd - sm.c0[dim]);
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
                if constexpr (Async)
                {
                    const int2 inf = raw.info[lane];
                    sm.cinfo[lane] = inf;
                    co             = (inf.y == -1) ? inf.x : 0;
#if NB_PFLEAF
                    if (co == 0 && uint32_t(inf.y) > uint32_t(inf.x))
                    {
                        prefetchL1(p.x + uint32_t(inf.x)), prefetchL1(p.y + uint32_t(inf.x));
                        prefetchL1(p.z + uint32_t(inf.x));
                    }
#endif
#if NB_PFNODE
                    if (co != 0)
                    {
                        prefetchL1(p.centers + 3 * size_t(co)), prefetchL1(p.centers + 3 * size_t(co) + 16);
                        prefetchL1(p.sizes + 3 * size_t(co)), prefetchL1(p.sizes + 3 * size_t(co) + 16);
                        prefetchL1(p.nodeInfo + co);
                    }
#endif
                }
                else if constexpr (Info)
                {
                    const int2 inf = p.nodeInfo[fc + int(lane)];
                    sm.cinfo[lane] = inf;
                    co             = (inf.y == -1) ? inf.x : 0;
                    if (co == 0)
                    {
                        if (uint32_t(inf.y) > uint32_t(inf.x))
                        {
                            prefetchL1(p.x + uint32_t(inf.x)), prefetchL1(p.y + uint32_t(inf.x));
                            prefetchL1(p.z + uint32_t(inf.x));
                        }
                    }
                    else
                    {
                        prefetchL1(p.centers + 3 * size_t(co)), prefetchL1(p.centers + 3 * size_t(co) + 16);
                        prefetchL1(p.sizes + 3 * size_t(co)), prefetchL1(p.sizes + 3 * size_t(co) + 16);
                        prefetchL1(p.nodeInfo + co);
                    }
                }
                else
                {
                    co             = p.childOffsets[fc + int(lane)];
                    const int leaf = p.internalToLeaf[fc + int(lane)]; // meaningful for leaf children only
                    if (co == 0)
                    {
                        const uint32_t jb = p.layout[leaf], je = p.layout[leaf + 1];
                        sm.cinfo[lane]    = make_int2(int(jb), int(je));
                        // the particles of this leaf are needed next: pull their first lines into L1
                        if (je > jb)
                        {
                            prefetchL1(p.x + jb), prefetchL1(p.y + jb), prefetchL1(p.z + jb);
                        }
                    }
                    else
                    {
                        sm.cinfo[lane] = make_int2(co, -1);
                        // records of the grandchildren, in case this child is pushed
                        prefetchL1(p.centers + 3 * size_t(co)), prefetchL1(p.centers + 3 * size_t(co) + 16);
                        prefetchL1(p.sizes + 3 * size_t(co)), prefetchL1(p.sizes + 3 * size_t(co) + 16);
                        prefetchL1(p.childOffsets + co), prefetchL1(p.internalToLeaf + co);
    }
                }
            }
            const unsigned leafBits = __ballot_sync(full, co == 0);
            __syncwarp();
            if constexpr (F)
            {
                if (lane < 8u)
                {
                    const float* r = sm.rec + lane * 8;
                    const float Q  = fmaxf(fabsf(r[0]) + r[3], fmaxf(fabsf(r[1]) + r[4], fabsf(r[2]) + r[5]));
                    const float e  = kU8 * (Q + sm.D);
                    sm.rec[lane * 8 + 6] = e;
                    sm.rec[lane * 8 + 7] = 8.0f * e * e;
                    // largest error term of the 8 siblings (non-negative floats order like their bit patterns)
                    const float eMax = __uint_as_float(__reduce_max_sync(0xffu, __float_as_uint(e)));
                    if (lane == 0) sm.eMax = make_float2(eMax, 8.0f * eMax * eMax);
                }
                __syncwarp();
            }
            return leafBits;
        };

        /* which of the 8 staged children fc..fc+7 does the search sphere of this lane overlap?  Bit o of the result.
         * Branch-free over the 8 children; decisions inside the filter's error band are collected and settled with
         * the exact expression afterwards (rare). */
        auto overlapChildren = [&](int fc, bool mine) -> uint32_t
        {
            uint32_t bits = 0;
            if constexpr (F)
            {
                float nax, nay, naz, dummy;
                unpack2(nax2, nax, dummy);
                unpack2(nay2, nay, dummy);
                unpack2(naz2, naz, dummy);
                // one error band [lo, hi] for the 8 siblings, from their largest error term
                const float2 em = sm.eMax;
                const float m   = __fmaf_rn(Ac, em.x, Bc) + em.y;
                const float lo = c2f - m, hi = c2f + m;
                uint32_t in = 0, out = 0; // sign bits of d2 - lo and hi - d2, child o at bit 7 - o
#pragma unroll
                for (int o = 0; o < 8; ++o)
                {
                    const float4 r0 = rec4[o * 2], r1 = rec4[o * 2 + 1];
                    const float dx  = fmaxf(fabsf(r0.x + nax) - r0.w, 0.f);
                    const float dy  = fmaxf(fabsf(r0.y + nay) - r1.x, 0.f);
                    const float dz  = fmaxf(fabsf(r0.z + naz) - r1.y, 0.f);
                    const float d2  = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
                    in              = __funnelshift_l(__float_as_uint(d2 - lo), in, 1);
                    out             = __funnelshift_l(__float_as_uint(hi - d2), out, 1);
                }
                bits          = __brev(in) >> 24;
                uint32_t band = ~(bits | (__brev(out) >> 24)) & 0xffu;
                if (exactLane) bits = 0, band = 0xffu;
                if (!mine) bits = 0, band = 0;
                if (__any_sync(full, band != 0))
                {
                    while (band)
                    {
                        const int o = __ffs(int(band)) - 1;
                        band &= band - 1u;
                        if (exactNodeSlow<T, AnyPbc>(p, fc + o, i)) bits |= 1u << o;
                    }
                }
            }
            else
            {
#pragma unroll
                for (int o = 0; o < 8; ++o)
                {
                    const float* r = sm.rec + o * 8; // raw float centre / size: the exact expression on staged data
                    T dx = T(r[0]) - tg.x, dy = T(r[1]) - tg.y, dz = T(r[2]) - tg.z;
                    if (AnyPbc && tg.usePbc)
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
                    bits |= (dx * dx + (dy * dy + dz * dz) < tg.cellRadiusSq) ? (1u << o) : 0u;
                }
                if (!mine) bits = 0;
            }
            return bits;
        };

        if constexpr (Async)
        {
            if (rootMask)
            {
                const int2 rootInf = p.nodeInfo[0];
                if (rootInf.y != -1)
                {
                    // single-node tree: decide everything exactly (infinite error terms put every pair in the band)
                    if (lane == 0)
                    {
                        sm.cinfo[0] = rootInf;
                        sm.rec[6] = sm.rec[7] = __int_as_float(0x7f800000);
                    }
                    __syncwarp();
                    unsigned lt = 1u, run = 0;
                    int o       = -1;
                    uint32_t base = 0, end = 0;
                    const bool have = advanceChunk(lt, o, run, base, end);
                    if (have) issueChunk(base, end);
                    cpAsyncCommit();
                    cpAsyncCommit(); // (empty) stands in for the node group
                    processLeaves(lt, o, run, base, end, have, (rootMask >> lane) & 1u);
                }
                else
                {
                    if (lane == 0)
                    {
                        sm.stackFc[0]   = 0; // sentinel: no internal node has first child 0
                        sm.stackMask[0] = 0;
                    }
                    int sp        = 1;
                    int fc        = rootInf.x;
                    unsigned mask = rootMask;
                    issueNode(fc);
                    cpAsyncCommit();
                    do
                    {
                        cpAsyncWait<0>();
                        __syncwarp();
                        const unsigned leafBits = stageChildren(fc);
                        const uint32_t bits     = overlapChildren(fc, (mask >> lane) & 1u);
                        const unsigned todo     = __reduce_or_sync(full, bits);
                        // first chunk of the reached leaf children: on its way before anything else
                        unsigned lt = todo & leafBits, run = 0;
                        int o       = -1;
                        uint32_t base = 0, end = 0;
                        const bool have = advanceChunk(lt, o, run, base, end);
                        if (have) issueChunk(base, end);
                        cpAsyncCommit();
                        // reached internal children go onto the stack (the leaves of this node are visited first)
                        unsigned pushTodo = todo & ~leafBits;
                        while (pushTodo)
                        {
                            const int q = __ffs(int(pushTodo)) - 1;
                            pushTodo &= pushTodo - 1u;
                            const unsigned cmask = __ballot_sync(full, (bits >> q) & 1u);
                            if (lane == 0)
                            {
                                sm.stackFc[sp]   = sm.cinfo[q].x;
                                sm.stackMask[sp] = cmask;
                            }
                            ++sp;
                        }
                        __syncwarp();
                        // the node that will be popped next is known now: fetch its records while the leaves are tested
                        --sp;
                        fc = sm.stackFc[sp];
                        if (fc != 0) issueNode(fc);
                        cpAsyncCommit();
                        processLeaves(lt, o, run, base, end, have, bits);
                        mask = sm.stackMask[sp];
                        __syncwarp();
                    } while (fc != 0);
                }
            }
        }
        else
        {
        if (rootMask)
        {
            const int rootFc = p.childOffsets[0];
            if (rootFc == 0)
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
                    sm.stackFc[0]   = 0; // sentinel: no internal node has first child 0
                    sm.stackMask[0] = 0;
                }
                int sp        = 1;
                int fc        = rootFc;
                unsigned mask = rootMask;
                do
                {
                    const unsigned leafBits = stageChildren(fc);
                    const uint32_t bits     = overlapChildren(fc, (mask >> lane) & 1u);
                    unsigned todo           = __reduce_or_sync(full, bits);
                    while (todo)
                    {
                        const int o = __ffs(int(todo)) - 1;
                        todo &= todo - 1u;
                        const unsigned cmask = __ballot_sync(full, (bits >> o) & 1u);
                        const int2 r         = sm.cinfo[o];
                        if ((leafBits >> o) & 1u)
                        {
                            processLeaf(uint32_t(r.x), uint32_t(r.y), cmask, sm.rec[o * 8 + 6], sm.rec[o * 8 + 7]);
                        }
                        else
                        {
                            if (lane == 0)
                            {
                                sm.stackFc[sp]   = r.x;
                                sm.stackMask[sp] = cmask;
                            }
                            ++sp;
                        }
                    }
                    __syncwarp();
                    --sp;
                    fc   = sm.stackFc[sp];
                    mask = sm.stackMask[sp];
                    __syncwarp();
                } while (fc != 0);
            }
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
                  uint32_t* neighbors, uint32_t* neighborsCount, int numLeafNodes, void* temp, void* stream)
{
    if (lastId <= firstId) return 0;
    if (!x || !y || !z || !h || !box || !childOffsets || !internalToLeaf || !layout || !centers || !sizes ||
        !neighborsCount || (ngmax && !neighbors) || (temp && numLeafNodes < 1))
        return CSTONE_E_INVALID_ARGUMENT;
    NbParams<T> p{x,       y,     z,         h,              childOffsets, internalToLeaf, layout, static_cast<const int2*>(temp),
                  centers, sizes, neighbors, neighborsCount, firstId,      lastId,         ngmax,  ext,
                  makeDBox<T>(box)};
    bool anyPbc       = p.box.pbc[0] || p.box.pbc[1] || p.box.pbc[2];
    uint32_t groups   = (lastId - firstId + 31u) / 32u;
    unsigned blocks   = gridFor(groups, kNbWarps, NB_MINBLOCKS);
    cudaStream_t s    = asStream(stream);
    if (temp)
    {
        static const bool carved = []()
        {
            // 7 blocks of 31 KB per SM: ask for the largest shared-memory carve-out
            cudaFuncSetAttribute(findNeighborsKernel<T, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
            cudaFuncSetAttribute(findNeighborsKernel<T, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
            return true;
        }();
        (void)carved;
        const int numNodes = numLeafNodes + (numLeafNodes - 1) / 7;
        nodeInfoKernel<<<gridFor(uint32_t(numNodes), 256, 8), 256, 0, s>>>(childOffsets, internalToLeaf, layout, numNodes,
                                                                           static_cast<int2*>(temp));
        CSB_LAUNCHED();
        if (anyPbc) findNeighborsKernel<T, true, true><<<blocks, kNbThreads, 0, s>>>(p);
        else findNeighborsKernel<T, false, true><<<blocks, kNbThreads, 0, s>>>(p);
    }
    else
    {
        if (anyPbc) findNeighborsKernel<T, true, false><<<blocks, kNbThreads, 0, s>>>(p);
        else findNeighborsKernel<T, false, false><<<blocks, kNbThreads, 0, s>>>(p);
    }
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
                                     sizes, ext, ngmax, neighbors, neighborsCount, 0, nullptr, stream);
}
int cstone_b200_find_neighbors_f64(const double* x, const double* y, const double* z, const double* h,
                                   uint32_t firstId, uint32_t lastId, const cstone_box_f64* box,
                                   const int32_t* childOffsets, const int32_t* internalToLeaf, const uint32_t* layout,
                                   const double* centers, const double* sizes, float ext, uint32_t ngmax,
                                   uint32_t* neighbors, uint32_t* neighborsCount, void* stream)
{
    return csb::findNeighbors<double>(x, y, z, h, firstId, lastId, box, childOffsets, internalToLeaf, layout, centers,
                                      sizes, ext, ngmax, neighbors, neighborsCount, 0, nullptr, stream);
}

uint64_t cstone_b200_find_neighbors_temp_bytes(int32_t numLeafNodes)
{
    if (numLeafNodes < 1) return 0;
    return 8ull * (uint64_t(numLeafNodes) + uint64_t(numLeafNodes - 1) / 7);
}
int cstone_b200_find_neighbors_ws_f32(const float* x, const float* y, const float* z, const float* h, uint32_t firstId,
                                      uint32_t lastId, const cstone_box_f32* box, const int32_t* childOffsets,
                                      const int32_t* internalToLeaf, const uint32_t* layout, const float* centers,
                                      const float* sizes, float ext, uint32_t ngmax, uint32_t* neighbors,
                                      uint32_t* neighborsCount, int32_t numLeafNodes, void* temp, void* stream)
{
    if (!temp) return CSTONE_E_INVALID_ARGUMENT;
    return csb::findNeighbors<float>(x, y, z, h, firstId, lastId, box, childOffsets, internalToLeaf, layout, centers,
                                     sizes, ext, ngmax, neighbors, neighborsCount, numLeafNodes, temp, stream);
}
int cstone_b200_find_neighbors_ws_f64(const double* x, const double* y, const double* z, const double* h,
                                      uint32_t firstId, uint32_t lastId, const cstone_box_f64* box,
                                      const int32_t* childOffsets, const int32_t* internalToLeaf, const uint32_t* layout,
                                      const double* centers, const double* sizes, float ext, uint32_t ngmax,
                                      uint32_t* neighbors, uint32_t* neighborsCount, int32_t numLeafNodes, void* temp,
                                      void* stream)
{
    if (!temp) return CSTONE_E_INVALID_ARGUMENT;
    return csb::findNeighbors<double>(x, y, z, h, firstId, lastId, box, childOffsets, internalToLeaf, layout, centers,
                                      sizes, ext, ngmax, neighbors, neighborsCount, numLeafNodes, temp, stream);
}

} // extern "C"
