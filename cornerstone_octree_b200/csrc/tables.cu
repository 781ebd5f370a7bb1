/*
 * tables.cu — Hilbert state-machine tables + library-wide bookkeeping (version, status strings, launch count).
 *
 * The reference computes Hilbert keys with a per-level loop of bit flips and coordinate rotations
 * [ref: sfc/hilbert.hpp:58-109].  After processing a level, the remaining lower bits of (x,y,z) have been
 * complemented per axis and the axes permuted; the accumulated transform is a signed axis permutation.
 * Exploring all transforms reachable from the identity yields 24 states.  We derive the single-level
 * transition function from the reference's level update and compose it three levels deep into lookup tables,
 * so that a 63-bit key costs 7 table look-ups instead of 21 dependent bit-twiddling iterations.
 */
#include <mutex>
#include <vector>
#include <array>
#include <cstring>

#include "common.cuh"

namespace csb
{

std::atomic<uint64_t> g_launchCount{0};

namespace
{

struct Machine
{
    // single-level transitions: digit[s][octant], next[s][octant]
    uint8_t digit[kHilbertStates][8];
    uint8_t next[kHilbertStates][8];
    std::vector<uint16_t> enc, dec;
};

//! what one level of the reference loop does to the REMAINING bits, as a map on 3-bit vectors (x<<2|y<<1|z)
unsigned levelTransform(unsigned octant, unsigned v)
{
    unsigned xi = (octant >> 2) & 1u, yi = (octant >> 1) & 1u, zi = octant & 1u;
    unsigned x = (v >> 2) & 1u, y = (v >> 1) & 1u, z = v & 1u;
    x ^= xi & ((!yi) | zi);
    y ^= (xi & (yi | zi)) | (yi & (!zi));
    z ^= (xi & (!yi) & (!zi)) | (yi & (!zi));
    if (zi)
    {
        unsigned t = x;
        x = y, y = z, z = t;
    }
    else if (!yi)
    {
        unsigned t = x;
        x = z, z = t;
    }
    return (x << 2) | (y << 1) | z;
}

Machine buildMachine()
{
    static const unsigned mortonToHilbert[8] = {0, 1, 3, 2, 7, 6, 4, 5};
    Machine m;
    typedef std::array<uint8_t, 8> Transform; // image of every 3-bit vector
    std::vector<Transform> states;
    Transform ident;
    for (unsigned v = 0; v < 8; ++v)
        ident[v] = uint8_t(v);
    states.push_back(ident);
    for (size_t s = 0; s < states.size(); ++s)
    {
        for (unsigned o = 0; o < 8; ++o)
        {
            unsigned seen = states[s][o]; // octant as the reference loop sees it at this level
            Transform t;
            for (unsigned v = 0; v < 8; ++v)
                t[v] = uint8_t(levelTransform(seen, states[s][v]));
            size_t idx = 0;
            while (idx < states.size() && states[idx] != t)
                ++idx;
            if (idx == states.size()) states.push_back(t);
            m.digit[s][o] = uint8_t(mortonToHilbert[seen]);
            m.next[s][o]  = uint8_t(idx);
        }
    }
    if (states.size() != size_t(kHilbertStates)) std::abort();

    m.enc.assign(kHilbertTabSize, 0);
    m.dec.assign(kHilbertTabSize, 0);
    for (unsigned s = 0; s < unsigned(kHilbertStates); ++s)
    {
        for (unsigned idx = 0; idx < 512; ++idx)
        {
            unsigned x3 = idx >> 6, y3 = (idx >> 3) & 7u, z3 = idx & 7u;
            unsigned st = s, digits = 0;
            for (int l = 2; l >= 0; --l)
            {
                unsigned o = (((x3 >> l) & 1u) << 2) | (((y3 >> l) & 1u) << 1) | ((z3 >> l) & 1u);
                digits     = (digits << 3) | m.digit[st][o];
                st         = m.next[st][o];
            }
            m.enc[s * 512 + idx]    = uint16_t(digits | (st << 9));
            m.dec[s * 512 + digits] = uint16_t(idx | (st << 9));
        }
    }
    return m;
}

const Machine& machine()
{
    static const Machine m = buildMachine();
    return m;
}

constexpr int kMaxDevices = 64;
std::mutex g_tabMutex;
uint16_t* g_devEnc[kMaxDevices] = {};
uint16_t* g_devDec[kMaxDevices] = {};

} // namespace

const uint16_t* hostEncTable() { return machine().enc.data(); }
const uint16_t* hostDecTable() { return machine().dec.data(); }

int getSfcTables(SfcTables* out)
{
    int dev = 0;
    CSB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return CSTONE_E_INVALID_ARGUMENT;
    std::lock_guard<std::mutex> lock(g_tabMutex);
    if (!g_devEnc[dev])
    {
        const Machine& m = machine();
        uint16_t* e      = nullptr;
        uint16_t* d      = nullptr;
        CSB_CUDA(cudaMalloc(&e, kHilbertTabBytes));
        CSB_CUDA(cudaMalloc(&d, kHilbertTabBytes));
        CSB_CUDA(cudaMemcpy(e, m.enc.data(), kHilbertTabBytes, cudaMemcpyHostToDevice));
        CSB_CUDA(cudaMemcpy(d, m.dec.data(), kHilbertTabBytes, cudaMemcpyHostToDevice));
        g_devEnc[dev] = e;
        g_devDec[dev] = d;
    }
    out->enc = g_devEnc[dev];
    out->dec = g_devDec[dev];
    return 0;
}

} // namespace csb

extern "C" {

int cstone_b200_abi_version(void) { return CSTONE_B200_ABI_VERSION; }

uint64_t cstone_b200_launch_count(void) { return csb::g_launchCount.load(); }

int cstone_b200_device_malloc(void** ptr, size_t bytes)
{
    if (!ptr) return CSTONE_E_INVALID_ARGUMENT;
    *ptr = nullptr;
    if (bytes == 0) return 0;
    return int(cudaMalloc(ptr, bytes));
}
int cstone_b200_device_free(void* ptr) { return ptr ? int(cudaFree(ptr)) : 0; }
int cstone_b200_device_memcpy(void* dst, const void* src, size_t bytes, int kind)
{
    if (bytes == 0) return 0;
    if (!dst || !src || kind < 0 || kind > 2) return CSTONE_E_INVALID_ARGUMENT;
    const cudaMemcpyKind k[3] = {cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice};
    return int(cudaMemcpy(dst, src, bytes, k[kind]));
}
int cstone_b200_device_memset(void* dst, int value, size_t bytes)
{
    if (bytes == 0) return 0;
    if (!dst) return CSTONE_E_INVALID_ARGUMENT;
    return int(cudaMemset(dst, value, bytes));
}
int cstone_b200_device_synchronize(void) { return int(cudaDeviceSynchronize()); }
int cstone_b200_set_device(int device) { return int(cudaSetDevice(device)); }

const char* cstone_b200_status_string(int status)
{
    switch (status)
    {
        case 0: return "success";
        case CSTONE_E_INVALID_ARGUMENT: return "cstone_b200: invalid argument";
        case CSTONE_E_TEMP_TOO_SMALL: return "cstone_b200: temporary storage too small";
        case CSTONE_E_TOO_LARGE: return "cstone_b200: problem size exceeds the supported range";
        case CSTONE_E_STACK_OVERFLOW: return "cstone_b200: traversal stack exhausted";
        case CSTONE_E_CAPACITY: return "cstone_b200: output capacity too small";
        case -6: return "cstone_b200: NCCL error";
        case -7: return "cstone_b200: array size differs from the previous sync";
        case -8: return "cstone_b200: focus tree / halo bookkeeping of the ranks is inconsistent (reference halo mode)";
        case -9: return "cstone_b200: a peer's halo data did not arrive (pack-and-push exchange)";
        default: break;
    }
    if (status > 0) return cudaGetErrorString(cudaError_t(status));
    return "cstone_b200: unknown status";
}

/* host-side table access for the table-generator unit test (no GPU needed) */
const uint16_t* cstone_b200_host_hilbert_table(int decode)
{
    return decode ? csb::hostDecTable() : csb::hostEncTable();
}

} // extern "C"
