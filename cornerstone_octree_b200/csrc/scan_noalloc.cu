/*
 * scan_noalloc.cu — exclusive u32 scan for callers that must not allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
 * reference.
 *
 * Same kernel as cstone_b200_exclusive_scan_u32 (scan.cuh: single-pass decoupled look-back); the only difference is
 * where the scratch (one ticket counter + one status word per tile) lives: in the file-scope device array below
 * instead of a caller-provided buffer.  Consequences, all documented in include/cstone_b200.h:
 *  - the array is sized for CSTONE_B200_SCAN_NOALLOC_MAX_N elements (8192 tiles -> 65 792 bytes); every GPU that
 *    loads the library holds it until the process exits; larger n is rejected with CSTONE_E_TOO_LARGE;
 *  - every call on a device shares the array, and it keeps the previous call's contents: the part this call uses is
 *    cleared (cudaMemsetAsync on the caller's stream) before the kernel; two calls that OVERLAP on one GPU (different
 *    streams without an ordering between them) race on it;
 *  - no cudaMalloc / cudaFree / allocating library routine is called: cudaGetSymbolAddress, cudaMemsetAsync and the
 *    kernel launch only.
 */
#include <mutex>

#include "scan.cuh"

namespace csb
{

constexpr size_t kNoAllocMaxN         = CSTONE_B200_SCAN_NOALLOC_MAX_N;
constexpr size_t kNoAllocScratchBytes = (128 + 8 * ((kNoAllocMaxN + kScanTile - 1) / kScanTile) + 8 + 255) / 256 * 256;

// In host code this name is a host-side stand-in, NOT a device address: the kernel gets cudaGetSymbolAddress's result.
static __device__ __align__(256) unsigned char g_scanNoAllocScratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;

//! device address of the scratch array on the CURRENT device (looked up once per device); status != 0 on failure
static int noAllocScratch(void** out)
{
    static std::mutex mtx;
    static void* addr[kMaxDevices] = {nullptr};
    int dev                        = 0;
    CSB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return CSTONE_E_INVALID_ARGUMENT;
    std::lock_guard<std::mutex> lock(mtx);
    if (!addr[dev])
    {
        void* p = nullptr;
        CSB_CUDA(cudaGetSymbolAddress(&p, g_scanNoAllocScratch));
        if (!p) return int(cudaErrorInvalidSymbol);
        addr[dev] = p;
    }
    *out = addr[dev];
    return 0;
}

} // namespace csb

extern "C" int cstone_b200_exclusive_scan_u32_noalloc(const uint32_t* in, uint32_t* out, size_t n, uint32_t init,
                                                      void* stream)
{
    if (n > csb::kNoAllocMaxN) return CSTONE_E_TOO_LARGE;
    if (n && (!in || !out)) return CSTONE_E_INVALID_ARGUMENT;
    if (n == 0) return 0;
    static_assert(csb::kNoAllocScratchBytes >= 128 + 8 * ((csb::kNoAllocMaxN + csb::kScanTile - 1) / csb::kScanTile) + 8,
                  "scratch array too small for the stated maximum");
    void* scratch = nullptr;
    int st        = csb::noAllocScratch(&scratch);
    if (st) return st;
    // launchScan clears the ticket and the status words of this call's tiles (cudaMemsetAsync), then launches
    return csb::launchScan<false>(csb::LoadPlain{in}, out, n, init, scratch, nullptr, csb::asStream(stream));
}
