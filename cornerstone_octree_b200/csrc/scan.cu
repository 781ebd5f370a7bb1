/*
 * scan.cu — C ABI of the plain u32 scans  [ref: primitives/primitives_gpu.h:112-122 exclusiveScanGpu / inclusiveScanGpu]
 */
#include "scan.cuh"

extern "C" {

uint64_t cstone_b200_scan_temp_bytes(uint64_t n) { return csb::scanTempBytes(n); }

int cstone_b200_exclusive_scan_u32(const uint32_t* in, uint32_t* out, size_t n, uint32_t init, void* temp,
                                   void* stream)
{
    if (n && (!in || !out || !temp)) return CSTONE_E_INVALID_ARGUMENT;
    return csb::launchScan<false>(csb::LoadPlain{in}, out, n, init, temp, nullptr, csb::asStream(stream));
}

int cstone_b200_inclusive_scan_u32(const uint32_t* in, uint32_t* out, size_t n, void* temp, void* stream)
{
    if (n && (!in || !out || !temp)) return CSTONE_E_INVALID_ARGUMENT;
    return csb::launchScan<true>(csb::LoadPlain{in}, out, n, 0u, temp, nullptr, csb::asStream(stream));
}

} // extern "C"
