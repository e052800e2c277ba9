// Shared helpers of the hive_b200 CUDA library (error plumbing, launch geometry, small device utilities).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/hive_b200.h"

namespace hv {

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

struct DeviceInfo {
    int device = -1;
    int sm_count = 0;
};
// Cached per device (queried once per device per process).
const DeviceInfo& device_info();

#define HV_CUDA(expr)                                            \
    do {                                                         \
        cudaError_t _e = (expr);                                 \
        if (_e != cudaSuccess) return hv::cuda_fail(_e, #expr);  \
    } while (0)

#define HV_REQUIRE(cond, ...)           \
    do {                                \
        if (!(cond)) {                  \
            hv::set_error(__VA_ARGS__); \
            return HV_E_INVALID;        \
        }                               \
    } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// hv_scan.cu: offs[i] = sum of counts[0..i) componentwise; totals[0..1] (device, optional) receive the grand totals.
// `bsum` is scratch of scan_scratch_pairs(n) pairs.
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;
long long scan_tiles(long long n);
long long scan_scratch_pairs(long long n);  // size of `bsum` in pairs
void exclusive_scan_u2(const uint2* counts, uint2* offs, uint2* bsum, long long n, unsigned long long* totals,
                       cudaStream_t st);

}  // namespace hv
