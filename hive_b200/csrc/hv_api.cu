// C-ABI plumbing: version, thread-local error text, device info cache.
#include "hv_common.cuh"
#include <string.h>
#include <mutex>

namespace hv {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return HV_E_CUDA;
}

const DeviceInfo& device_info() {
    static DeviceInfo cache[64];
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    std::lock_guard<std::mutex> lock(mu);
    DeviceInfo& d = cache[dev];
    if (d.device != dev) {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0) {
            d.sm_count = sms;
            d.device = dev;
        } else {
            d.sm_count = 148;  // B200
        }
    }
    return d;
}

}  // namespace hv

extern "C" int hv_version(void) { return HV_VERSION; }

#ifndef HV_BUILD_ID
#define HV_BUILD_ID "unknown"
#endif
extern "C" int hv_build_id(char* buf, int n) {
    static const char id[] = HV_BUILD_ID;
    const int len = (int)sizeof(id) - 1;
    if (buf && n > 0) {
        const int c = len < n - 1 ? len : n - 1;
        memcpy(buf, id, c);
        buf[c] = 0;
    }
    return len;
}

extern "C" int hv_last_error(char* buf, int n) {
    int len = (int)strlen(hv::g_err);
    if (buf && n > 0) {
        int c = len < n - 1 ? len : n - 1;
        memcpy(buf, hv::g_err, c);
        buf[c] = 0;
    }
    return len;
}

extern "C" int hv_sm_count(void) {
    int dev = 0;
    HV_CUDA(cudaGetDevice(&dev));
    return hv::device_info().sm_count;
}
