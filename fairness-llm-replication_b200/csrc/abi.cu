// Library-wide pieces of the C ABI: version, thread-local error text, device query.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace flr {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int device_sm_count() {
    static thread_local int cached_dev = -1, cached = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached = n;
        cached_dev = dev;
    }
    return cached;
}

}  // namespace flr

extern "C" int flr_version(void) { return 100; }

extern "C" const char *flr_last_error(void) { return flr::g_err; }

extern "C" int flr_device_info(int *sm_count, int *cc_major, int *cc_minor) {
    int dev = 0;
    FLR_CUDA(cudaGetDevice(&dev));
    int sm = 0, maj = 0, min = 0;
    FLR_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    FLR_CUDA(cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, dev));
    FLR_CUDA(cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, dev));
    if (sm_count) *sm_count = sm;
    if (cc_major) *cc_major = maj;
    if (cc_minor) *cc_minor = min;
    return FLR_OK;
}
