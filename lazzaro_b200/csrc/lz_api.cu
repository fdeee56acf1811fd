// Library-level entry points: version, per-thread error text, device check.
#include "lz_common.cuh"

#include <mutex>

namespace {
thread_local char g_err[512] = "";
std::mutex g_dev_mu;
int g_sm_count[64] = {0};      // per device ordinal; 0 = not probed, -1 = not an sm_100 device
}  // namespace

void lz_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int lz_sm_count() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        lz_set_error("no usable CUDA device (cudaGetDevice failed) — this library has no CPU fallback");
        return -1;
    }
    std::lock_guard<std::mutex> lock(g_dev_mu);
    if (g_sm_count[dev] == 0) {
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
            lz_set_error("cudaGetDeviceProperties failed for device %d", dev);
            return -1;
        }
        g_sm_count[dev] = (prop.major == 10) ? prop.multiProcessorCount : -1;
        if (g_sm_count[dev] < 0)
            lz_set_error("device %d is sm_%d%d; this library is built for sm_100a only", dev, prop.major, prop.minor);
    }
    return g_sm_count[dev];
}

extern "C" int lz_version(void) { return 100; }   // 0.1.0

extern "C" const char* lz_last_error(void) { return g_err; }

extern "C" int lz_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0;
    LZ_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    LZ_CUDA(cudaGetDeviceProperties(&prop, dev));
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    if (prop.major != 10) {
        lz_set_error("device %d is sm_%d%d; this library is built for sm_100a only", dev, prop.major, prop.minor);
        return LZ_ERR_DEVICE;
    }
    return LZ_OK;
}
