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


// ---- the exchange step of the multi-GPU search behind the C-ABI ----------------------------------------
// ncclAllGather of every rank's packed local top-k block, then the strided merge, on the caller's stream.  NCCL is
// resolved at first use with dlopen/dlsym (no link-time dependency: a binding that never calls this needs no NCCL at
// all), from the library image the process already has — the caller's ncclComm_t must come from that same library.
#include <dlfcn.h>

namespace {
typedef int (*NcclAllGatherFn)(const void*, void*, size_t, int /*ncclDataType_t*/, void* /*ncclComm_t*/, cudaStream_t);
typedef const char* (*NcclErrFn)(int);
std::mutex g_nccl_mu;
NcclAllGatherFn g_allgather = nullptr;
NcclErrFn g_nccl_err = nullptr;

int resolve_nccl() {
    std::lock_guard<std::mutex> lock(g_nccl_mu);
    if (g_allgather) return LZ_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);              // the copy already mapped (e.g. torch's), if any
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        lz_set_error("lz_allgather_merge: libnccl.so.2 could not be loaded (%s)", dlerror());
        return LZ_ERR_CUDA;
    }
    g_allgather = (NcclAllGatherFn)dlsym(h, "ncclAllGather");
    g_nccl_err = (NcclErrFn)dlsym(h, "ncclGetErrorString");
    if (!g_allgather) {
        lz_set_error("lz_allgather_merge: ncclAllGather not found in libnccl");
        return LZ_ERR_CUDA;
    }
    return LZ_OK;
}
}  // namespace

extern "C" int lz_allgather_merge(void* nccl_comm, int world, const int64_t* local_block, int64_t* gathered, int n_queries,
                                  int k, int metric, int64_t* out_ids, double* out_scores, int32_t* out_counts,
                                  void* stream) {
    LZ_REQUIRE(nccl_comm && local_block && gathered && out_ids && out_scores && out_counts, "lz_allgather_merge: NULL pointer");
    LZ_REQUIRE(world >= 1 && n_queries >= 1 && k >= 1 && k <= LZ_MAX_K, "lz_allgather_merge: world=%d n_queries=%d k=%d", world, n_queries, k);
    if (int rc = resolve_nccl()) return rc;
    const size_t count = (size_t)2 * n_queries * k;                           // int64 words per rank: [ids | score bits]
    const int nrc = g_allgather(local_block, gathered, count, 4 /* ncclInt64 */, nccl_comm, (cudaStream_t)stream);
    if (nrc != 0) {
        lz_set_error("lz_allgather_merge: ncclAllGather failed: %s", g_nccl_err ? g_nccl_err(nrc) : "?");
        return LZ_ERR_CUDA;
    }
    return lz_merge_topk_strided(gathered, reinterpret_cast<const double*>(gathered + (size_t)n_queries * k), (int64_t)count, world,
                                 n_queries, k, metric, out_ids, out_scores, out_counts, stream);
}
