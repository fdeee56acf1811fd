// Shared helpers for the lazzaro_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/lazzaro_b200.h"

typedef unsigned long long u64;
typedef unsigned int u32;

// ---- host-side error reporting ---------------------------------------------------------------
void lz_set_error(const char* fmt, ...);

#define LZ_CUDA(call)                                                                         \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            lz_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__,   \
                         __LINE__);                                                           \
            return LZ_ERR_CUDA;                                                               \
        }                                                                                     \
    } while (0)

#define LZ_REQUIRE(cond, ...)                                                                 \
    do {                                                                                      \
        if (!(cond)) {                                                                        \
            lz_set_error(__VA_ARGS__);                                                        \
            return LZ_ERR_ARG;                                                                \
        }                                                                                     \
    } while (0)

int lz_sm_count();   // SM count of the current device (cached per device), <= 0 on failure

// lz_search_exact with an optional device-side work list: int32 indices of the queries to answer (the
// others' outputs are left untouched) and their count.  Used by the batched search to re-run only the
// queries that failed certification; an empty list costs one small launch.
int lz_search_exact_impl(const float* rows, int64_t ld, int dim, const double* norms, const int32_t* tags,
                         int64_t row_lo, int64_t row_hi, const int64_t* group_lo, const int64_t* group_hi,
                         const int32_t* group_tag, const float* queries, int64_t q_ld, int n_queries,
                         int group_size, int k, int metric, int n_chunks, int64_t id_base, void* scratch,
                         size_t scratch_bytes, int64_t* out_ids, double* out_scores, int32_t* out_counts,
                         const int32_t* work_list, const int32_t* work_count, void* stream, int n_ranges = 1);

// ---- device helpers ----------------------------------------------------------------------------

// Order-preserving map double -> u64 (larger double = larger key). NaN must be filtered before.
__device__ __forceinline__ u64 lz_ord(double v) {
    u64 b = (u64)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double lz_unord(u64 k) {
    u64 b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}

// (key, id) total order used everywhere: better = larger key, then smaller id.
__device__ __forceinline__ bool lz_better(u64 ka, u32 ia, u64 kb, u32 ib) {
    return ka > kb || (ka == kb && ia < ib);
}

// streaming 128-bit load: read-only path, do not allocate in L1 (rows are read exactly once)
__device__ __forceinline__ float4 lz_ldg_stream(const float* p) {
    float4 r;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ int4 lz_ldg_stream_i4(const void* p) {
    int4 r;
    asm("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ double lz_warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__host__ __device__ __forceinline__ int lz_next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}
