// K2: batched cosine search as a tensor-core contraction with a fused candidate epilogue.
//
//   coarse pass   scores[q, row] = fp16(q/|q|) . fp16(x/|x|), fp32 accumulate in TMEM: a persistent,
//                 warp-specialised tcgen05 GEMM (TMA 128B-swizzled tiles -> shared memory ring ->
//                 tcgen05.mma 128x256x16 -> double-buffered TMEM accumulators).  The B x N score matrix
//                 is never written: four epilogue warps read each accumulator tile with tcgen05.ld
//                 (one query per thread, i.e. per TMEM lane) and keep, per query, the best KP coarse
//                 candidates of the CTA's slice of the rows (threshold compare, append, warp-cooperative
//                 bitonic compaction).
//   exact pass    per query: merge the slices' candidates by coarse score, rescore the best KP in
//                 float64 from the float32 rows (same summation order as the exact scan, so certified
//                 results are bit-identical to it), select the top-k with the id tie-break, and CERTIFY:
//                 every row dropped anywhere had coarse score <= T, hence exact score <= T + eps; if the
//                 k-th exact score exceeds T + eps the top-k is complete.
//   fallback      queries that fail certification are re-run by the exact float64 scan (still on the
//                 GPU), so the call always returns the exact answer.
//
// Stands behind LanceDBStore.search_nodes (core/vector_store.py:132-140) for query batches, and behind
// the dedup / linking scans of core/memory_system.py:719-742, 838-891 (same contraction, small k).
#include <cuda_fp16.h>
#include <stdlib.h>

#include "lz_common.cuh"
#include "lz_sm100.cuh"
#include "lz_topk.cuh"

namespace {

using namespace sm100;

constexpr int BM = 128;                 // queries per tile  (MMA M, one TMEM lane per query)
constexpr int BN = 256;                 // rows per tile     (MMA N, one TMEM column per row)
constexpr int BK = 64;                  // fp16 elements per pipeline stage = one 128-byte swizzle row
constexpr int A_BYTES = BM * BK * 2;    // 16 KiB per stage; the rows tile adds 32 KiB
constexpr int kGemmThreads = 192;       // warp 0: TMA, warp 1: MMA + TMEM owner, warps 2-5: epilogue
constexpr int kTmemCols = 512;          // two 128x256 fp32 accumulators
constexpr int kSelThreads = 256;
#ifndef LZ_SUB
#define LZ_SUB 8
#endif
constexpr int kSub = LZ_SUB;            // rows per candidate group in the coarse pass (8, 16 or 32)

__device__ __forceinline__ u32 ord32(float v) {
    u32 b = __float_as_uint(v);
    return (b >> 31) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float unord32(u32 k) {
    u32 b = (k >> 31) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(b);
}

// ---- preparation: the 16-bit shadow --------------------------------------------------------------------
// The coarse pass reads an fp16 copy of the rows ("shadow") and of the query batch.  fp16, not bf16: the tensor
// cores run both at the same rate, the operands are normalised (cosine) or of embedding scale (L2) so fp16's
// range is ample, and its 11 significant bits make the rounding error 8x smaller than bf16's 8 — small enough
// that the DETERMINISTIC certification margin below (no independence assumption) certifies ordinary data.
//
// LZ_COSINE: out[r] = fp16(x / |x|), dim columns.  LZ_L2: out[r] = [fp16(x), h, m, l, 0, 0, 0, 0], dim + 8 columns,
// where h + m + l is a three-way fp16 split of -|x|^2 / 2: against the query [fp16(q), 1, 1, 1, 0, ...] the
// contraction gives x.q - |x|^2/2 = -(|x - q|^2 - |q|^2) / 2, i.e. the L2 ranking comes out of the same GEMM,
// largest = nearest.
//
// row_err (optional) receives, per row, an upper bound of the 2-norm of the rounding error of its shadow:
// |shadow - x/|x|| for LZ_COSINE (the normalisation's own fp32 rounding included), |fp16(x) - x| for LZ_L2 —
// +inf when a component or the norm term leaves fp16's range (such a row certifies nothing: every query then
// takes the exact fallback, slow but right).  The largest of them over the scanned rows is the row-side term of
// the deterministic margin.
__global__ void __launch_bounds__(256) rows_to_shadow_kernel(const float* __restrict__ rows, int64_t n, int dim, int64_t ld,
                                                             const double* __restrict__ norms, int metric,
                                                             __half* __restrict__ out, float* __restrict__ row_err) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * 8;
    const int odim = metric == LZ_L2 ? dim + 8 : dim;
    for (int64_t r = warp; r < n; r += nwarps) {
        const double nr = norms[r];
        const float inv = metric == LZ_L2 ? 1.f : (nr > 0.0 ? (float)(1.0 / nr) : 0.f);
        const double inv_d = metric == LZ_L2 ? 1.0 : (nr > 0.0 ? 1.0 / nr : 0.0);
        const float* x = rows + r * ld;
        __half* o = out + r * odim;
        double err2 = 0.0;
        for (int e = lane * 4; e < dim; e += 128) {
            const float4 v = lz_ldg_stream(x + e);
            __half2 a = __floats2half2_rn(v.x * inv, v.y * inv);
            __half2 b = __floats2half2_rn(v.z * inv, v.w * inv);
            uint2 pk;
            pk.x = *reinterpret_cast<u32*>(&a);
            pk.y = *reinterpret_cast<u32*>(&b);
            *reinterpret_cast<uint2*>(o + e) = pk;
            if (row_err != nullptr) {
                const double d0 = (double)__low2float(a) - (double)v.x * inv_d, d1 = (double)__high2float(a) - (double)v.y * inv_d;
                const double d2 = (double)__low2float(b) - (double)v.z * inv_d, d3 = (double)__high2float(b) - (double)v.w * inv_d;
                err2 = fma(d0, d0, err2); err2 = fma(d1, d1, err2); err2 = fma(d2, d2, err2); err2 = fma(d3, d3, err2);
            }
        }
        bool split_ok = true;
        if (metric == LZ_L2) {
            const double v = -0.5 * nr * nr;
            const __half h = __double2half(v);
            const __half m = __double2half(v - (double)__half2float(h));
            const __half l = __double2half(v - (double)__half2float(h) - (double)__half2float(m));
            if (lane < 8) o[dim + lane] = lane == 0 ? h : lane == 1 ? m : lane == 2 ? l : __float2half_rn(0.f);
            const double back = (double)__half2float(h) + (double)__half2float(m) + (double)__half2float(l);
            split_ok = fabs(back - v) <= 1e-6 * fabs(v) + 1e-12;          // false for inf / NaN too
        }
        if (row_err != nullptr) {
            err2 = lz_warp_sum(err2);                                       // inf / NaN if a component overflowed
            const double e = sqrt(err2) * (1.0 + 1e-9);
            if (lane == 0) row_err[r] = (split_ok && e == e) ? __double2float_ru(e) : INFINITY;
        }
    }
}

// one warp per query: fp16 copy (normalised for cosine) + the certification margin eps[q]
__global__ void __launch_bounds__(256) queries_to_shadow_kernel(const float* __restrict__ q, int n, int dim, int64_t q_ld,
                                                                __half* __restrict__ out, double* __restrict__ eps,
                                                                int32_t* __restrict__ gtag, u32* __restrict__ gthr, int32_t* __restrict__ fb_count,
                                                                int want_tag, double sigmas, int metric, double max_norm,
                                                                double det_row_err) {
    const int lane = threadIdx.x & 31;
    const int w = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (blockIdx.x == 0 && threadIdx.x == 0) *fb_count = 0;
    if (w >= n) return;
    const float* v = q + (int64_t)w * q_ld;
    double ss = 0.0;
    float mx = 0.f;
    for (int e = lane; e < dim; e += 32) {
        const float x = v[e];
        ss = fma((double)x, (double)x, ss);
        mx = fmaxf(mx, fabsf(x));
    }
    ss = lz_warp_sum(ss);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const double nr = sqrt(ss);
    // Two margins (u = 2^-11, fp16's unit roundoff).
    // Statistical (det_row_err < 0): each operand carries an independent relative rounding error uniform in +-u,
    // so the coarse dot errs with std <= u*sqrt(2/3)*max|q_i|*|x|; eps = sigmas * that + an accumulation term.  It
    // holds for embeddings whose rounding errors are uncorrelated and FAILS for correlated ones (quantised or
    // near-constant vectors whose components all round the same way).
    // Deterministic (det_row_err >= 0, an upper bound of the rows' shadow error norms dx = |shadow - x^| from
    // lz_rows_to_shadow): with dq = |fp16(q^) - q^| measured here, Cauchy-Schwarz gives
    // |coarse - exact| <= dx |q^| + dq |x^| + dx dq for ANY data, plus the fp32 accumulation of dim products
    // (dim * 2^-23 * |q^||x^|: a truncating add per product assumed).
    constexpr double kU = 4.8828125e-4;                // 2^-11
    constexpr double kAcc = 1.1920928955078125e-7;     // 2^-23
    if (metric == LZ_L2) {
        __half* o = out + (int64_t)w * (dim + 8);
        double e2 = 0.0;
        for (int e = lane; e < dim; e += 32) {
            const __half b = __float2half_rn(v[e]);
            o[e] = b;
            const double d = (double)__half2float(b) - (double)v[e];
            e2 = fma(d, d, e2);
        }
        if (lane < 8) o[dim + lane] = __float2half_rn(lane < 3 ? 1.f : 0.f);
        const double dq = sqrt(lz_warp_sum(e2)) * (1.0 + 1e-9);
        if (lane == 0) {  // resid: the norm split's residue (<= 1e-6 |x|^2 / 2, checked per row) and float32 inputs' own spacing
            const double resid = 2e-6 * (max_norm * nr + max_norm * max_norm) + 1e-30;
            double ev;
            if (det_row_err >= 0.0)
                ev = det_row_err * nr + dq * max_norm + det_row_err * dq +
                     (double)(dim + 8) * kAcc * (max_norm * nr + max_norm * max_norm) + resid;
            else
                ev = sigmas * kU * 0.816496580927726 * (double)mx * max_norm + resid;
            eps[w] = ev == ev ? ev : INFINITY;
        }
    } else {
        const float inv = nr > 0.0 ? (float)(1.0 / nr) : 0.f;
        const double inv_d = nr > 0.0 ? 1.0 / nr : 0.0;
        double e2 = 0.0;
        for (int e = lane; e < dim; e += 32) {
            const __half b = __float2half_rn(v[e] * inv);
            out[(int64_t)w * dim + e] = b;
            const double d = (double)__half2float(b) - (double)v[e] * inv_d;
            e2 = fma(d, d, e2);
        }
        const double dq = sqrt(lz_warp_sum(e2)) * (1.0 + 1e-9);
        if (lane == 0) {
            double ev;
            if (det_row_err >= 0.0)
                ev = det_row_err + dq + det_row_err * dq + (double)dim * kAcc * 1.01 + 1e-9;
            else
                ev = sigmas * kU * 0.816496580927726 * (double)(mx * inv) + 8e-6;
            eps[w] = ev == ev ? ev : INFINITY;
        }
    }
    if (lane == 0) {
        gtag[w] = want_tag;
        gthr[w] = 0u;
    }
}

#ifdef LZ_STATS
__device__ unsigned long long g_stats[8];
__device__ unsigned long long g_sel_stats[4];   // exact pass: sum of merged groups, of groups kept by the prune, queries   // 0 epi busy, 1 epi max, 2 mma wait tempty, 3 mma wait full, 4 prod wait empty, 5 tiles
#define LZ_T0 long long t0_ = clock64();
#define LZ_ACC(i) atomicAdd(&g_stats[i], (unsigned long long)(clock64() - t0_));
#else
#define LZ_T0
#define LZ_ACC(i)
#endif

// ---- the coarse pass ----------------------------------------------------------------------------------
struct GemmParams {
    int n_queries, n_qtiles, n_slices, tiles_per_slice, n_tiles, kblocks;
    int tile_mul;       // rows tile t starts at row_lo + t * tile_mul * BN (1 normally; a stride in the sampling pass)
    int cs, n_qgroups;  // cluster size (CTAs that share one rows tile by TMA multicast) and q-tile groups of that size
    int64_t row_lo, row_hi;
    int kp, cap;
    const int32_t* tags;
    int want_tag;
    u64* cand;          // [n_items][BM][cap]  key = ord32(best coarse of the group) << 32 | (0xffffffff - first row)
    int32_t* cand_cnt;  // [n_items][BM]
    float* cand_thr;    // [n_items][BM]  every row dropped by this (item, query) had coarse <= thr
    u32* gthr;          // [n_queries] best threshold any CTA has reached for the query (ord32 key, 0 = none)
    const double* top1_eps;   // top-1 against a tiny table only (else null): per-query eps; the threshold then
                              // follows the running best group at a distance of 3 eps (see scan_chunk)
    // Grouped mode (two-level search; items != null): the work is a device-built list of (q-tile of the PERMUTED
    // query array, row range of one shard slice) items instead of the (q-group, slice) grid; cs == 1.
    const int4* items;        // [*n_items_dev] x = q-tile, y = first row, z = end row (of the shard slice)
    const int* n_items_dev;
    const int32_t* qmap;      // [perm rows] real query index of a permuted query row, -1 = padding
    int debug_skip;           // profiling only (LZ_DEBUG_SKIP): 1 = the epilogue drains the accumulators without reading them
    int raw_items;            // sampling pass of the grouped search: items are short enough never to overflow a candidate
                              // buffer; leave the buffers as pushed (no compaction to kp at the end of an item)
};

// What one CTA works on: a q-tile against rows tiles row_begin, row_begin + stride, ... (n_tiles of them); rows at or
// beyond row_end are not candidates; its candidate slots are slot_base + lane.
struct WorkItem {
    int qt, n_tiles;
    int64_t row_begin, row_end, stride;
    size_t slot_base;
};

__device__ __forceinline__ WorkItem decode_item(const GemmParams& p, int item, int rank) {
    WorkItem w;
    if (p.items != nullptr) {
        const int4 it = p.items[item];
        w.qt = it.x;
        w.row_begin = it.y;
        w.row_end = it.z;
        w.stride = BN;
        w.n_tiles = (it.z - it.y + BN - 1) / BN;
        w.slot_base = (size_t)item * BM;
    } else {
        const int sl = item / p.n_qgroups;
        w.qt = (item % p.n_qgroups) * p.cs + rank;
        const int t0 = sl * p.tiles_per_slice;
        const int t1 = min(p.n_tiles, t0 + p.tiles_per_slice);
        w.stride = (int64_t)p.tile_mul * BN;
        w.row_begin = p.row_lo + (int64_t)t0 * w.stride;
        w.row_end = p.row_hi;
        w.n_tiles = t1 - t0;
        w.slot_base = ((size_t)sl * p.n_qgroups * p.cs + w.qt) * BM;
    }
    return w;
}

// Warp-cooperative compaction of the candidate buffers of the lanes in `need` (each over kp entries):
// the warp loads one lane's buffer into registers (EPL keys per lane), finds the kp-th largest coarse
// score by a 32-step bisection on the order-preserving key (compare + warp reduce-add per step), and
// writes back the entries above it plus enough of the ties to make kp.  No sort: the buffer stays
// unordered, which is all the later merge needs.  Which of several rows tied AT the threshold survive
// is immaterial — the threshold itself is what the certification accounts for.
template <int EPL>
__device__ __forceinline__ void warp_select(unsigned need, u64* my_buf, int& count, float& thr, int kp, int lane) {
    const unsigned lt = (1u << lane) - 1u;
    __syncwarp();
    while (need) {
        const int L = __ffs(need) - 1;
        need &= need - 1;
        u64* src = reinterpret_cast<u64*>(__shfl_sync(0xffffffffu, (unsigned long long)my_buf, L));
        const int n = __shfl_sync(0xffffffffu, count, L);
        u64 key[EPL];
#pragma unroll
        for (int i = 0; i < EPL; ++i) key[i] = (lane + 32 * i) < n ? src[lane + 32 * i] : 0ull;
        // the kp-th largest key lies between the smallest and the largest live key: their common leading
        // bits are settled, so the bisection starts at the first bit in which they differ (scores of one
        // query span a narrow range, which saves about a third of the 32 steps)
        u32 mx = 0u, mn = 0xffffffffu;
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            const u32 hi = (u32)(key[i] >> 32);
            mx = max(mx, hi);
            mn = min(mn, hi != 0u ? hi : 0xffffffffu);
        }
        mx = __reduce_max_sync(0xffffffffu, mx);
        mn = __reduce_min_sync(0xffffffffu, mn);
        const u32 diff = mx ^ mn;
        const int top = diff ? 31 - __clz(diff) : -1;
        u32 prefix = top >= 0 ? (mx & ~((2u << top) - 1u)) : mx;
#pragma unroll 1
        for (int bit = top; bit >= 0; --bit) {
            const u32 cand = prefix | (1u << bit);
            int c = 0;
#pragma unroll
            for (int i = 0; i < EPL; ++i) c += (u32)(key[i] >> 32) >= cand;
            c = __reduce_add_sync(0xffffffffu, c);
            if (c >= kp) prefix = cand;
        }
        int gt = 0;
#pragma unroll
        for (int i = 0; i < EPL; ++i) gt += (u32)(key[i] >> 32) > prefix;
        gt = __reduce_add_sync(0xffffffffu, gt);
        const int room = kp - gt;
        __syncwarp();
        int out = 0, eq_seen = 0;
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            const u32 hi = (u32)(key[i] >> 32);
            const bool e = hi == prefix && prefix != 0;
            const unsigned me = __ballot_sync(0xffffffffu, e);
            const bool keep = hi > prefix || (e && eq_seen + __popc(me & lt) < room);
            const unsigned mk = __ballot_sync(0xffffffffu, keep);
            if (keep) src[out + __popc(mk & lt)] = key[i];
            out += __popc(mk);
            eq_seen += __popc(me);
        }
        if (lane == L) {
            count = out;
            thr = unord32(prefix);
        }
        __syncwarp();
    }
}

// One TMEM chunk (32 consecutive rows of this thread's query): reduce each group of kSub rows to its best
// coarse score and append the groups that beat the query's threshold.
template <int EPL>
__device__ __forceinline__ void scan_chunk(const uint32_t (&v)[32], int64_t row_first, int64_t row_hi, const GemmParams& p, u64* buf,
                                           int& count, float& thr, float slack, int q, int lane) {
    float g[32 / kSub];
#pragma unroll
    for (int j = 0; j < 32 / kSub; ++j) {
        float m = __uint_as_float(v[j * kSub]);
#pragma unroll
        for (int i = 1; i < kSub; ++i) m = fmaxf(m, __uint_as_float(v[j * kSub + i]));
        g[j] = m;
    }
    float m = g[0];
#pragma unroll
    for (int j = 1; j < 32 / kSub; ++j) m = fmaxf(m, g[j]);
    // top-1 mode (slack = 3 eps, else +inf): only groups within 2 eps of the best group can hold the winner, so
    // the threshold trails the running best by 3 eps — a handful of candidates per query instead of all of them,
    // and the final threshold T = c_1 - 3 eps certifies by construction (exact best >= c_1 - eps > T + eps).
    // Chunks reaching past the range are left out: their rows are not candidates.
    if (row_first + 32 <= row_hi) thr = fmaxf(thr, m - slack);
    const bool top1 = slack != INFINITY;                          // single-row candidates
    if (m > thr) {
        if (top1) {
            // row mode keeps single ROWS: the group max is tested first, so the common case (nothing passes) costs
            // the same, and the exact pass rescores the rows themselves instead of kSub rows per candidate
#pragma unroll
            for (int j = 0; j < 32 / kSub; ++j) {
                if (g[j] > thr) {
#pragma unroll
                    for (int i = 0; i < kSub; ++i) {
                        const float val = __uint_as_float(v[j * kSub + i]);
                        const int64_t r = row_first + j * kSub + i;
                        if (val > thr && r < row_hi) buf[count++] = ((u64)ord32(val) << 32) | (u64)(0xffffffffu - (u32)r);
                    }
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32 / kSub; ++j) {
                const int64_t r0 = row_first + j * kSub;
                if (g[j] > thr && r0 < row_hi) buf[count++] = ((u64)ord32(g[j]) << 32) | (u64)(0xffffffffu - (u32)r0);
            }
        }
    }
    const unsigned need = __ballot_sync(0xffffffffu, count > p.cap - (top1 ? 32 : 32 / kSub));
    if (need) {
        warp_select<EPL>(need, buf, count, thr, p.kp, lane);
        if (((need >> lane) & 1u) && q >= 0) atomicMax(p.gthr + q, ord32(thr));
    }
}

// cta_group::1 MMAs; a cluster of cs (1 or 2) CTAs shares each rows tile.  (A cta_group::2 form — M = 256 across a
// CTA pair, rows tile split over the two shared memories — and clusters of 4 and 8 were built and measured in
// round 1: 12.50 ms, 12.7 ms and 13.5 ms against 12.0-12.3 ms for this form at 10M x 768, batch 1024; removed.)
//
// The rows tile is loaded with TMA MULTICAST when cs == 2: tmap_x must cover memory of THIS GPU only.  TMA
// multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's
// memory hung the GPU; lz_search_batched rejects a shadow that is not an allocation of the current device.
template <int EPL>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_topk_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x, GemmParams p) {
    constexpr int kStages = 4;
    constexpr int B_BYTES = BN * BK * 2;
    constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ unsigned char smem_raw[];
    unsigned char* tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-byte aligned
    uint64_t* full = reinterpret_cast<uint64_t*>(tiles + kStages * STAGE_BYTES);
    uint64_t* empty = full + kStages;
    uint64_t* tfull = empty + kStages;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // A cluster of cs CTAs walks the same (slice, tile) sequence; CTA `rank` owns q-tile qg*cs + rank and
    // loads 1/cs of every rows tile, multicast into all cs CTAs.
    const int cs = p.cs;
    const int rank = cs > 1 ? (int)cluster_rank() : 0;
    const uint16_t mc_mask = (uint16_t)((1u << cs) - 1u);
    const int n_items = p.items != nullptr ? *p.n_items_dev : p.n_qgroups * p.n_slices;
    const int cluster_id = blockIdx.x / cs, n_clusters = gridDim.x / cs;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_q);
        prefetch_tmap(&tmap_x);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], cs); }
        for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 4); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
    tc_fence_before();
    __syncthreads();
    if (cs > 1) cluster_sync();          // every CTA's barriers exist before anyone multicasts into them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            const int slice_rows = BN / cs;
            for (int item = cluster_id; item < n_items; item += n_clusters) {
                const WorkItem w = decode_item(p, item, rank);
                const int qt = w.qt;
                for (int t = 0; t < w.n_tiles; ++t) {
                    const int row0 = (int)(w.row_begin + (int64_t)t * w.stride);
                    for (int kb = 0; kb < p.kblocks; ++kb) {
                        { LZ_T0 mbar_wait(&empty[stage], phase ^ 1); LZ_ACC(4) }   // all cs consumers have released the stage
                        unsigned char* st = tiles + stage * STAGE_BYTES;
                        mbar_expect_tx(&full[stage], STAGE_BYTES);
                        tma_load_2d(st, &tmap_q, &full[stage], kb * BK, qt * BM);
                        if (cs > 1)
                            tma_load_2d_mc(st + A_BYTES + rank * slice_rows * BK * 2, &tmap_x, &full[stage], kb * BK,
                                           row0 + rank * slice_rows, mc_mask);
                        else
                            tma_load_2d(st + A_BYTES, &tmap_x, &full[stage], kb * BK, row0);
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            constexpr uint32_t idesc = idesc_f16_f32(BM, BN);
            uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
            for (int item = cluster_id; item < n_items; item += n_clusters) {
                const WorkItem w = decode_item(p, item, rank);
                for (int t = 0; t < w.n_tiles; ++t) {
                    { LZ_T0 mbar_wait(&tempty[acc], acc_phase ^ 1); LZ_ACC(2) }   // epilogue has drained this accumulator
                    tc_fence_after();
                    const uint32_t d = tmem_base + acc * BN;
                    for (int kb = 0; kb < p.kblocks; ++kb) {
                        { LZ_T0 mbar_wait(&full[stage], phase); LZ_ACC(3) }
                        tc_fence_after();
                        const uint64_t da = smem_desc_sw128(tiles + stage * STAGE_BYTES);
                        const uint64_t db = smem_desc_sw128(tiles + stage * STAGE_BYTES + A_BYTES);
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k) {     // 32 bytes along K per MMA: +2 in 16-byte units
                            mma_f16_ss(d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                        }
                        if (cs > 1) mma_commit_mc(&empty[stage], mc_mask);   // release the stage in every CTA of the cluster
                        else mma_commit(&empty[stage]);          // frees the smem stage when these MMAs retire
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                    mma_commit(&tfull[acc]);                      // accumulator complete
                    acc ^= 1;
                    if (acc == 0) acc_phase ^= 1;
                }
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue: one query per thread =====
        const int quarter = warp & 3;                            // the TMEM lane quarter this warp may read
        const int qlane = quarter * 32 + lane;
        uint32_t acc = 0, acc_phase = 0;
        for (int item = cluster_id; item < n_items; item += n_clusters) {
            const WorkItem w = decode_item(p, item, rank);
            const size_t slot0 = w.slot_base + qlane;   // flat: [slice][q-tile][query lane]; grouped: [item][query lane]
            u64* buf = p.cand + slot0 * p.cap;
            // grouped mode: the lane's row of the permuted query array stands for real query qmap[...] (thresholds are
            // shared between all items of the REAL query, whichever shard they scan)
            const int q = p.qmap != nullptr ? p.qmap[w.qt * BM + qlane] : (w.qt * BM + qlane < p.n_queries ? w.qt * BM + qlane : -1);
            const bool valid = q >= 0;
            int count = 0;
            float thr = valid ? -INFINITY : INFINITY;
            const float slack = (p.top1_eps != nullptr && valid) ? fmaxf(3.0f * (float)p.top1_eps[q], 1e-30f) : INFINITY;
            for (int t = 0; t < w.n_tiles; ++t) {
                const int64_t row0 = w.row_begin + (int64_t)t * w.stride;
                if (valid) {
                    // adopt the best threshold any CTA working on this query has reached: a row below some
                    // CTA's kp-th best coarse score cannot be among the kp best overall
                    u32 g;
                    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(g) : "l"(p.gthr + q) : "memory");
                    if (g != 0u) thr = fmaxf(thr, unord32(g));
                }
                mbar_wait(&tfull[acc], acc_phase);
                tc_fence_after();
#ifdef LZ_STATS
                long long te0 = clock64();
#endif
                // Candidates are GROUPS of kSub consecutive rows keyed by the group's best coarse score: one
                // compare per group, no per-row branch, so a tile costs the same whether or not it holds a
                // query's neighbours.  TMEM loads are double-buffered in registers (load chunk c+1, scan c).
                const uint32_t tacc = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BN;
                uint32_t va[32], vb[32];
                if (p.debug_skip & 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty[acc]);
                    acc ^= 1;
                    if (acc == 0) acc_phase ^= 1;
                    continue;
                }
                tmem_ld32(tacc, va);
#pragma unroll 1
                for (int c = 0; c < BN / 32; c += 2) {
                    tmem_ld_wait();
                    tmem_ld32(tacc + (c + 1) * 32, vb);
                    scan_chunk<EPL>(va, row0 + c * 32, w.row_end, p, buf, count, thr, slack, q, lane);
                    tmem_ld_wait();
                    if (c + 2 < BN / 32) tmem_ld32(tacc + (c + 2) * 32, va);
                    scan_chunk<EPL>(vb, row0 + (c + 1) * 32, w.row_end, p, buf, count, thr, slack, q, lane);
                }
                tc_fence_before();
                __syncwarp();
#ifdef LZ_STATS
                if (lane == 0) {
                    unsigned long long d = (unsigned long long)(clock64() - te0);
                    atomicAdd(&g_stats[0], d); atomicMax(&g_stats[1], d); atomicAdd(&g_stats[5], 1ull);
                    if (d > 6000) atomicAdd(&g_stats[6], 1ull);
                    if (d > 6000) atomicAdd(&g_stats[7], d - 6000);
                }
#endif
                if (lane == 0) mbar_arrive(&tempty[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
            const unsigned need = p.raw_items ? 0u : __ballot_sync(0xffffffffu, count > p.kp);
            if (need) warp_select<EPL>(need, buf, count, thr, p.kp, lane);
            p.cand_cnt[slot0] = count;
            p.cand_thr[slot0] = thr;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (cs > 1) cluster_sync();          // no CTA leaves while a peer may still signal its barriers
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ---- the exact pass: one CTA per query ---------------------------------------------------------------
struct SelectParams {
    const float* rows; int64_t ld; int dim; const double* norms;
    const float* queries; int64_t q_ld; const double* eps;
    const int32_t* tags; int want_tag; int64_t row_lo, row_hi;
    const u64* cand; const int32_t* cand_cnt; const float* cand_thr;
    int n_qtiles_pad, n_slices, kp, cap, k, cap_a, cap_b;   // n_qtiles_pad = n_qgroups * cs
    int sub;                // rows per candidate: kSub, or 1 in the top-1 tiny-table mode (small kernels only)
    int64_t id_base;
    int64_t* out_ids; double* out_scores; int32_t* out_counts; int32_t* certified;
    int32_t* fb_list; int32_t* fb_count;     // queries that failed certification, for the exact fallback
    const u32* gthr;                         // [n_queries] best threshold any CTA reached (lower bound of the kp-th best)
    int metric;                              // LZ_COSINE, or LZ_L2 (coarse key = x.q - |x|^2/2 = -(d - |q|^2)/2)
    // grouped (two-level) mode, probe != null: a query's candidates sit in the items of its probed shards
    const int64_t* probe;                    // [n_queries][nprobe] shard of each probe (-1 = none)
    const int32_t* pair_row;                 // [n_queries][nprobe] the pair's row in the permuted query array
    const int32_t* item_base;                // [n_shards] first item of a shard
    const int32_t* qt_base;                  // [n_shards] first permuted q-tile of a shard
    const int32_t* nsl;                      // [n_shards] row slices (items) per q-tile of a shard
    const int64_t* shard_off;                // [n_shards + 1]
    int nprobe, n_shards;
};

constexpr int kMaxSlices = 1024;             // candidate sources (slices) per query the exact pass can index

__global__ void __launch_bounds__(kSelThreads, 2) select_rescore_kernel(SelectParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    u64* key_a = reinterpret_cast<u64*>(smem);
    u64* key_b = key_a + p.cap_a;
    u32* id_a = reinterpret_cast<u32*>(key_b + p.cap_b);
    u32* id_b = id_a + p.cap_a;
    __shared__ LzSelector sa, sb;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q = blockIdx.x, qt = q / BM, ql = q % BM;
    if (tid == 0) {
        lz_sel_init(&sa, key_a, id_a, p.cap_a, p.kp);
        lz_sel_init(&sb, key_b, id_b, p.cap_b, p.k);
        // every CTA's threshold is a lower bound of the kp-th best group key overall, so groups below the
        // best of them cannot be among the kp best: start the merge from there instead of from -inf
        const u32 g = p.gthr[q];
        if (g != 0u) { sa.thr_key = (u64)g; sa.thr_id = 0xffffffffu; }
    }
    __syncthreads();
    // 1. merge the slices' candidates by coarse score.  The per-slice counts and thresholds are fetched by all
    // threads at once, then the (slice, entry) pairs are walked flat, 512 per round, so the global loads of a
    // round are independent (a slice-by-slice loop is one dependent round trip per slice).
    __shared__ float s_tw[kSelThreads / 32];
    __shared__ int s_nw[kSelThreads / 32];
    __shared__ int s_ps[kSelThreads / 32];
    __shared__ int s_slot[kMaxSlices];        // the (item, lane) slots that hold this query's candidates
    __shared__ int s_off[kMaxSlices + 1];
    __shared__ int s_nsrc;
    // 0. where this query's candidates are: flat mode — slot (slice, q-tile, lane) of every slice; grouped mode — the
    // items of the probed shards (slices of that shard for the q-tile the pair was placed in), laid end to end
    if (p.probe != nullptr) {
        const int per = (p.nprobe + kSelThreads - 1) / kSelThreads;
        const int j0 = min(p.nprobe, tid * per), j1 = min(p.nprobe, j0 + per);
        int mine = 0;
        for (int j = j0; j < j1; ++j) {
            const int64_t sh = p.probe[(size_t)q * p.nprobe + j];
            mine += sh >= 0 ? p.nsl[sh] : 0;
        }
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_ps[warp] = incl;
        __syncthreads();
        int run = incl - mine;
#pragma unroll
        for (int w = 0; w < kSelThreads / 32; ++w) if (w < warp) run += s_ps[w];
        for (int j = j0; j < j1; ++j) {
            const int64_t sh = p.probe[(size_t)q * p.nprobe + j];
            if (sh < 0) continue;
            const int pr = p.pair_row[(size_t)q * p.nprobe + j];
            const int n = p.nsl[sh];
            const int first = p.item_base[sh] + (pr / BM - p.qt_base[sh]) * n;
            for (int i = 0; i < n; ++i)
                if (run + i < kMaxSlices) s_slot[run + i] = (first + i) * BM + pr % BM;
            run += n;
        }
        if (tid == kSelThreads - 1) s_nsrc = min(run, kMaxSlices);     // the host sized the items so that this fits
        __syncthreads();
    } else {
        for (int sl = tid; sl < min(p.n_slices, kMaxSlices); sl += kSelThreads)
            s_slot[sl] = (int)(((size_t)sl * p.n_qtiles_pad + qt) * BM + ql);
        if (tid == 0) s_nsrc = p.n_slices;
        __syncthreads();
    }
    const int n_src = s_nsrc;
    {
        float t_loc = -INFINITY;
        int n_loc = 0;
        for (int sl = tid; sl < n_src; sl += kSelThreads) {
            const size_t slot = n_src <= kMaxSlices ? (size_t)s_slot[sl] : ((size_t)sl * p.n_qtiles_pad + qt) * BM + ql;
            t_loc = fmaxf(t_loc, p.cand_thr[slot]);
            n_loc += p.cand_cnt[slot];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            t_loc = fmaxf(t_loc, __shfl_xor_sync(0xffffffffu, t_loc, o));
            n_loc += __shfl_xor_sync(0xffffffffu, n_loc, o);
        }
        if (lane == 0) { s_tw[warp] = t_loc; s_nw[warp] = n_loc; }
    }
    __syncthreads();
    float T = -INFINITY;
    int total = 0;
#pragma unroll
    for (int w = 0; w < kSelThreads / 32; ++w) { T = fmaxf(T, s_tw[w]); total += s_nw[w]; }
    if (n_src <= kMaxSlices) {
        // With shared thresholds most slices leave only a few entries: lay the slices' counts end to end (block
        // scan) and walk the ENTRIES, 512 per round — not the n_slices * kp slots, most of them empty, each of
        // which cost a dependent count-then-entry round trip.
        const int per = (n_src + kSelThreads - 1) / kSelThreads;
        const int c0 = min(n_src, tid * per), c1 = min(n_src, c0 + per);
        int mine = 0;
        for (int sl = c0; sl < c1; ++sl) {
            const int c = p.cand_cnt[s_slot[sl]];
            s_off[sl + 1] = c;
            mine += c;
        }
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        __syncthreads();                                   // s_ps may still be read by the source layout above
        if (lane == 31) s_ps[warp] = incl;
        __syncthreads();
        int run = incl - mine;
#pragma unroll
        for (int w = 0; w < kSelThreads / 32; ++w) if (w < warp) run += s_ps[w];
        for (int sl = c0; sl < c1; ++sl) { const int c = s_off[sl + 1]; s_off[sl + 1] = run + c; run += c; }
        if (tid == 0) s_off[0] = 0;
        __syncthreads();
        const int Q = s_off[n_src];
        for (int base = 0; base < Q; base += 2 * kSelThreads) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int t = base + h * kSelThreads + tid;
                if (t < Q) {
                    int lo = 0, hi = n_src - 1;                                 // slice with s_off[sl] <= t < s_off[sl+1]
                    while (lo < hi) {
                        const int mid = (lo + hi + 1) >> 1;
                        if (s_off[mid] <= t) lo = mid; else hi = mid - 1;
                    }
                    const u64 kk = p.cand[(size_t)s_slot[lo] * p.cap + (t - s_off[lo])];
                    lz_sel_push(&sa, kk >> 32, 0xffffffffu - (u32)kk);
                }
            }
            lz_sel_maybe_compact<kSelThreads>(&sa, 2 * kSelThreads);
        }
    } else {
    const int64_t flat = (int64_t)p.n_slices * p.kp;                          // each slice left at most kp entries
    for (int64_t base = 0; base < flat; base += 2 * kSelThreads) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t i = base + h * kSelThreads + tid;
            if (i < flat) {
                const int sl = (int)(i / p.kp), j = (int)(i - (int64_t)sl * p.kp);
                const size_t slot = ((size_t)sl * p.n_qtiles_pad + qt) * BM + ql;
                if (j < p.cand_cnt[slot]) {
                    const u64 kk = p.cand[slot * p.cap + j];
                    lz_sel_push(&sa, kk >> 32, 0xffffffffu - (u32)kk);
                }
            }
        }
        lz_sel_maybe_compact<kSelThreads>(&sa, 2 * kSelThreads);
    }
    }
    __syncthreads();
    lz_sel_compact<kSelThreads>(&sa);
    int m = min(sa.count, p.kp);
    if (total > p.kp && sa.count >= p.kp) T = fmaxf(T, unord32((u32)key_a[p.kp - 1]));
    // Each surviving group's rows are eligible up to the end of ITS range: the scanned range (flat), or the shard the
    // group lies in (grouped: a binary search of the shard offsets).  Kept in the upper half of the group's key word.
    __syncthreads();
    for (int j = tid; j < m; j += kSelThreads) {
        u64 hi_row = (u64)p.row_hi;
        if (p.probe != nullptr) {
            int lo = 0, hi = p.n_shards - 1;                                    // shard with off[s] <= row < off[s+1]
            const int64_t r = (int64_t)id_a[j];
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (p.shard_off[mid] <= r) lo = mid; else hi = mid - 1;
            }
            hi_row = (u64)p.shard_off[lo + 1];
        }
        key_a[j] = (key_a[j] & 0xffffffffull) | (hi_row << 32);
    }
    __syncthreads();
    // Prune before rescoring: the k best groups lying wholly inside the range each hold an eligible row with
    // coarse >= c_k (the k-th such group's key), i.e. k rows with exact >= c_k - eps; a group whose best
    // coarse is below c_k - 2 eps holds only rows with exact < c_k - eps and cannot reach the top-k.
    // (Not applicable under a tag filter, where a group's best row may be ineligible.)
    __shared__ int m_eff;
    if (tid == 0) {
        int keep = m;
        if (p.tags == nullptr) {
            int full = 0;
            for (int j = 0; j < m; ++j) {
                if ((int64_t)id_a[j] + p.sub <= (int64_t)(key_a[j] >> 32) && ++full == p.k) {
                    const float cutoff = unord32((u32)key_a[j]) - 2.0f * (float)p.eps[q];
                    keep = j + 1;
                    while (keep < m && unord32((u32)key_a[keep]) >= cutoff) ++keep;
                    break;
                }
            }
        }
        m_eff = keep;
#ifdef LZ_STATS
        atomicAdd(&g_sel_stats[0], (unsigned long long)m);
        atomicAdd(&g_sel_stats[1], (unsigned long long)keep);
        atomicAdd(&g_sel_stats[2], 1ull);
#endif
    }
    __syncthreads();
    m = m_eff;
    // 2. exact float64 rescoring of the survivors, canonical summation order
    const float* qv = p.queries + (int64_t)q * p.q_ld;
    double qs = 0.0;
    for (int e = lane * 4; e < p.dim; e += 128) {
        const float4 x = *reinterpret_cast<const float4*>(qv + e);
        qs = fma((double)x.x, (double)x.x, qs);
        qs = fma((double)x.y, (double)x.y, qs);
        qs = fma((double)x.z, (double)x.z, qs);
        qs = fma((double)x.w, (double)x.w, qs);
    }
    const double qn = sqrt(lz_warp_sum(qs));
    // (group, row) pairs are walked flat, four per warp and step, and a step issues all of its loads (4 rows x up to
    // 3 segments of 128 floats) before the first use: the rows come from HBM, and with two rows and one segment in
    // flight this loop was a chain of memory round trips (the kernel's top stall)
    constexpr int kRow = 4, kSeg = 3, kW = kSelThreads / 32, kSteps = 4;
    const int n_pairs = m * p.sub;
    for (int base = 0; base < n_pairs; base += kRow * kW * kSteps) {
        for (int it = 0; it < kSteps; ++it) {
            int64_t row[kRow];
#pragma unroll
            for (int h = 0; h < kRow; ++h) {
                const int pi = base + (it * kRow + h) * kW + warp;
                row[h] = pi < n_pairs ? (int64_t)id_a[pi / p.sub] + pi % p.sub : -1;
                bool ok = row[h] >= 0 && row[h] < (int64_t)(key_a[pi / p.sub] >> 32);
                if (ok && p.tags) {
                    const int tg = p.tags[row[h]];
                    ok = tg >= 0 && (p.want_tag < 0 || tg == p.want_tag);
                }
                if (!ok) row[h] = -1;
            }
            if (row[0] < 0 && row[1] < 0 && row[2] < 0 && row[3] < 0) continue;      // warp-uniform
            double acc[kRow];
#pragma unroll
            for (int h = 0; h < kRow; ++h) acc[h] = 0.0;
            for (int e0 = lane * 4; e0 < p.dim; e0 += 128 * kSeg) {
                float4 xv[kRow][kSeg];
#pragma unroll
                for (int sg = 0; sg < kSeg; ++sg) {
                    const int e = e0 + 128 * sg;
#pragma unroll
                    for (int h = 0; h < kRow; ++h)
                        xv[h][sg] = (row[h] >= 0 && e < p.dim) ? lz_ldg_stream(p.rows + row[h] * p.ld + e)
                                                               : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int sg = 0; sg < kSeg; ++sg) {
                    const int e = e0 + 128 * sg;
                    if (e >= p.dim) break;
                    const float4 qq = *reinterpret_cast<const float4*>(qv + e);
#pragma unroll
                    for (int h = 0; h < kRow; ++h) {
                        if (p.metric == LZ_COSINE) {
                            acc[h] = fma((double)xv[h][sg].x, (double)qq.x, acc[h]);
                            acc[h] = fma((double)xv[h][sg].y, (double)qq.y, acc[h]);
                            acc[h] = fma((double)xv[h][sg].z, (double)qq.z, acc[h]);
                            acc[h] = fma((double)xv[h][sg].w, (double)qq.w, acc[h]);
                        } else {
                            const double d0 = (double)xv[h][sg].x - (double)qq.x, d1 = (double)xv[h][sg].y - (double)qq.y;
                            const double d2 = (double)xv[h][sg].z - (double)qq.z, d3 = (double)xv[h][sg].w - (double)qq.w;
                            acc[h] = fma(d0, d0, acc[h]);
                            acc[h] = fma(d1, d1, acc[h]);
                            acc[h] = fma(d2, d2, acc[h]);
                            acc[h] = fma(d3, d3, acc[h]);
                        }
                    }
                }
            }
#pragma unroll
            for (int h = 0; h < kRow; ++h) {
                const double dot = lz_warp_sum(acc[h]);
                if (lane == 0 && row[h] >= 0) {
                    double s;
                    if (p.metric == LZ_COSINE) {
                        const double den = p.norms[row[h]] * qn;
                        s = den > 0.0 ? dot / den : 0.0;
                        s += 0.0;
                    } else {
                        s = 0.0 - dot;                     // key order: larger = nearer
                    }
                    if (s == s) lz_sel_push(&sb, lz_ord(s), (u32)row[h]);
                }
            }
        }
        lz_sel_maybe_compact<kSelThreads>(&sb, kRow * kW * kSteps);
    }
    __syncthreads();
    lz_sel_compact<kSelThreads>(&sb);
    const int n = min(sb.count, p.k);
    for (int j = tid; j < p.k; j += kSelThreads) {
        if (j < n) {
            const double v = lz_unord(key_b[j]);
            p.out_ids[(size_t)q * p.k + j] = (int64_t)id_b[j] + p.id_base;
            p.out_scores[(size_t)q * p.k + j] = p.metric == LZ_COSINE ? v : 0.0 - v;
        } else {
            p.out_ids[(size_t)q * p.k + j] = -1;
            p.out_scores[(size_t)q * p.k + j] = __longlong_as_double(0x7ff8000000000000ll);
        }
    }
    if (tid == 0) {
        p.out_counts[q] = n;
        // 3. certify: rows dropped anywhere had coarse <= T, i.e. exact <= T + eps
        bool cert = T == -INFINITY;
        if (!cert && n >= p.k) {
            const double ek = lz_unord(key_b[p.k - 1]);                       // cosine, or -distance
            const double e_coarse = p.metric == LZ_COSINE ? ek : 0.5 * (ek + qn * qn);   // -(d - |q|^2)/2
            cert = e_coarse > (double)T + p.eps[q];
        }
        p.certified[q] = cert ? 1 : 0;
        if (!cert) p.fb_list[atomicAdd(p.fb_count, 1)] = q;
    }
}


// Small-k variant of the exact pass (k <= 3: dedup, linking, centroid assignment — often millions of
// "queries"): one WARP per query, no shared memory.  Same rule as above: find c_k among the candidate
// groups, rescore every group within 2 eps of it, keep the best k rows in lane 0, certify against the
// largest threshold any slice dropped under.  Nothing is discarded here, so T is just that maximum.
constexpr int kSmallK = 3;

// pass 1: c_k = k-th largest group key among groups wholly inside the range (k rounds of warp max); returns the
// rescoring cutoff c_k - 2 eps and the largest threshold any slice dropped candidates under (T).
__device__ __forceinline__ float small_cutoff(const SelectParams& p, int q, int lane, float eps, float* T_out) {
    const int qt = q / BM, ql = q % BM;
    float T = -INFINITY;
    u64 below = ~0ull;                     // keys must be < below to be picked in this round
    float ck = -INFINITY;
    bool have_k = true;
    for (int round = 0; round < p.k; ++round) {
        u64 best = 0ull;
        for (int sl = 0; sl < p.n_slices; ++sl) {
            const size_t slot = ((size_t)sl * p.n_qtiles_pad + qt) * BM + ql;
            const int cnt = p.cand_cnt[slot];
            if (round == 0) T = fmaxf(T, p.cand_thr[slot]);
            const u64* src = p.cand + slot * p.cap;
            for (int j = lane; j < cnt; j += 32) {
                const u64 kk = src[j];
                const int64_t r0 = (int64_t)(0xffffffffu - (u32)kk);
                if (kk < below && kk > best && (p.tags != nullptr || r0 + p.sub <= p.row_hi)) best = kk;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other > best ? other : best;
        }
        if (best == 0ull) { have_k = false; break; }
        below = best;
        ck = unord32((u32)(best >> 32));
    }
    *T_out = T;
    return (have_k && p.tags == nullptr) ? ck - 2.0f * eps : -INFINITY;
}

// pass 2, one warp: rescore the rows of every group at or above the cutoff straight from global memory; lane 0
// keeps the best k, writes the result and certifies it.
__device__ __forceinline__ void small_rescore_direct(const SelectParams& p, int q, int lane, float eps, float T, float cutoff) {
    const int qt = q / BM, ql = q % BM;
    // pass 2: rescore the rows of every group at or above the cutoff; lane 0 keeps the best k
    const float* qv = p.queries + (int64_t)q * p.q_ld;
    double qs = 0.0;
    for (int e = lane * 4; e < p.dim; e += 128) {
        const float4 x = *reinterpret_cast<const float4*>(qv + e);
        qs = fma((double)x.x, (double)x.x, qs);
        qs = fma((double)x.y, (double)x.y, qs);
        qs = fma((double)x.z, (double)x.z, qs);
        qs = fma((double)x.w, (double)x.w, qs);
    }
    const double qn = sqrt(lz_warp_sum(qs));
    u64 bk[kSmallK];
    u32 bi[kSmallK];
#pragma unroll
    for (int i = 0; i < kSmallK; ++i) { bk[i] = 0ull; bi[i] = 0xffffffffu; }
    int found = 0;
    for (int sl = 0; sl < p.n_slices; ++sl) {
        const size_t slot = ((size_t)sl * p.n_qtiles_pad + qt) * BM + ql;
        const int cnt = p.cand_cnt[slot];
        const u64* src = p.cand + slot * p.cap;
        // 32 entries are fetched at once (one per lane); the few at or above the cutoff are then walked by
        // ballot — a lane-uniform loop over all entries would be one dependent L2 round trip per entry
        for (int jb = 0; jb < cnt; jb += 32) {
            const u64 mine = jb + lane < cnt ? src[jb + lane] : 0ull;
            unsigned live = __ballot_sync(0xffffffffu, jb + lane < cnt && unord32((u32)(mine >> 32)) >= cutoff);
            while (live) {
                const int b = __ffs(live) - 1;
                live &= live - 1;
                const u64 kk = __shfl_sync(0xffffffffu, mine, b);
                const int64_t r0 = (int64_t)(0xffffffffu - (u32)kk);
                // the group's rows four at a time: 4 x dim/128 independent 16-byte loads per lane in flight
                for (int rb = 0; rb < p.sub; rb += 4) {
                    int64_t row[4];
                    double acc[4];
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        row[h] = r0 + rb + h;
                        bool ok = rb + h < p.sub && row[h] < p.row_hi;
                        if (ok && p.tags) {
                            const int tg = p.tags[row[h]];
                            ok = tg >= 0 && (p.want_tag < 0 || tg == p.want_tag);
                        }
                        if (!ok) row[h] = -1;
                        acc[h] = 0.0;
                    }
                    for (int e = lane * 4; e < p.dim; e += 128) {
                        const float4 qq = *reinterpret_cast<const float4*>(qv + e);
                        float4 xv[4];
#pragma unroll
                        for (int h = 0; h < 4; ++h)
                            xv[h] = row[h] >= 0 ? lz_ldg_stream(p.rows + row[h] * p.ld + e) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                        for (int h = 0; h < 4; ++h) {
                            if (p.metric == LZ_COSINE) {
                                acc[h] = fma((double)xv[h].x, (double)qq.x, acc[h]);
                                acc[h] = fma((double)xv[h].y, (double)qq.y, acc[h]);
                                acc[h] = fma((double)xv[h].z, (double)qq.z, acc[h]);
                                acc[h] = fma((double)xv[h].w, (double)qq.w, acc[h]);
                            } else {
                                const double d0 = (double)xv[h].x - (double)qq.x, d1 = (double)xv[h].y - (double)qq.y;
                                const double d2 = (double)xv[h].z - (double)qq.z, d3 = (double)xv[h].w - (double)qq.w;
                                acc[h] = fma(d0, d0, acc[h]);
                                acc[h] = fma(d1, d1, acc[h]);
                                acc[h] = fma(d2, d2, acc[h]);
                                acc[h] = fma(d3, d3, acc[h]);
                            }
                        }
                    }
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        const double dot = lz_warp_sum(acc[h]);
                        if (row[h] < 0) continue;                                  // warp-uniform
                        double s;
                        if (p.metric == LZ_COSINE) {
                            const double den = p.norms[row[h]] * qn;
                            s = den > 0.0 ? dot / den : 0.0;
                            s += 0.0;
                        } else {
                            s = 0.0 - dot;
                        }
                        if (s != s) continue;
                        u64 key = lz_ord(s);
                        u32 id = (u32)row[h];
                        ++found;
#pragma unroll
                        for (int i = 0; i < kSmallK; ++i) {                     // insertion into the sorted best-k
                            if (i < p.k && lz_better(key, id, bk[i], bi[i])) {
                                const u64 tk = bk[i]; const u32 ti = bi[i];
                                bk[i] = key; bi[i] = id; key = tk; id = ti;
                            }
                        }
                    }
                }
            }
        }
    }
    if (lane == 0) {
        const int n = found < p.k ? found : p.k;
        for (int j = 0; j < p.k; ++j) {
            const bool has = j < n;
            const double v = has ? lz_unord(bk[j]) : 0.0;
            p.out_ids[(size_t)q * p.k + j] = has ? (int64_t)bi[j] + p.id_base : -1;
            p.out_scores[(size_t)q * p.k + j] = has ? (p.metric == LZ_COSINE ? v : 0.0 - v)
                                                    : __longlong_as_double(0x7ff8000000000000ll);
        }
        p.out_counts[q] = n;
        bool cert = T == -INFINITY;
        if (!cert && n >= p.k) {
            const double ek = lz_unord(bk[p.k - 1]);
            const double e_coarse = p.metric == LZ_COSINE ? ek : 0.5 * (ek + qn * qn);
            cert = e_coarse > (double)T + (double)eps;
        }
        p.certified[q] = cert ? 1 : 0;
        if (!cert) p.fb_list[atomicAdd(p.fb_count, 1)] = q;
    }
}

__global__ void __launch_bounds__(kSelThreads) select_rescore_small_kernel(SelectParams p, int n_queries) {
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * (kSelThreads / 32) + (threadIdx.x >> 5);
    if (q >= n_queries) return;
    const float eps = (float)p.eps[q];
    float T;
    const float cutoff = small_cutoff(p, q, lane, eps, &T);
    small_rescore_direct(p, q, lane, eps, T, cutoff);
}

// Top-1 against a TINY table (shard reassignment: a million rows against ~1000 centroids).  The coarse pass
// runs in its top-1 mode there (single-row candidates within 3 eps of the running best: a handful per
// query).  Rescoring one warp per query straight from global memory would re-read a centroid row from L2
// for every query — as many bytes again as the queries themselves — so a CTA takes kBucketQ consecutive
// queries, finds each one's single surviving row (one thread per query), counting-sorts the queries by the
// row's group of kSub in shared memory, and walks the groups: the group's rows are staged in shared memory
// once and each query is scored against its row in float64, canonical summation order (scores equal the
// exact scan's bit for bit).  Queries with several surviving rows take the direct path.
__device__ __noinline__ void small_direct_call(const SelectParams& p, int q, int lane) {
    const float eps = (float)p.eps[q];
    float T;
    const float cutoff = small_cutoff(p, q, lane, eps, &T);
    small_rescore_direct(p, q, lane, eps, T, cutoff);
}

constexpr int kBucketQ = 1024;
constexpr int kBucketGroups = 256;

__global__ void __launch_bounds__(kSelThreads) assign_rescore_kernel(SelectParams p, int n_queries) {
    extern __shared__ __align__(16) unsigned char smem[];
    float* s_rows = reinterpret_cast<float*>(smem);                          // [kSub][dim]
    __shared__ int s_pick[kBucketQ];                                          // surviving row - row_lo, or -2
    __shared__ int s_sorted[kBucketQ];
    __shared__ float s_T[kBucketQ];
    __shared__ int s_hist[kBucketGroups], s_off[kBucketGroups + 1], s_cur[kBucketGroups];
    __shared__ double s_norm[kSub];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kSelThreads / 32;
    const int q0 = blockIdx.x * kBucketQ;
    const int nq = min(kBucketQ, n_queries - q0);
    for (int g = tid; g < kBucketGroups; g += kSelThreads) { s_hist[g] = 0; s_cur[g] = 0; }
    __syncthreads();
    // A. one THREAD per query: best candidate row, cutoff = its key - 2 eps, survivors counted.  A single
    // survivor is bucketed; anything else is scored by a warp straight from global memory afterwards.
    for (int i = tid; i < nq; i += kSelThreads) {
        const int q = q0 + i;                                                 // n_slices == 1: slot == query
        const int cnt = p.cand_cnt[q];
        const u64* src = p.cand + (size_t)q * p.cap;
        u64 best = 0ull;
        for (int j = 0; j < cnt; ++j) {
            const u64 kk = src[j];
            best = kk > best ? kk : best;
        }
        int pick = -2;
        if (best != 0ull) {
            const float cutoff = unord32((u32)(best >> 32)) - 2.0f * (float)p.eps[q];
            int n_live = 0;
            for (int j = 0; j < cnt; ++j) n_live += unord32((u32)(src[j] >> 32)) >= cutoff;
            const int rel = (int)((0xffffffffu - (u32)best) - (u32)p.row_lo);
            if (n_live == 1 && rel < kBucketGroups * kSub) pick = rel;
        }
        s_pick[i] = pick;
        s_T[i] = p.cand_thr[q];
        // neighbouring rows mostly share a group: one shared-memory atomic per distinct group and warp
        const int g = pick >= 0 ? pick / kSub : -1;
        const unsigned act = __activemask();
        const unsigned same = __match_any_sync(act, g);
        if (g >= 0 && lane == __ffs(same) - 1) atomicAdd(&s_hist[g], __popc(same));
    }
    __syncthreads();
    for (int i = warp; i < nq; i += kWarps) {
        if (s_pick[i] != -2) continue;                                        // warp-uniform
        small_direct_call(p, q0 + i, lane);                                   // rare: kept out of line
    }
    // B. counting sort of the bucketed queries by group
    if (tid == 0) {
        int run = 0;
        for (int g = 0; g < kBucketGroups; ++g) { s_off[g] = run; run += s_hist[g]; }
        s_off[kBucketGroups] = run;
    }
    __syncthreads();
    for (int i = tid; i < nq; i += kSelThreads) {
        const int g = s_pick[i] >= 0 ? s_pick[i] / kSub : -1;
        const unsigned act = __activemask();
        const unsigned same = __match_any_sync(act, g);
        const int leader = __ffs(same) - 1;
        int base = 0;
        if (g >= 0 && lane == leader) base = atomicAdd(&s_cur[g], __popc(same));
        base = __shfl_sync(act, base, leader);
        if (g >= 0) s_sorted[s_off[g] + base + __popc(same & ((1u << lane) - 1u))] = i;
    }
    __syncthreads();
    // C. group by group
    const int n4 = p.dim / 4;
    for (int g = 0; g < kBucketGroups; ++g) {
        const int lo = s_off[g], hi = s_off[g + 1];
        if (lo == hi) continue;                                               // CTA-uniform
        __syncthreads();                                                      // previous group's readers done
        const int64_t r0 = p.row_lo + (int64_t)g * kSub;
        const int n_in = (int)min((int64_t)kSub, p.row_hi - r0);
        for (int i = tid; i < n_in * n4; i += kSelThreads) {
            const int r = i / n4, e = i - r * n4;
            reinterpret_cast<float4*>(s_rows)[i] = lz_ldg_stream(p.rows + (r0 + r) * p.ld + e * 4);
        }
        if (tid < n_in) s_norm[tid] = p.norms[r0 + tid];
        __syncthreads();
        // two queries per warp and step: both queries' loads are in flight together
        for (int j = lo + 2 * warp; j < hi; j += 2 * kWarps) {
            const bool two = j + 1 < hi;
            const int iq[2] = {s_sorted[j], two ? s_sorted[j + 1] : s_sorted[j]};
            const float* qp[2] = {p.queries + (int64_t)(q0 + iq[0]) * p.q_ld, p.queries + (int64_t)(q0 + iq[1]) * p.q_ld};
            const int rr[2] = {s_pick[iq[0]] - g * kSub, s_pick[iq[1]] - g * kSub};
            double dot[2] = {0.0, 0.0}, qs[2] = {0.0, 0.0};
            constexpr int kSeg = 6;                      // segments of 128 floats fetched per round trip
            for (int e0 = lane * 4; e0 < p.dim; e0 += 128 * kSeg) {
                float4 qv[2][kSeg];
#pragma unroll
                for (int sg = 0; sg < kSeg; ++sg) {
                    const int e = e0 + 128 * sg;
#pragma unroll
                    for (int h = 0; h < 2; ++h)
                        qv[h][sg] = e < p.dim ? lz_ldg_stream(qp[h] + e) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int sg = 0; sg < kSeg; ++sg) {
                    const int e = e0 + 128 * sg;
                    if (e >= p.dim) break;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const float4 x = *reinterpret_cast<const float4*>(s_rows + rr[h] * p.dim + e);
                        const double q0d = qv[h][sg].x, q1d = qv[h][sg].y, q2d = qv[h][sg].z, q3d = qv[h][sg].w;
                        qs[h] = fma(q0d, q0d, qs[h]); qs[h] = fma(q1d, q1d, qs[h]);
                        qs[h] = fma(q2d, q2d, qs[h]); qs[h] = fma(q3d, q3d, qs[h]);
                        if (p.metric == LZ_COSINE) {
                            dot[h] = fma((double)x.x, q0d, dot[h]); dot[h] = fma((double)x.y, q1d, dot[h]);
                            dot[h] = fma((double)x.z, q2d, dot[h]); dot[h] = fma((double)x.w, q3d, dot[h]);
                        } else {
                            const double d0 = (double)x.x - q0d, d1 = (double)x.y - q1d;
                            const double d2 = (double)x.z - q2d, d3 = (double)x.w - q3d;
                            dot[h] = fma(d0, d0, dot[h]); dot[h] = fma(d1, d1, dot[h]);
                            dot[h] = fma(d2, d2, dot[h]); dot[h] = fma(d3, d3, dot[h]);
                        }
                    }
                }
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const double d = lz_warp_sum(dot[h]);
                const double qn = sqrt(lz_warp_sum(qs[h]));
                if (lane != 0 || (h == 1 && !two)) continue;
                const int q = q0 + iq[h];
                double sc;
                if (p.metric == LZ_COSINE) {
                    const double den = s_norm[rr[h]] * qn;
                    sc = den > 0.0 ? d / den : 0.0;
                    sc += 0.0;
                } else {
                    sc = 0.0 - d;
                }
                const bool has = sc == sc;
                p.out_ids[q] = has ? r0 + rr[h] + p.id_base : -1;
                p.out_scores[q] = has ? (p.metric == LZ_COSINE ? sc : 0.0 - sc) : __longlong_as_double(0x7ff8000000000000ll);
                p.out_counts[q] = has ? 1 : 0;
                const float T = s_T[iq[h]];
                bool cert = T == -INFINITY;
                if (!cert && has) {
                    const double e_coarse = p.metric == LZ_COSINE ? sc : 0.5 * (sc + qn * qn);
                    cert = e_coarse > (double)T + p.eps[q];
                }
                p.certified[q] = cert ? 1 : 0;
                if (!cert) p.fb_list[atomicAdd(p.fb_count, 1)] = q;
            }
        }
    }
}

// Top-1 against a tiny table, any number of surviving candidates per query.  The kernel above handles queries with ONE
// surviving row; everything else used to fall to a warp per query reading its candidate rows from L2 — as many bytes
// again per candidate as the query itself, which made nearly coincident centroids (the normal k-means shape, k above
// the number of clusters) 5-10x slower than separated ones.  Here a CTA takes kMultiQ consecutive queries, collects
// every query's surviving rows (up to kMultiLive), stages the DISTINCT table rows the block needs in shared memory
// (rows of one block mostly share their candidates: the store keeps rows shard by shard), and a warp then holds one
// query in registers as float64 and scores it against each of its candidates from shared memory: the query is read
// from HBM once, a candidate costs dim*4 bytes of shared-memory traffic.  When a block needs more distinct table rows
// than fit, they are staged in rounds and a query is re-read in the rounds that hold some of its candidates.
// Summation order per pair is the exact scan's, so scores are bit-identical to it.
constexpr int kMultiQ = 512;             // queries per CTA
constexpr int kMultiLive = 16;           // surviving candidates kept per query (more: the direct path)
constexpr int kMultiThreads = 512;
constexpr int kMultiTable = kBucketGroups * kSub;     // table rows the bitmap covers (2048)

// float64 scores of N (1..4) cached table rows against one query held in registers: N independent accumulation chains,
// each in the exact scan's summation order; lane-level partial sums are reduced across the warp.
template <int NSEG, int METRIC, int N>
__device__ __forceinline__ void assign_dots(const float* const (&xr)[4], const double (&qd)[NSEG][4], double (&acc)[4]) {
#pragma unroll
    for (int c = 0; c < N; ++c) acc[c] = 0.0;
#pragma unroll
    for (int sg = 0; sg < NSEG; ++sg) {
        float4 x[N];
#pragma unroll
        for (int c = 0; c < N; ++c) x[c] = *reinterpret_cast<const float4*>(xr[c] + 128 * sg);
#pragma unroll
        for (int c = 0; c < N; ++c) {
            if (METRIC == LZ_COSINE) {
                acc[c] = fma((double)x[c].x, qd[sg][0], acc[c]); acc[c] = fma((double)x[c].y, qd[sg][1], acc[c]);
                acc[c] = fma((double)x[c].z, qd[sg][2], acc[c]); acc[c] = fma((double)x[c].w, qd[sg][3], acc[c]);
            } else {
                const double d0 = (double)x[c].x - qd[sg][0], d1 = (double)x[c].y - qd[sg][1];
                const double d2 = (double)x[c].z - qd[sg][2], d3 = (double)x[c].w - qd[sg][3];
                acc[c] = fma(d0, d0, acc[c]); acc[c] = fma(d1, d1, acc[c]);
                acc[c] = fma(d2, d2, acc[c]); acc[c] = fma(d3, d3, acc[c]);
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
        for (int c = 0; c < N; ++c) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
}

template <int NSEG, int METRIC>
__global__ void __launch_bounds__(kMultiThreads) assign_multi_kernel(SelectParams p, int n_queries, int cache_rows) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int kStride = NSEG * 128;                                                // floats per cached row, zero padded
    float* s_rows = reinterpret_cast<float*>(smem);                                    // [cache_rows][kStride]
    double* s_cnorm = reinterpret_cast<double*>(s_rows + (size_t)cache_rows * kStride); // [cache_rows]
    u64* s_bkey = reinterpret_cast<u64*>(s_cnorm + cache_rows);                         // [kMultiQ] best exact key so far
    float* s_T = reinterpret_cast<float*>(s_bkey + kMultiQ);                            // [kMultiQ]
    u32* s_bid = reinterpret_cast<u32*>(s_T + kMultiQ);                                 // [kMultiQ]
    unsigned short* s_list = reinterpret_cast<unsigned short*>(s_bid + kMultiQ);        // [kMultiQ][kMultiLive] table row - row_lo
    signed char* s_nlive = reinterpret_cast<signed char*>(s_list + (size_t)kMultiQ * kMultiLive);   // [kMultiQ], -1 = direct path
    // per warp two staging buffers for query rows: the next query is fetched (cp.async) while the current one is scored
    float* s_stage = reinterpret_cast<float*>(smem + ((size_t)cache_rows * (kStride * 4 + 8) + (size_t)kMultiQ * (8 + 4 + 4 + 2 * kMultiLive + 1) + 15) / 16 * 16);
    __shared__ u32 s_bits[kMultiTable / 32];
    __shared__ int s_pref[kMultiTable / 32 + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kMultiThreads / 32;
    const int q0 = blockIdx.x * kMultiQ;
    const int nq = min(kMultiQ, n_queries - q0);
    for (int i = tid; i < kMultiTable / 32; i += kMultiThreads) s_bits[i] = 0u;
    for (int i = tid; i < kWarps * 2 * kStride; i += kMultiThreads) s_stage[i] = 0.f;     // the pads beyond dim stay zero
    __syncthreads();
    // A. one thread per query: best coarse candidate, the rows within 2 eps of it, the block's bitmap of needed rows
    for (int i = tid; i < nq; i += kMultiThreads) {
        const int q = q0 + i;                                                 // n_slices == 1: slot == query
        const int cnt = p.cand_cnt[q];
        const u64* src = p.cand + (size_t)q * p.cap;
        u64 best = 0ull;
        for (int j = 0; j < cnt; ++j) {
            const u64 kk = src[j];
            best = kk > best ? kk : best;
        }
        int n_live = 0;
        if (best != 0ull) {
            const float cutoff = unord32((u32)(best >> 32)) - 2.0f * (float)p.eps[q];
            for (int j = 0; j < cnt; ++j) {
                const u64 kk = src[j];
                if (unord32((u32)(kk >> 32)) >= cutoff) {
                    const u32 rel = (0xffffffffu - (u32)kk) - (u32)p.row_lo;
                    if (n_live < kMultiLive && rel < (u32)kMultiTable) s_list[i * kMultiLive + n_live] = (unsigned short)rel;
                    else n_live = kMultiLive + 64;                            // too many (or out of the bitmap): direct path
                    ++n_live;
                }
            }
        }
        if (n_live > kMultiLive) n_live = -1;
        s_nlive[i] = (signed char)n_live;
        for (int j = 0; j < n_live; ++j) {
            const int rel = s_list[i * kMultiLive + j];
            atomicOr(&s_bits[rel >> 5], 1u << (rel & 31));
        }
        s_T[i] = p.cand_thr[q];
        s_bkey[i] = 0ull;
        s_bid[i] = 0xffffffffu;
    }
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int w = 0; w < kMultiTable / 32; ++w) { s_pref[w] = run; run += __popc(s_bits[w]); }
        s_pref[kMultiTable / 32] = run;
    }
    __syncthreads();
    const int n_distinct = s_pref[kMultiTable / 32];
    const int n4 = p.dim / 4;
    for (int r0 = 0; r0 < n_distinct; r0 += cache_rows) {
        const int r1 = min(n_distinct, r0 + cache_rows);
        __syncthreads();                                                       // the previous round's readers are done
        // stage the table rows with rank [r0, r1): warp w takes slots w, w + kWarps, ...
        for (int slot = warp; slot < r1 - r0; slot += kWarps) {
            const int want = r0 + slot;                                        // rank of the row to fetch
            int w = 0;                                                         // word with s_pref[w] <= want < s_pref[w+1]
            {
                int lo = 0, hi = kMultiTable / 32 - 1;
                while (lo < hi) {
                    const int mid = (lo + hi + 1) >> 1;
                    if (s_pref[mid] <= want) lo = mid; else hi = mid - 1;
                }
                w = lo;
            }
            u32 bits = s_bits[w];
            for (int skip = want - s_pref[w]; skip > 0; --skip) bits &= bits - 1;
            const int rel = w * 32 + __ffs(bits) - 1;
            const float* src = p.rows + (p.row_lo + rel) * p.ld;
            for (int e = lane; e < kStride / 4; e += 32)
                reinterpret_cast<float4*>(s_rows + (size_t)slot * kStride)[e] =
                    e < n4 ? lz_ldg_stream(src + e * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (lane == 0) s_cnorm[slot] = METRIC == LZ_COSINE ? p.norms[p.row_lo + rel] : 0.0;
        }
        __syncthreads();
        // every warp walks the block's queries; a query with candidates in this round is fetched once — asynchronously
        // into the warp's staging buffer while the previous query is being scored — and held as float64 in registers
        float* stage = s_stage + (size_t)warp * 2 * kStride;
        const bool one_round = r0 == 0 && r1 == n_distinct;
        auto has_work = [&](int i) -> bool {
            const int nl = s_nlive[i];
            if (nl <= 0) return false;
            if (one_round) return true;
            for (int j = 0; j < nl; ++j) {
                const int rel = s_list[i * kMultiLive + j];
                const int rank = s_pref[rel >> 5] + __popc(s_bits[rel >> 5] & ((1u << (rel & 31)) - 1u));
                if (rank >= r0 && rank < r1) return true;
            }
            return false;
        };
        auto next_row = [&](int i) -> int {
            i += kWarps;
            while (i < nq && !has_work(i)) i += kWarps;
            return i;
        };
        auto issue = [&](int i, int b) {
            const float* qp = p.queries + (int64_t)(q0 + i) * p.q_ld;
#pragma unroll
            for (int sg = 0; sg < NSEG; ++sg) {
                const int e = lane * 4 + 128 * sg;
                if (e < p.dim)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stage + b * kStride + e)), "l"(qp + e) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        // two queries in flight per warp: a staging buffer is handed to the query after next as soon as the current
        // one sits in registers, so 6 KB per warp (96 KB per SM) are on their way while a query is being scored
        auto skip_issue = [&]() { asm volatile("cp.async.commit_group;" ::: "memory"); };
        int cur = next_row(warp - kWarps), buf = 0;
        int nxt = cur < nq ? next_row(cur) : nq;
        if (cur < nq) issue(cur, 0); else skip_issue();
        if (nxt < nq) issue(nxt, 1); else skip_issue();
        while (cur < nq) {
            const int i = cur;
            asm volatile("cp.async.wait_group 1;" ::: "memory");             // the older of the two has landed
            const int nl = s_nlive[i];
            int slots[kMultiLive];
            int n_here = 0;
#pragma unroll
            for (int j = 0; j < kMultiLive; ++j) {
                if (j < nl) {
                    const int rel = s_list[i * kMultiLive + j];
                    const int rank = s_pref[rel >> 5] + __popc(s_bits[rel >> 5] & ((1u << (rel & 31)) - 1u));
                    if (rank >= r0 && rank < r1) slots[n_here++] = (rank - r0) | (rel << 16);
                }
            }
            double qd[NSEG][4];
            double qs = 0.0;
#pragma unroll
            for (int sg = 0; sg < NSEG; ++sg) {
                const float4 v = *reinterpret_cast<const float4*>(stage + buf * kStride + lane * 4 + 128 * sg);   // pads stay zero
                qd[sg][0] = v.x; qd[sg][1] = v.y; qd[sg][2] = v.z; qd[sg][3] = v.w;
            }
            __syncwarp();                                                      // every lane has its part of the query
            {
                const int after = nxt < nq ? next_row(nxt) : nq;
                if (after < nq) issue(after, buf); else skip_issue();
                cur = nxt;
                nxt = after;
                buf ^= 1;
            }
#pragma unroll
            for (int sg = 0; sg < NSEG; ++sg)
#pragma unroll
                for (int c = 0; c < 4; ++c) qs = fma(qd[sg][c], qd[sg][c], qs);
            const double qn = sqrt(lz_warp_sum(qs));
            u64 bkey = s_bkey[i];
            u32 bid = s_bid[i];
            constexpr int kAt = 4;
            // (A float32 pre-pass over the candidates, keeping only those within 2e-5 of the best for the float64 pass, was
            // tried: parity-green but slower, 4.62 vs 4.24 ms with ~10 candidates per row — the loop is bound by the
            // shared-memory reads of the cached table rows, which a pre-pass repeats, not by the float64 arithmetic.)
            // Up to four candidates at a time: independent accumulation chains, each in the exact scan's order
            for (int j0 = 0; j0 < n_here; j0 += kAt) {
                const int n_valid = min(kAt, n_here - j0);
                const float* xr[kAt];
                int slot[kAt], rel[kAt];
#pragma unroll
                for (int c = 0; c < kAt; ++c) {
                    const int sv = slots[min(j0 + c, n_here - 1)];
                    slot[c] = sv & 0xffff;
                    rel[c] = sv >> 16;
                    xr[c] = s_rows + (size_t)slot[c] * kStride + lane * 4;
                }
                double acc[kAt];
                if (n_valid == 4) assign_dots<NSEG, METRIC, 4>(xr, qd, acc);          // warp-uniform dispatch: no work on padding
                else if (n_valid == 1) assign_dots<NSEG, METRIC, 1>(xr, qd, acc);
                else if (n_valid == 2) assign_dots<NSEG, METRIC, 2>(xr, qd, acc);
                else assign_dots<NSEG, METRIC, 3>(xr, qd, acc);
#pragma unroll
                for (int c = 0; c < kAt; ++c) {
                    if (c >= n_valid) break;                                   // warp-uniform
                    double sc;
                    if (METRIC == LZ_COSINE) {
                        const double den = s_cnorm[slot[c]] * qn;
                        sc = den > 0.0 ? acc[c] / den : 0.0;
                        sc += 0.0;
                    } else {
                        sc = 0.0 - acc[c];
                    }
                    if (sc == sc) {
                        const u64 key = lz_ord(sc);
                        const u32 id = (u32)(p.row_lo + rel[c]);
                        if (lz_better(key, id, bkey, bid)) { bkey = key; bid = id; }
                    }
                }
            }
            if (lane == 0) { s_bkey[i] = bkey; s_bid[i] = bid; }
        }
    }
    __syncthreads();
    // results + certification: every row dropped by the coarse pass had coarse <= T = c_1 - 3 eps
    for (int i = tid; i < nq; i += kMultiThreads) {
        if (s_nlive[i] < 0) continue;
        const int q = q0 + i;
        const bool has = s_bkey[i] != 0ull;
        const double v = has ? lz_unord(s_bkey[i]) : 0.0;
        p.out_ids[q] = has ? (int64_t)s_bid[i] + p.id_base : -1;
        p.out_scores[q] = has ? (METRIC == LZ_COSINE ? v : 0.0 - v) : __longlong_as_double(0x7ff8000000000000ll);
        p.out_counts[q] = has ? 1 : 0;
        const float T = s_T[i];
        bool cert = T == -INFINITY;
        if (!cert && has) {
            double e_coarse = v;
            if (METRIC != LZ_COSINE) {                                          // -(d - |q|^2)/2 needs |q|^2
                const float* qp = p.queries + (int64_t)q * p.q_ld;
                double qq = 0.0;
                for (int e = 0; e < p.dim; ++e) qq = fma((double)qp[e], (double)qp[e], qq);
                e_coarse = 0.5 * (v + qq);
            }
            cert = e_coarse > (double)T + p.eps[q];
        }
        p.certified[q] = cert ? 1 : 0;
        if (!cert) p.fb_list[atomicAdd(p.fb_count, 1)] = q;
    }
    __syncthreads();
    for (int i = warp; i < nq; i += kWarps) {
        if (s_nlive[i] >= 0) continue;                                        // warp-uniform
        small_direct_call(p, q0 + i, lane);                                   // rare: more than kMultiLive survivors
    }
}

// Threshold seeding: the coarse pass is first run over a strided sample of the rows tiles (64 tiles, one per
// item, so nothing is compacted); here each query's kp-th best sampled group key becomes its starting
// threshold.  The kp-th best of a subset can only be lower than the kp-th best of all groups, so the seed is
// a valid lower bound; it lets the real pass reject ~94 % of the groups from its first tile on instead of
// accepting everything until every lane's buffer has filled and been compacted a few times.
__global__ void __launch_bounds__(kSelThreads) seed_threshold_kernel(SelectParams p, u32* gthr) {
    extern __shared__ __align__(16) unsigned char smem[];
    u64* key_a = reinterpret_cast<u64*>(smem);
    u32* id_a = reinterpret_cast<u32*>(key_a + p.cap_a);
    __shared__ LzSelector sa;
    const int tid = threadIdx.x;
    const int q = blockIdx.x, qt = q / BM, ql = q % BM;
    if (tid == 0) lz_sel_init(&sa, key_a, id_a, p.cap_a, p.kp);
    __syncthreads();
    const int64_t flat = (int64_t)p.n_slices * (BN / kSub);                   // one tile = BN / kSub groups per slice
    for (int64_t base = 0; base < flat; base += 2 * kSelThreads) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t i = base + h * kSelThreads + tid;
            if (i < flat) {
                const int sl = (int)(i / (BN / kSub)), j = (int)(i % (BN / kSub));
                const size_t slot = ((size_t)sl * p.n_qtiles_pad + qt) * BM + ql;
                if (j < p.cand_cnt[slot]) {
                    const u64 kk = p.cand[slot * p.cap + j];
                    lz_sel_push(&sa, kk >> 32, 0xffffffffu - (u32)kk);
                }
            }
        }
        lz_sel_maybe_compact<kSelThreads>(&sa, 2 * kSelThreads);
    }
    __syncthreads();
    lz_sel_compact<kSelThreads>(&sa);
    if (tid == 0 && sa.count >= p.kp) gthr[q] = (u32)key_a[p.kp - 1];
}

// ---- two-level search: grouping the (query, probed shard) pairs by shard -----------------------------------------------
// Level 1 leaves probe[q][j] = the j-th best shard of query q.  The member pass runs the same GEMM kernel over a work
// list: the queries that probe shard s are copied next to each other into a PERMUTED query array (padded to whole
// q-tiles of 128 per shard), and an item is (one such q-tile, one slice of the shard's rows).  Everything below is
// built on the device, so the search never waits for the host.
struct GroupParams {
    const int64_t* probe; int n_queries, nprobe, n_shards;
    const int64_t* shard_off;
    int32_t* cnt;         // [n_shards] pairs per shard, then reused as the scatter cursor
    int32_t* qt_base;     // [n_shards] first permuted q-tile of the shard
    int32_t* n_qt;        // [n_shards]
    int32_t* nsl;         // [n_shards] slices per q-tile
    int32_t* item_base;   // [n_shards]
    int32_t* pair_row;    // [n_queries * nprobe]
    int32_t* qmap;        // [max_ptiles * BM], pre-set to -1
    int64_t* range_lo;    // [n_queries * nprobe] the pair's row range (for the exact fallback)
    int64_t* range_hi;
    int4* items; int* n_items; int max_items, max_ptiles, grid, max_nsl, per_cta;
    int4* sample_items;   // [max_ptiles] one per permuted q-tile: the first sample_tiles rows tiles of its shard
    int* n_sample; int sample_tiles;
};

__global__ void __launch_bounds__(256) group_count_kernel(GroupParams g) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= g.n_queries * g.nprobe) return;
    const int64_t sh = g.probe[i];
    const bool ok = sh >= 0 && sh < g.n_shards;
    g.range_lo[i] = ok ? g.shard_off[sh] : 0;
    g.range_hi[i] = ok ? g.shard_off[sh + 1] : 0;
    if (ok) atomicAdd(&g.cnt[sh], 1);
}

// One CTA: per shard the number of q-tiles and row tiles, the tiles-per-item T that makes ~3 items per CTA of the GEMM
// grid (and keeps a query's candidate sources within what the exact pass indexes), the prefix sums, the item list.
__global__ void __launch_bounds__(1024) group_layout_kernel(GroupParams g) {
    __shared__ long long s_red[32];
    __shared__ int s_scan[2][32];
    __shared__ int s_T;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (g.n_shards + 1023) / 1024;
    const int s0 = min(g.n_shards, tid * per), s1 = min(g.n_shards, s0 + per);
    long long work = 0;
    int max_rt = 0;
    for (int s = s0; s < s1; ++s) {
        const int nq = (g.cnt[s] + BM - 1) / BM;
        const int rt = (int)((g.shard_off[s + 1] - g.shard_off[s] + BN - 1) / BN);
        g.n_qt[s] = nq;
        work += (long long)nq * rt;
        if (nq > 0) max_rt = max(max_rt, rt);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        work += __shfl_xor_sync(0xffffffffu, work, o);
        max_rt = max(max_rt, __shfl_xor_sync(0xffffffffu, max_rt, o));
    }
    if (lane == 0) { s_red[warp] = work; s_scan[0][warp] = max_rt; }
    __syncthreads();
    if (tid == 0) {
        long long w = 0;
        int mr = 0;
        for (int i = 0; i < 32; ++i) { w += s_red[i]; mr = max(mr, s_scan[0][i]); }
        long long T = (w + (long long)g.per_cta * g.grid - 1) / ((long long)g.per_cta * g.grid);   // ~per_cta items per CTA
        const long long t_min = (mr + g.max_nsl - 1) / g.max_nsl;            // at most max_nsl slices per q-tile
        if (T < t_min) T = t_min;
        if (T < 1) T = 1;
        s_T = (int)T;
    }
    __syncthreads();
    const int T = s_T;
    int qt_sum = 0, it_sum = 0;
    for (int s = s0; s < s1; ++s) {
        const int rt = (int)((g.shard_off[s + 1] - g.shard_off[s] + BN - 1) / BN);
        const int n = g.n_qt[s] > 0 ? (rt + T - 1) / T : 0;
        g.nsl[s] = n;
        qt_sum += g.n_qt[s];
        it_sum += g.n_qt[s] * n;
    }
    int qi = qt_sum, ii = it_sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int a = __shfl_up_sync(0xffffffffu, qi, o), b = __shfl_up_sync(0xffffffffu, ii, o);
        if (lane >= o) { qi += a; ii += b; }
    }
    __syncthreads();
    if (lane == 31) { s_scan[0][warp] = qi; s_scan[1][warp] = ii; }
    __syncthreads();
    int q_run = qi - qt_sum, i_run = ii - it_sum;
    for (int w = 0; w < warp; ++w) { q_run += s_scan[0][w]; i_run += s_scan[1][w]; }
    for (int s = s0; s < s1; ++s) {
        g.qt_base[s] = q_run;
        g.item_base[s] = i_run;
        const int64_t lo = g.shard_off[s], hi = g.shard_off[s + 1];
        for (int qt = 0; qt < g.n_qt[s]; ++qt) {
            if (q_run + qt < g.max_ptiles)
                g.sample_items[q_run + qt] = make_int4(q_run + qt, (int)lo, (int)min(hi, lo + (int64_t)g.sample_tiles * BN), s);
        }
        for (int qt = 0; qt < g.n_qt[s]; ++qt)
            for (int sl = 0; sl < g.nsl[s]; ++sl) {
                const int item = i_run + qt * g.nsl[s] + sl;
                const int64_t rb = lo + (int64_t)sl * T * BN;
                const int64_t re = min(hi, rb + (int64_t)T * BN);
                if (item < g.max_items) g.items[item] = make_int4(q_run + qt, (int)rb, (int)re, s);
            }
        q_run += g.n_qt[s];
        i_run += g.n_qt[s] * g.nsl[s];
        g.cnt[s] = 0;                                                       // becomes the scatter cursor
    }
    if (tid == 1023) {                                                      // the last thread holds the totals
        *g.n_items = min(i_run, g.max_items);
        *g.n_sample = min(q_run, g.max_ptiles);
    }
}

__global__ void __launch_bounds__(256) group_scatter_kernel(GroupParams g) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= g.n_queries * g.nprobe) return;
    const int64_t sh = g.probe[i];
    int row = -1;
    if (sh >= 0 && sh < g.n_shards) {
        row = g.qt_base[sh] * BM + atomicAdd(&g.cnt[sh], 1);
        if (row < g.max_ptiles * BM) g.qmap[row] = i / g.nprobe; else row = -1;
    }
    g.pair_row[i] = row;
}

// one warp per row of the permuted query array: copy the real query's fp16 row (zeros for padding rows)
__global__ void __launch_bounds__(256) group_gather_kernel(const __half* __restrict__ q16, int dim_s, const int32_t* __restrict__ qmap,
                                                           int n_rows, __half* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= n_rows) return;
    const int q = qmap[r];
    const uint4* src = reinterpret_cast<const uint4*>(q16 + (size_t)(q < 0 ? 0 : q) * dim_s);
    uint4* dst = reinterpret_cast<uint4*>(out + (size_t)r * dim_s);
    for (int e = lane; e < dim_s / 8; e += 32) dst[e] = q >= 0 ? src[e] : make_uint4(0u, 0u, 0u, 0u);
}

// The same seeding with one WARP per query and the keys in registers: 64 sampled tiles x 32 groups = 2048 coarse keys,
// 64 per lane, and the kp-th largest found by bisection on the order-preserving 32-bit key (a compare per key and a
// warp reduce-add per bit, starting at the first bit in which the largest and smallest key differ).  The CTA-per-query
// selector above took 0.11 ms per batch of 1024 — a fixed cost that does not shrink when the rows are split over more
// GPUs; this form takes about a tenth of that.
constexpr int kSeedTiles = 64;
__global__ void __launch_bounds__(256) seed_threshold_warp_kernel(SelectParams p, u32* gthr, int n_queries) {
    constexpr int kPer = BN / kSub;                                          // keys per sampled tile (32)
    constexpr int kMine = kSeedTiles * kPer / 32;                            // keys per lane (64)
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (q >= n_queries) return;
    const int qt = q / BM, ql = q % BM;
    u32 key[kMine];
    // lane l holds entry (l % kPer) of tiles (l / kPer) * (kSeedTiles / (32 / kPer)) ... : any fixed assignment works; this one
    // makes a warp's loads of one tile contiguous
#pragma unroll
    for (int i = 0; i < kMine; ++i) {
        const int flat = i * 32 + lane;                                      // 0 .. 2047
        const int sl = flat / kPer, j = flat % kPer;
        const size_t slot = ((size_t)sl * p.n_qtiles_pad + qt) * BM + ql;
        key[i] = j < p.cand_cnt[slot] ? (u32)(p.cand[slot * p.cap + j] >> 32) : 0u;
    }
    u32 mx = 0u, mn = 0xffffffffu;
    int live = 0;
#pragma unroll
    for (int i = 0; i < kMine; ++i) {
        mx = max(mx, key[i]);
        mn = min(mn, key[i] != 0u ? key[i] : 0xffffffffu);
        live += key[i] != 0u;
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    mn = __reduce_min_sync(0xffffffffu, mn);
    live = __reduce_add_sync(0xffffffffu, live);
    if (live < p.kp) return;                                                 // too few sampled groups to rank: no seed
    const u32 diff = mx ^ mn;
    const int top = diff ? 31 - __clz(diff) : -1;
    u32 prefix = top >= 0 ? (mx & ~((2u << top) - 1u)) : mx;
#pragma unroll 1
    for (int bit = top; bit >= 0; --bit) {
        const u32 cand = prefix | (1u << bit);
        int c = 0;
#pragma unroll
        for (int i = 0; i < kMine; ++i) c += key[i] >= cand;
        c = __reduce_add_sync(0xffffffffu, c);
        if (c >= p.kp) prefix = cand;
    }
    if (lane == 0) gthr[q] = prefix;                                         // the kp-th largest sampled key
}

// Seeding for the grouped (two-level) search.  Its items are short — a q-tile against a few thousand rows of one shard —
// so without a starting threshold every lane accepts everything, overflows its buffer once per item and compacts it
// again at the end (32 warp-wide selections per warp and item: the member GEMM then runs at 40 % of the flat one's
// speed), and the exact pass receives kp groups per item.  A sampling pass over the first sample_tiles rows tiles of
// every (q-tile, shard) leaves its candidates as pushed (slot = the permuted query row); here a warp gathers a real
// query's sampled keys over its probed shards and bisects for the kp-th largest, exactly as the flat seeding does.  The
// kp-th best of a subset of the probed rows cannot exceed the kp-th best of all of them, so the seed is a valid bound.
__global__ void __launch_bounds__(256) seed_threshold_grouped_kernel(const u64* __restrict__ cand, const int32_t* __restrict__ cand_cnt,
                                                                     int cap, int kp, const int32_t* __restrict__ pair_row, int nprobe,
                                                                     int sample_tiles, u32* gthr, int n_queries) {
    constexpr int kMine = 64;                                                // keys per lane: 2048 sampled groups per query at most
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (q >= n_queries) return;
    const int per = sample_tiles * (BN / kSub);                              // keys a sampled slot can hold
    const int np = min(nprobe, (kMine * 32) / per);
    u32 key[kMine];
#pragma unroll
    for (int i = 0; i < kMine; ++i) {
        const int flat = i * 32 + lane;
        const int j = flat / per, e = flat - j * per;
        u32 kv = 0u;
        if (j < np) {
            const int row = pair_row[(size_t)q * nprobe + j];
            if (row >= 0 && e < cand_cnt[row]) kv = (u32)(cand[(size_t)row * cap + e] >> 32);
        }
        key[i] = kv;
    }
    u32 mx = 0u, mn = 0xffffffffu;
    int live = 0;
#pragma unroll
    for (int i = 0; i < kMine; ++i) {
        mx = max(mx, key[i]);
        mn = min(mn, key[i] != 0u ? key[i] : 0xffffffffu);
        live += key[i] != 0u;
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    mn = __reduce_min_sync(0xffffffffu, mn);
    live = __reduce_add_sync(0xffffffffu, live);
    if (live < kp) return;                                                   // too few sampled groups to rank: no seed
    const u32 diff = mx ^ mn;
    const int top = diff ? 31 - __clz(diff) : -1;
    u32 prefix = top >= 0 ? (mx & ~((2u << top) - 1u)) : mx;
#pragma unroll 1
    for (int bit = top; bit >= 0; --bit) {
        const u32 c0 = prefix | (1u << bit);
        int c = 0;
#pragma unroll
        for (int i = 0; i < kMine; ++i) c += key[i] >= c0;
        c = __reduce_add_sync(0xffffffffu, c);
        if (c >= kp) prefix = c0;
    }
    if (lane == 0) atomicMax(gthr + q, prefix);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
    return fn;
}

// fp16 [n_rows, dim] row-major, box = 64 elements x box_rows rows, 128-byte swizzle, zero fill out of bounds
int make_tmap(CUtensorMap* m, const void* base, int64_t n_rows, int dim, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) { lz_set_error("cuTensorMapEncodeTiled is not available from the driver"); return LZ_ERR_CUDA; }
    cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)n_rows};
    cuuint64_t gstride[1] = {(cuuint64_t)dim * 2};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { lz_set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return LZ_ERR_CUDA; }
    return LZ_OK;
}

// The rows tile of the coarse pass is fetched by TMA multicast, and TMA multicast loads here read only this
// GPU's memory (on B200 a TMA multicast load from another GPU's memory hung the GPU).  So the shadow must be
// plain device memory allocated on the CURRENT device: peer, IPC-imported, managed and host-mapped pointers
// are rejected before any tensor map is built over them.
int require_local_device_memory(const void* p, const char* what) {
    int dev = -1;
    LZ_CUDA(cudaGetDevice(&dev));
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        lz_set_error("%s is not a CUDA allocation", what);
        return LZ_ERR_ARG;
    }
    if (at.type != cudaMemoryTypeDevice || at.device != dev) {
        lz_set_error("%s must be device memory allocated on the current GPU (device %d); got memory type %d on device %d",
                     what, dev, (int)at.type, at.device);
        return LZ_ERR_ARG;
    }
    return LZ_OK;
}

struct Plan {
    int n_qtiles, cs, n_qgroups, n_tiles, n_slices, tiles_per_slice, kp, cap, fb_chunks, grid;
};

int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v ? atoi(v) : dflt;
}

// cs = CTAs per cluster sharing each rows tile by TMA multicast (one q-tile each).  Measured at 10M x 768,
// batch 1024 on B200 (ms per batch): cs=1 13.8, cs=2 12.0 — so the default is 2; LZ_CLUSTER=1 overrides.
constexpr size_t kGemmSmem = 1024 + (size_t)196608 + (2 * 4 + 4) * 8 + 16;   // 4 x 48 KiB of stages

// How many clusters of `cs` CTAs the device can hold at once.  Clusters are placed inside one GPC, and GPCs do
// not all have a multiple of `cs` usable SMs, so this can be less than SMs / cs — and a persistent grid with more
// clusters than that runs its surplus clusters as a second wave (measured: cs = 4 took 20.1 ms instead of ~12).
int active_clusters(int cs) {
    static int cache[16][3];                                  // [device][cs]
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16 || cs < 1 || cs > 2) return 0;
    if (cache[dev][cs] > 0) return cache[dev][cs];
    const int sms = lz_sm_count();
    int n = sms / cs;
    if (cs > 1) {
        auto kern = gemm_topk_kernel<8>;
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGemmSmem);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(sms / cs * cs));
        cfg.blockDim = dim3(kGemmThreads);
        cfg.dynamicSmemBytes = kGemmSmem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        int got = 0;
        if (cudaOccupancyMaxActiveClusters(&got, kern, &cfg) == cudaSuccess && got > 0 && got < n) n = got;
        else cudaGetLastError();
    }
    if (n < 1) n = 1;
    cache[dev][cs] = n;
    return n;
}

int make_plan(int64_t n_rows, int n_queries, int k, Plan* pl) {
    const int sms = lz_sm_count();
    if (sms <= 0) return LZ_ERR_DEVICE;
    pl->n_qtiles = (n_queries + BM - 1) / BM;
    int cs = env_int("LZ_CLUSTER", 2);
    if (cs != 1 && cs != 2) cs = 2;
    while (cs > 1 && cs / 2 >= pl->n_qtiles) cs /= 2;             // no larger than the q-tiles there are
    pl->cs = cs;
    pl->n_qgroups = (pl->n_qtiles + cs - 1) / cs;
    pl->n_tiles = (int)((n_rows + BN - 1) / BN);
    const int n_clusters = active_clusters(cs) > 0 ? active_clusters(cs) : 1;
    int want = (2 * n_clusters) / pl->n_qgroups;
    if (want < 1) want = 1;
    if (want > pl->n_tiles) want = pl->n_tiles;
    if (want < 1) want = 1;
    pl->tiles_per_slice = (pl->n_tiles + want - 1) / want;
    if (pl->tiles_per_slice < 1) pl->tiles_per_slice = 1;
    pl->n_slices = (pl->n_tiles + pl->tiles_per_slice - 1) / pl->tiles_per_slice;
    if (pl->n_slices < 1) pl->n_slices = 1;
    pl->kp = k <= 3 ? 32 : (k <= 16 ? 128 : (k <= 48 ? 256 : 512));
    if (const int kp_env = env_int("LZ_KP", 0)) { if (kp_env >= k && (kp_env == 128 || kp_env == 256 || kp_env == 512)) pl->kp = kp_env; }   // measurements
    // a table this small (centroids) has at most 2*128 candidate groups per query: keep them all and never
    // compact — compaction per item would otherwise dominate when the query side is the large one
    if (n_rows <= 2 * 128 * kSub && pl->kp < 128) pl->kp = 128;
    pl->cap = 2 * pl->kp;
    int64_t fb = (n_rows + 255) / 256;
    pl->fb_chunks = (int)(fb < sms ? (fb < 1 ? 1 : fb) : sms);
    const int items = pl->n_qgroups * pl->n_slices;
    pl->grid = (items < n_clusters ? items : n_clusters) * cs;
    return LZ_OK;
}

struct ScratchLayout {
    size_t q16, eps, gtag, gthr, fb_list, fb_count, cand, cand_cnt, cand_thr, certified, fallback, total;
};

ScratchLayout layout(const Plan& pl, int n_queries, int k, int dim, bool with_q16 = true) {
    ScratchLayout L;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t at = off; off = (off + bytes + 255) & ~(size_t)255; return at; };
    const size_t items = (size_t)pl.n_qgroups * pl.cs * (pl.n_slices > 64 ? pl.n_slices : 64);
    L.q16 = take(with_q16 ? (size_t)((n_queries + BM - 1) / BM * BM) * (dim + 8) * 2 : 0);      // whole q-tiles (zero rows behind the queries)
    L.eps = take((size_t)n_queries * 8);
    L.gtag = take((size_t)n_queries * 4);
    L.gthr = take((size_t)n_queries * 4);
    L.fb_list = take((size_t)n_queries * 4);
    L.fb_count = take(4);
    L.cand = take(items * BM * pl.cap * 8);
    L.cand_cnt = take(items * BM * 4);
    L.cand_thr = take(items * BM * 4);
    L.certified = take((size_t)n_queries * 4);
    L.fallback = take(lz_search_exact_scratch_bytes(n_queries, k, pl.fb_chunks));
    L.total = off;
    return L;
}

}  // namespace

#ifdef LZ_STATS
extern "C" int lz_debug_stats(unsigned long long* host_out, int reset) {
    LZ_CUDA(cudaDeviceSynchronize());
    LZ_CUDA(cudaMemcpyFromSymbol(host_out, g_stats, sizeof(unsigned long long) * 8));
    unsigned long long ss[4];
    LZ_CUDA(cudaMemcpyFromSymbol(ss, g_sel_stats, sizeof(ss)));
    if (ss[2]) fprintf(stderr, "[lz stats] exact pass: %.1f groups merged, %.1f kept by the prune, per query (%llu queries)\n",
                       (double)ss[0] / ss[2], (double)ss[1] / ss[2], ss[2]);
    if (reset) { unsigned long long z[8] = {0}; LZ_CUDA(cudaMemcpyToSymbol(g_stats, z, sizeof(z))); LZ_CUDA(cudaMemcpyToSymbol(g_sel_stats, z, sizeof(ss))); }
    return LZ_OK;
}
#endif

extern "C" int lz_rows_to_shadow(const float* rows, int64_t n, int dim, int64_t ld, const double* norms, int metric,
                                 void* out_shadow, float* row_err, void* stream) {
    if (n <= 0) return LZ_OK;
    LZ_REQUIRE(rows && norms && out_shadow, "lz_rows_to_shadow: NULL pointer");
    LZ_REQUIRE(dim > 0 && dim % 8 == 0 && ld >= dim && ld % 4 == 0, "lz_rows_to_shadow: dim=%d must be a multiple of 8", dim);
    LZ_REQUIRE((((uintptr_t)rows | (uintptr_t)out_shadow) & 15) == 0, "lz_rows_to_shadow: pointers must be 16-byte aligned");
    const int sms = lz_sm_count();
    if (sms <= 0) return LZ_ERR_DEVICE;
    int64_t blocks = (n + 7) / 8;
    if (blocks > (int64_t)sms * 16) blocks = (int64_t)sms * 16;
    LZ_REQUIRE(metric == LZ_COSINE || metric == LZ_L2, "lz_rows_to_shadow: unknown metric %d", metric);
    rows_to_shadow_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(rows, n, dim, ld, norms, metric,
                                                                              (__half*)out_shadow, row_err);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

extern "C" size_t lz_search_batched_scratch_bytes(int64_t n_rows, int n_queries, int k, int dim) {
    if (n_rows <= 0 || n_queries <= 0 || k <= 0) return 0;
    Plan pl;
    if (make_plan(n_rows, n_queries, k, &pl)) return 0;
    return layout(pl, n_queries, k, dim).total;
}

namespace {
// eps[q] for queries whose fp16 shadow already exists (rows of a resident store used as the query side): the
// deterministic margin from the row-side bound and the query's own measured rounding-error norm.
__global__ void __launch_bounds__(256) queries_eps_kernel(const float* __restrict__ q_err, int n, int dim, double det_row_err,
                                                          double* __restrict__ eps, int32_t* __restrict__ gtag, u32* __restrict__ gthr,
                                                          int32_t* __restrict__ fb_count, int want_tag) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i == 0) *fb_count = 0;
    if (i >= n) return;
    const double dq = (double)q_err[i];
    const double ev = det_row_err + dq + det_row_err * dq + (double)dim * 1.1920928955078125e-7 * 1.01 + 1e-9;
    eps[i] = ev == ev ? ev : INFINITY;
    gtag[i] = want_tag;
    gthr[i] = 0u;
}

int search_batched_core(const void* shadow, const float* rows, int64_t ld, int dim, const double* norms,
                        const int32_t* tags, int want_tag, int64_t n_rows_total, int64_t row_lo,
                        int64_t row_hi, const float* queries, int64_t q_ld, int n_queries, int k,
                        double margin_sigmas, int metric, double max_row_norm, double det_row_err,
                        int64_t id_base, void* scratch, size_t scratch_bytes,
                        int64_t* out_ids, double* out_scores, int32_t* out_counts, int32_t* out_certified,
                        void* stream, const void* query_shadow, const float* query_err);
}  // namespace

extern "C" int lz_search_batched(const void* shadow, const float* rows, int64_t ld, int dim, const double* norms,
                                 const int32_t* tags, int want_tag, int64_t n_rows_total, int64_t row_lo,
                                 int64_t row_hi, const float* queries, int64_t q_ld, int n_queries, int k,
                                 double margin_sigmas, int metric, double max_row_norm, double det_row_err,
                                 int64_t id_base, void* scratch, size_t scratch_bytes,
                                 int64_t* out_ids, double* out_scores, int32_t* out_counts, int32_t* out_certified,
                                 void* stream) {
    return search_batched_core(shadow, rows, ld, dim, norms, tags, want_tag, n_rows_total, row_lo, row_hi, queries, q_ld,
                               n_queries, k, margin_sigmas, metric, max_row_norm, det_row_err, id_base, scratch, scratch_bytes,
                               out_ids, out_scores, out_counts, out_certified, stream, nullptr, nullptr);
}

extern "C" size_t lz_assign_nearest_scratch_bytes(int n_table, int64_t n_rows, int dim, int have_row_shadow) {
    if (n_table <= 0 || n_rows <= 0 || n_rows > 0x7fffffffLL) return 0;
    Plan pl;
    if (make_plan(n_table, (int)n_rows, 1, &pl)) return 0;
    return layout(pl, (int)n_rows, 1, dim, have_row_shadow == 0).total;
}

extern "C" int lz_assign_nearest(const void* table_shadow, const float* table, int64_t table_ld, int dim,
                                 const double* table_norms, int n_table, double table_det_row_err,
                                 const float* rows, int64_t ld, int64_t n_rows, const void* row_shadow, const float* row_err,
                                 double margin_sigmas, void* scratch, size_t scratch_bytes, int64_t* out_idx,
                                 double* out_cos, int32_t* out_counts, int32_t* out_certified, void* stream) {
    LZ_REQUIRE(n_rows >= 0 && n_rows <= 0x7fffffffLL, "lz_assign_nearest: n_rows=%lld", (long long)n_rows);
    LZ_REQUIRE((row_shadow == nullptr) == (row_err == nullptr), "lz_assign_nearest: row_shadow and row_err come together");
    LZ_REQUIRE(row_shadow == nullptr || table_det_row_err >= 0.0,
               "lz_assign_nearest: a resident row shadow needs the deterministic margin (table_det_row_err >= 0)");
    return search_batched_core(table_shadow, table, table_ld, dim, table_norms, nullptr, -1, n_table, 0, n_table, rows, ld,
                               (int)n_rows, 1, margin_sigmas, LZ_COSINE, 0.0, table_det_row_err, 0, scratch, scratch_bytes,
                               out_idx, out_cos, out_counts, out_certified, stream, row_shadow, row_err);
}

namespace {
int search_batched_core(const void* shadow, const float* rows, int64_t ld, int dim, const double* norms,
                        const int32_t* tags, int want_tag, int64_t n_rows_total, int64_t row_lo,
                        int64_t row_hi, const float* queries, int64_t q_ld, int n_queries, int k,
                        double margin_sigmas, int metric, double max_row_norm, double det_row_err,
                        int64_t id_base, void* scratch, size_t scratch_bytes,
                        int64_t* out_ids, double* out_scores, int32_t* out_counts, int32_t* out_certified,
                        void* stream, const void* query_shadow, const float* query_err) {
    if (n_queries <= 0) return LZ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    LZ_REQUIRE(shadow && rows && norms && queries && scratch && out_ids && out_scores && out_counts,
               "lz_search_batched: NULL pointer");
    LZ_REQUIRE(dim > 0 && dim % 8 == 0 && ld >= dim && ld % 4 == 0, "lz_search_batched: dim=%d must be a multiple of 8", dim);
    LZ_REQUIRE(q_ld >= dim && q_ld % 4 == 0 && ((uintptr_t)queries & 15) == 0,
               "lz_search_batched: queries must be 16-byte aligned with q_ld >= dim, q_ld %% 4 == 0");
    LZ_REQUIRE((((uintptr_t)shadow | (uintptr_t)rows | (uintptr_t)scratch) & 15) == 0,
               "lz_search_batched: pointers must be 16-byte aligned");
    LZ_REQUIRE(k >= 1 && k <= LZ_BATCHED_MAX_K, "lz_search_batched: k=%d outside 1..%d", k, LZ_BATCHED_MAX_K);
    LZ_REQUIRE(0 <= row_lo && row_lo <= row_hi && row_hi <= n_rows_total && n_rows_total <= 0x7fffff00LL,
               "lz_search_batched: bad row range [%lld, %lld) of %lld", (long long)row_lo, (long long)row_hi,
               (long long)n_rows_total);
    LZ_REQUIRE(margin_sigmas > 0.0, "lz_search_batched: margin_sigmas must be positive");
    LZ_REQUIRE(metric == LZ_COSINE || metric == LZ_L2, "lz_search_batched: unknown metric %d", metric);
    LZ_REQUIRE(metric == LZ_COSINE || max_row_norm >= 0.0, "lz_search_batched: LZ_L2 needs max_row_norm >= 0");
    if (int rc = require_local_device_memory(shadow, "lz_search_batched: shadow")) return rc;
    const int dim_s = metric == LZ_L2 ? dim + 8 : dim;         // columns of the shadow (and K of the contraction)
    Plan pl;
    if (int rc = make_plan(row_hi - row_lo, n_queries, k, &pl)) return rc;
    const ScratchLayout L = layout(pl, n_queries, k, dim, query_shadow == nullptr);
    if (scratch_bytes < L.total) {
        lz_set_error("lz_search_batched: scratch has %zu bytes, needs %zu", scratch_bytes, L.total);
        return LZ_ERR_SCRATCH;
    }
    unsigned char* s = (unsigned char*)scratch;
    __half* q16 = (__half*)(s + L.q16);
    int q_rows = n_queries;                        // rows of the query-side tensor map
    double* eps = (double*)(s + L.eps);
    int32_t* cert = out_certified ? out_certified : (int32_t*)(s + L.certified);

    int32_t* gtag = (int32_t*)(s + L.gtag);
    u32* gthr = (u32*)(s + L.gthr);
    int32_t* fb_list = (int32_t*)(s + L.fb_list);
    int32_t* fb_count = (int32_t*)(s + L.fb_count);
    if (query_shadow != nullptr) {
        // the query side already has its fp16 shadow (rows of a resident store): no conversion pass over the queries
        if (int rc = require_local_device_memory(query_shadow, "lz_assign_nearest: row_shadow")) return rc;
        q16 = (__half*)const_cast<void*>(query_shadow);
        queries_eps_kernel<<<(n_queries + 255) / 256, 256, 0, st>>>(query_err, n_queries, dim, det_row_err, eps, gtag, gthr,
                                                                    fb_count, want_tag);
    } else {
        queries_to_shadow_kernel<<<(n_queries + 7) / 8, 256, 0, st>>>(queries, n_queries, dim, q_ld, q16, eps, gtag, gthr,
                                                                    fb_count, want_tag, margin_sigmas, metric, max_row_norm,
                                                                    det_row_err);
        // the last q-tile is filled up with zero rows in memory, so that no TMA box of the query side reaches out of bounds
        q_rows = (n_queries + BM - 1) / BM * BM;
        if (q_rows > n_queries && env_int("LZ_QPAD", 1))
            LZ_CUDA(cudaMemsetAsync(q16 + (size_t)n_queries * dim_s, 0, (size_t)(q_rows - n_queries) * dim_s * 2, st));
        else
            q_rows = n_queries;
    }
    LZ_CUDA(cudaGetLastError());

    // top-1 against a tiny table (shard reassignment): trailing threshold in the coarse pass, bucketed exact pass
    const size_t rows_smem = (size_t)kSub * dim * 4;
    const bool tiny_top1 = k == 1 && tags == nullptr && pl.n_slices == 1 && row_hi - row_lo <= (int64_t)kBucketGroups * kSub &&
                           rows_smem <= 160 * 1024 && n_queries >= 4 * kBucketQ;
    // (Single-row candidates for every k were tried to shrink the exact pass and measured slower overall — 18.0 vs
    // 15.6 ms at k = 100, 14.6 vs 12.0 ms at k = 10 at 10M x 768: more entries pass the group-level seeds and the
    // epilogue compacts more often, which stalls the MMA pipeline.  The top-1 tiny-table mode is the one user.)
    if (row_hi > row_lo) {
        CUtensorMap tq, tx;
        if (int rc = make_tmap(&tq, q16, q_rows, dim_s, BM)) return rc;
        if (int rc = make_tmap(&tx, shadow, n_rows_total, dim_s, BN / pl.cs)) return rc;
        GemmParams g = {};
        g.n_queries = n_queries; g.n_qtiles = pl.n_qtiles; g.n_slices = pl.n_slices;
        g.tiles_per_slice = pl.tiles_per_slice; g.n_tiles = pl.n_tiles; g.kblocks = (dim_s + BK - 1) / BK;
        g.cs = pl.cs; g.n_qgroups = pl.n_qgroups; g.tile_mul = 1;
        g.row_lo = row_lo; g.row_hi = row_hi; g.kp = pl.kp; g.cap = pl.cap;
        g.tags = tags; g.want_tag = want_tag; g.gthr = gthr; g.top1_eps = tiny_top1 ? eps : nullptr;
        g.cand = (u64*)(s + L.cand); g.cand_cnt = (int32_t*)(s + L.cand_cnt); g.cand_thr = (float*)(s + L.cand_thr);
        g.debug_skip = env_int("LZ_DEBUG_SKIP", 0);
        const size_t smem = kGemmSmem;
        auto kern = pl.kp == 32 ? gemm_topk_kernel<2> : pl.kp == 128 ? gemm_topk_kernel<8>
                    : pl.kp == 256 ? gemm_topk_kernel<16> : gemm_topk_kernel<32>;
        LZ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaLaunchConfig_t cfg = {};
        cfg.blockDim = dim3(kGemmThreads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)pl.cs;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        // sampling pass: kSampleTiles tiles spread evenly over the range seed every query's threshold
        // (skipped when the table is small, or when kp exceeds what the sample can rank)
        constexpr int kSampleTiles = 64;
        if (pl.n_tiles >= 8 * kSampleTiles && kSampleTiles * (BN / kSub) >= 4 * pl.kp && env_int("LZ_SEED", 1) != 0) {
            GemmParams gs = g;
            gs.n_tiles = kSampleTiles; gs.n_slices = kSampleTiles; gs.tiles_per_slice = 1;
            gs.tile_mul = pl.n_tiles / kSampleTiles;
            const int n_clusters = pl.grid / pl.cs;
            const int items = pl.n_qgroups * kSampleTiles;
            cfg.gridDim = dim3((unsigned)((items < n_clusters ? items : n_clusters) * pl.cs));
            LZ_CUDA(cudaLaunchKernelEx(&cfg, kern, tq, tx, gs));
            SelectParams ss = {};
            ss.cand = (const u64*)(s + L.cand); ss.cand_cnt = (const int32_t*)(s + L.cand_cnt);
            ss.n_qtiles_pad = pl.n_qgroups * pl.cs; ss.n_slices = kSampleTiles; ss.kp = pl.kp; ss.cap = pl.cap;
            ss.cap_a = lz_next_pow2(2 * pl.kp + 2 * kSelThreads);     // room for kp survivors + a round of pushes + slack, so not every round compacts
            const size_t ssm = (size_t)ss.cap_a * 12;
            if (kSampleTiles == kSeedTiles && env_int("LZ_SEED_OLD", 0) == 0) {
                seed_threshold_warp_kernel<<<(n_queries + 7) / 8, 256, 0, st>>>(ss, gthr, n_queries);
            } else {
                LZ_CUDA(cudaFuncSetAttribute(seed_threshold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
                seed_threshold_kernel<<<n_queries, kSelThreads, ssm, st>>>(ss, gthr);
            }
            LZ_CUDA(cudaGetLastError());
        }
        cfg.gridDim = dim3((unsigned)pl.grid);
        LZ_CUDA(cudaLaunchKernelEx(&cfg, kern, tq, tx, g));
    } else {
        const size_t items = (size_t)pl.n_qgroups * pl.cs * pl.n_slices;
        LZ_CUDA(cudaMemsetAsync(s + L.cand_cnt, 0, items * BM * 4, st));
        LZ_CUDA(cudaMemsetAsync(s + L.cand_thr, 0xff, items * BM * 4, st));     // NaN: ignored by fmaxf
    }

    SelectParams sp = {};
    sp.rows = rows; sp.ld = ld; sp.dim = dim; sp.norms = norms; sp.queries = queries; sp.q_ld = q_ld; sp.eps = eps;
    sp.tags = tags; sp.want_tag = want_tag; sp.row_lo = row_lo; sp.row_hi = row_hi;
    sp.cand = (const u64*)(s + L.cand); sp.cand_cnt = (const int32_t*)(s + L.cand_cnt);
    sp.cand_thr = (const float*)(s + L.cand_thr);
    sp.n_qtiles_pad = pl.n_qgroups * pl.cs; sp.n_slices = pl.n_slices; sp.kp = pl.kp; sp.cap = pl.cap; sp.k = k;
    sp.cap_a = lz_next_pow2(2 * pl.kp + 2 * kSelThreads);
    // the rescoring loop pushes at most 128 rows between two compaction checks; for small k a 256-entry selector keeps the
    // kernel's shared memory under what is left beside a resident GEMM CTA, so the exact pass of one search can run
    // while the next search's coarse pass holds the SMs (ShardedSearch lanes)
    sp.cap_b = k <= 16 ? lz_next_pow2(k + 128 + 32) : lz_next_pow2(k + 2 * kSelThreads);
    sp.id_base = id_base; sp.sub = tiny_top1 ? 1 : kSub;
    sp.out_ids = out_ids; sp.out_scores = out_scores; sp.out_counts = out_counts; sp.certified = cert;
    sp.fb_list = fb_list; sp.fb_count = fb_count; sp.gthr = gthr; sp.metric = metric;
    if (tiny_top1 && dim <= 1024 && env_int("LZ_ASSIGN_OLD", 0) == 0) {
        // per-CTA fixed part of the shared memory, then as many cached table rows as fit (at most 64)
        const size_t fixed = (size_t)kMultiQ * (8 + 4 + 4 + 2 * kMultiLive + 1) + 64;
        const int nseg0 = (dim + 127) / 128, nseg = nseg0 <= 2 ? 2 : nseg0 <= 4 ? 4 : nseg0 <= 6 ? 6 : 8;
        const size_t row_bytes = (size_t)nseg * 512 + 8;
        const size_t stage_bytes = (size_t)(kMultiThreads / 32) * 2 * nseg * 512 + 16;     // two query rows per warp
        int cache_rows = (int)((200 * 1024 - fixed - stage_bytes) / row_bytes);
        if (cache_rows > 64) cache_rows = 64;
        const size_t msmem = fixed + stage_bytes + (size_t)cache_rows * row_bytes;
        auto mk = metric == LZ_COSINE
                      ? (nseg == 2 ? assign_multi_kernel<2, LZ_COSINE> : nseg == 4 ? assign_multi_kernel<4, LZ_COSINE>
                         : nseg == 6 ? assign_multi_kernel<6, LZ_COSINE> : assign_multi_kernel<8, LZ_COSINE>)
                      : (nseg == 2 ? assign_multi_kernel<2, LZ_L2> : nseg == 4 ? assign_multi_kernel<4, LZ_L2>
                         : nseg == 6 ? assign_multi_kernel<6, LZ_L2> : assign_multi_kernel<8, LZ_L2>);
        LZ_CUDA(cudaFuncSetAttribute(mk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
        mk<<<(n_queries + kMultiQ - 1) / kMultiQ, kMultiThreads, msmem, st>>>(sp, n_queries, cache_rows);
    } else if (tiny_top1) {
        LZ_CUDA(cudaFuncSetAttribute(assign_rescore_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rows_smem));
        assign_rescore_kernel<<<(n_queries + kBucketQ - 1) / kBucketQ, kSelThreads, rows_smem, st>>>(sp, n_queries);
    } else if (k <= kSmallK) {
        select_rescore_small_kernel<<<(n_queries + kSelThreads / 32 - 1) / (kSelThreads / 32), kSelThreads, 0, st>>>(sp, n_queries);
    } else {
        const size_t ssmem = (size_t)(sp.cap_a + sp.cap_b) * 12;
        LZ_CUDA(cudaFuncSetAttribute(select_rescore_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssmem));
        select_rescore_kernel<<<n_queries, kSelThreads, ssmem, st>>>(sp);
    }
    LZ_CUDA(cudaGetLastError());

    // exact float64 scan for the queries that could not be certified (a no-op launch when all were)
    return lz_search_exact_impl(rows, ld, dim, norms, tags, row_lo, row_hi, nullptr, nullptr, tags ? gtag : nullptr, queries, q_ld,
                                n_queries, 1, k, metric, pl.fb_chunks, id_base, s + L.fallback,
                                L.total - L.fallback, out_ids, out_scores, out_counts, fb_list, fb_count, stream);
}
}  // namespace


// ---- two-level search (north_star (1); the reference's level 1 is core/memory_system.py:464-470) ---------------------
namespace {
struct TwoLevelLayout {
    size_t l1, probe_ids, probe_sc, probe_cnt, shard_i32, pair_row, qmap, range_lo, range_hi, items, n_items, sample_items, n_sample,
           q16, qperm, eps, gtag, gthr, fb_list, fb_count, cand, cand_cnt, cand_thr, certified, fallback, total;
    int max_ptiles, max_items, kp, cap, grid, fb_cps, max_nsl;
    size_t l1_bytes, fb_bytes;
    bool l1_batched;
};

int two_level_layout(int n_shards, int n_queries, int k, int nprobe, int dim, TwoLevelLayout* L) {
    const int sms = lz_sm_count();
    if (sms <= 0) return LZ_ERR_DEVICE;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t at = off; off = (off + bytes + 255) & ~(size_t)255; return at; };
    const int64_t pairs = (int64_t)n_queries * nprobe;
    L->grid = sms;
    L->max_ptiles = (int)((pairs + BM - 1) / BM + (pairs < n_shards ? pairs : n_shards));
    L->max_items = 3 * L->grid + L->max_ptiles;
    L->max_nsl = kMaxSlices / nprobe < 1 ? 1 : kMaxSlices / nprobe;
    L->kp = k <= 3 ? 32 : (k <= 16 ? 128 : (k <= 48 ? 256 : 512));
    L->cap = 2 * L->kp;
    L->l1_batched = nprobe <= LZ_BATCHED_MAX_K && n_queries >= 8;
    L->l1_bytes = L->l1_batched ? lz_search_batched_scratch_bytes(n_shards, n_queries, nprobe, dim)
                                : lz_search_exact_scratch_bytes(n_queries, nprobe, lz_search_exact_default_chunks(n_shards, (n_queries + 3) / 4));
    L->fb_cps = (2 * sms) / nprobe < 1 ? 1 : (2 * sms) / nprobe;
    L->fb_bytes = lz_search_exact_scratch_bytes(n_queries, k, nprobe * L->fb_cps);
    L->l1 = take(L->l1_bytes);
    L->probe_ids = take((size_t)pairs * 8);
    L->probe_sc = take((size_t)pairs * 8);
    L->probe_cnt = take((size_t)n_queries * 4);
    L->shard_i32 = take((size_t)n_shards * 4 * 5);                 // cnt, qt_base, n_qt, nsl, item_base
    L->pair_row = take((size_t)pairs * 4);
    L->qmap = take((size_t)L->max_ptiles * BM * 4);
    L->range_lo = take((size_t)pairs * 8);
    L->range_hi = take((size_t)pairs * 8);
    L->items = take((size_t)L->max_items * 16);
    L->n_items = take(4);
    L->sample_items = take((size_t)L->max_ptiles * 16);
    L->n_sample = take(4);
    L->q16 = take((size_t)n_queries * (dim + 8) * 2);
    L->qperm = take((size_t)L->max_ptiles * BM * (dim + 8) * 2);
    L->eps = take((size_t)n_queries * 8);
    L->gtag = take((size_t)n_queries * 4);
    L->gthr = take((size_t)n_queries * 4);
    L->fb_list = take((size_t)n_queries * 4);
    L->fb_count = take(4);
    L->cand = take((size_t)L->max_items * BM * L->cap * 8);
    L->cand_cnt = take((size_t)L->max_items * BM * 4);
    L->cand_thr = take((size_t)L->max_items * BM * 4);
    L->certified = take((size_t)n_queries * 4);
    L->fallback = take(L->fb_bytes);
    L->total = off;
    return LZ_OK;
}
}  // namespace

extern "C" size_t lz_search_two_level_scratch_bytes(int n_shards, int n_queries, int k, int nprobe, int dim) {
    if (n_shards <= 0 || n_queries <= 0 || k <= 0 || nprobe <= 0) return 0;
    if (nprobe > n_shards) nprobe = n_shards;
    TwoLevelLayout L;
    if (two_level_layout(n_shards, n_queries, k, nprobe, dim, &L)) return 0;
    return L.total;
}

extern "C" int lz_search_two_level(const void* shadow, const float* rows, int64_t ld, int dim, const double* norms,
                                   int64_t n_rows_total, const int64_t* shard_offsets, int n_shards,
                                   const void* cen_shadow, const float* centroids, int64_t cen_ld, const double* cen_norms,
                                   double cen_det_row_err, const float* queries, int64_t q_ld, int n_queries, int k,
                                   int nprobe, double margin_sigmas, int metric, double max_row_norm, double det_row_err,
                                   int64_t id_base, void* scratch, size_t scratch_bytes, int64_t* out_ids,
                                   double* out_scores, int32_t* out_counts, int32_t* out_certified, int64_t* out_probe,
                                   void* stream) {
    if (n_queries <= 0) return LZ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    LZ_REQUIRE(shadow && rows && norms && shard_offsets && cen_shadow && centroids && cen_norms && queries && scratch &&
               out_ids && out_scores && out_counts, "lz_search_two_level: NULL pointer");
    LZ_REQUIRE(dim > 0 && dim % 8 == 0 && ld >= dim && ld % 4 == 0 && cen_ld >= dim && cen_ld % 4 == 0,
               "lz_search_two_level: dim=%d must be a multiple of 8", dim);
    LZ_REQUIRE(q_ld >= dim && q_ld % 4 == 0 && ((uintptr_t)queries & 15) == 0,
               "lz_search_two_level: queries must be 16-byte aligned with q_ld >= dim, q_ld %% 4 == 0");
    LZ_REQUIRE(k >= 1 && k <= LZ_BATCHED_MAX_K, "lz_search_two_level: k=%d outside 1..%d", k, LZ_BATCHED_MAX_K);
    LZ_REQUIRE(n_shards >= 1 && nprobe >= 1 && nprobe <= LZ_MAX_K, "lz_search_two_level: n_shards=%d nprobe=%d", n_shards, nprobe);
    LZ_REQUIRE(n_rows_total >= 0 && n_rows_total <= 0x7fffff00LL, "lz_search_two_level: bad row count %lld", (long long)n_rows_total);
    LZ_REQUIRE(metric == LZ_COSINE || metric == LZ_L2, "lz_search_two_level: unknown metric %d", metric);
    LZ_REQUIRE(margin_sigmas > 0.0, "lz_search_two_level: margin_sigmas must be positive");
    if (int rc = require_local_device_memory(shadow, "lz_search_two_level: shadow")) return rc;
    if (nprobe > n_shards) nprobe = n_shards;
    TwoLevelLayout L;
    if (int rc = two_level_layout(n_shards, n_queries, k, nprobe, dim, &L)) return rc;
    if (scratch_bytes < L.total) {
        lz_set_error("lz_search_two_level: scratch has %zu bytes, needs %zu", scratch_bytes, L.total);
        return LZ_ERR_SCRATCH;
    }
    unsigned char* s = (unsigned char*)scratch;
    int64_t* probe = (int64_t*)(s + L.probe_ids);
    const int dim_s = metric == LZ_L2 ? dim + 8 : dim;

    // level 1: the nprobe best shards of every query by cosine against the shard centroids (memory_system.py:466-470)
    if (L.l1_batched) {
        if (int rc = lz_search_batched(cen_shadow, centroids, cen_ld, dim, cen_norms, nullptr, -1, n_shards, 0, n_shards, queries,
                                       q_ld, n_queries, nprobe, margin_sigmas, LZ_COSINE, 0.0, cen_det_row_err, 0, s + L.l1,
                                       L.l1_bytes, probe, (double*)(s + L.probe_sc), (int32_t*)(s + L.probe_cnt), nullptr, stream))
            return rc;
    } else {
        const int gs = n_queries >= 4 ? 4 : (n_queries >= 2 ? 2 : 1);
        if (int rc = lz_search_exact(centroids, cen_ld, dim, cen_norms, nullptr, 0, n_shards, nullptr, nullptr, nullptr, queries, q_ld,
                                     n_queries, gs, nprobe, LZ_COSINE, lz_search_exact_default_chunks(n_shards, (n_queries + gs - 1) / gs),
                                     0, s + L.l1, L.l1_bytes, probe, (double*)(s + L.probe_sc), (int32_t*)(s + L.probe_cnt), stream))
            return rc;
    }
    if (out_probe) LZ_CUDA(cudaMemcpyAsync(out_probe, probe, (size_t)n_queries * nprobe * 8, cudaMemcpyDeviceToDevice, st));

    // group the (query, shard) pairs by shard; build the permuted query array and the item list
    GroupParams g;
    g.probe = probe; g.n_queries = n_queries; g.nprobe = nprobe; g.n_shards = n_shards; g.shard_off = shard_offsets;
    int32_t* si = (int32_t*)(s + L.shard_i32);
    g.cnt = si; g.qt_base = si + n_shards; g.n_qt = si + 2 * (size_t)n_shards; g.nsl = si + 3 * (size_t)n_shards;
    g.item_base = si + 4 * (size_t)n_shards;
    g.pair_row = (int32_t*)(s + L.pair_row); g.qmap = (int32_t*)(s + L.qmap);
    g.range_lo = (int64_t*)(s + L.range_lo); g.range_hi = (int64_t*)(s + L.range_hi);
    g.items = (int4*)(s + L.items); g.n_items = (int*)(s + L.n_items);
    g.max_items = L.max_items; g.max_ptiles = L.max_ptiles; g.grid = L.grid; g.max_nsl = L.max_nsl;
    g.per_cta = env_int("LZ_TL_ITEMS_PER_CTA", 3);
    if (g.per_cta < 1 || g.per_cta > 3) g.per_cta = 3;                       // the item buffer is sized for 3
    // sampling pass for the threshold seeds: up to 64 rows tiles per query over its probed shards, never more per item
    // than a candidate buffer holds without compaction; skipped when the shards are hardly larger than the sample
    g.sample_items = (int4*)(s + L.sample_items); g.n_sample = (int*)(s + L.n_sample);
    g.sample_tiles = 64 / nprobe < 1 ? 1 : 64 / nprobe;
    if (g.sample_tiles > L.cap / (BN / kSub) - 1) g.sample_tiles = L.cap / (BN / kSub) - 1;
    // Measured: 1M x 768 / 64 shards / batch 256 (16 pairs per shard): the two extra launches cost 64 us and pay from nprobe 4
    // on — 0.598 -> 0.572 ms at 4, 0.664 -> 0.553 at 16, 0.914 -> 0.763 at 64; at nprobe 1-2 they cost more than they save.
    // 10M x 768 / 1024 shards / batch 1024 (4 pairs per shard, every q-tile nearly empty): 3.13 -> 3.60 ms at nprobe 4 — the
    // sample costs a fifth of the member pass and spares few lanes a compaction.  Hence: from nprobe 4 on and only when the
    // q-tiles are reasonably filled.  LZ_TL_SEED=0/2: never/always.
    const int seed_mode = env_int("LZ_TL_SEED", 1);
    const bool seeded = seed_mode != 0 && g.sample_tiles >= 1 && n_rows_total / n_shards >= (int64_t)4 * g.sample_tiles * BN &&
                        (seed_mode == 2 || (nprobe >= 4 && (int64_t)n_queries * nprobe >= (int64_t)8 * n_shards));
    const int pairs = n_queries * nprobe;
    LZ_CUDA(cudaMemsetAsync(g.cnt, 0, (size_t)n_shards * 4, st));
    LZ_CUDA(cudaMemsetAsync(g.qmap, 0xff, (size_t)L.max_ptiles * BM * 4, st));
    group_count_kernel<<<(pairs + 255) / 256, 256, 0, st>>>(g);
    group_layout_kernel<<<1, 1024, 0, st>>>(g);
    group_scatter_kernel<<<(pairs + 255) / 256, 256, 0, st>>>(g);
    LZ_CUDA(cudaGetLastError());

    __half* q16 = (__half*)(s + L.q16);
    __half* qperm = (__half*)(s + L.qperm);
    double* eps = (double*)(s + L.eps);
    u32* gthr = (u32*)(s + L.gthr);
    int32_t* fb_list = (int32_t*)(s + L.fb_list);
    int32_t* fb_count = (int32_t*)(s + L.fb_count);
    int32_t* cert = out_certified ? out_certified : (int32_t*)(s + L.certified);
    queries_to_shadow_kernel<<<(n_queries + 7) / 8, 256, 0, st>>>(queries, n_queries, dim, q_ld, q16, eps, (int32_t*)(s + L.gtag),
                                                                  gthr, fb_count, -1, margin_sigmas, metric, max_row_norm, det_row_err);
    const int perm_rows = L.max_ptiles * BM;
    group_gather_kernel<<<(perm_rows + 7) / 8, 256, 0, st>>>(q16, dim_s, g.qmap, perm_rows, qperm);
    LZ_CUDA(cudaGetLastError());

    // level 2: the member rows of the probed shards — the same tcgen05 GEMM + candidate epilogue, over the item list
    CUtensorMap tq, tx;
    if (int rc = make_tmap(&tq, qperm, perm_rows, dim_s, BM)) return rc;
    if (int rc = make_tmap(&tx, shadow, n_rows_total, dim_s, BN)) return rc;
    GemmParams gp = {};
    gp.n_queries = n_queries; gp.kblocks = (dim_s + BK - 1) / BK; gp.cs = 1; gp.n_qgroups = 1; gp.tile_mul = 1;
    gp.kp = L.kp; gp.cap = L.cap; gp.tags = nullptr; gp.want_tag = -1; gp.gthr = gthr; gp.top1_eps = nullptr;
    gp.cand = (u64*)(s + L.cand); gp.cand_cnt = (int32_t*)(s + L.cand_cnt); gp.cand_thr = (float*)(s + L.cand_thr);
    gp.items = g.items; gp.n_items_dev = g.n_items; gp.qmap = g.qmap;
    auto kern = L.kp == 32 ? gemm_topk_kernel<2> : L.kp == 128 ? gemm_topk_kernel<8>
                : L.kp == 256 ? gemm_topk_kernel<16> : gemm_topk_kernel<32>;
    LZ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGemmSmem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)L.grid);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = kGemmSmem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (seeded) {
        // the same kernel over the sample items (slot = permuted query row), candidates left as pushed; then one warp per
        // real query turns its sampled keys into the starting threshold of every item that scans for it
        GemmParams gs = gp;
        gs.items = g.sample_items; gs.n_items_dev = g.n_sample; gs.raw_items = 1;
        LZ_CUDA(cudaLaunchKernelEx(&cfg, kern, tq, tx, gs));
        seed_threshold_grouped_kernel<<<(n_queries + 7) / 8, 256, 0, st>>>(gp.cand, gp.cand_cnt, L.cap, L.kp, g.pair_row, nprobe,
                                                                           g.sample_tiles, gthr, n_queries);
        LZ_CUDA(cudaGetLastError());
    }
    LZ_CUDA(cudaLaunchKernelEx(&cfg, kern, tq, tx, gp));

    SelectParams sp = {};
    sp.rows = rows; sp.ld = ld; sp.dim = dim; sp.norms = norms; sp.queries = queries; sp.q_ld = q_ld; sp.eps = eps;
    sp.tags = nullptr; sp.want_tag = -1; sp.row_lo = 0; sp.row_hi = n_rows_total;
    sp.cand = gp.cand; sp.cand_cnt = gp.cand_cnt; sp.cand_thr = gp.cand_thr;
    sp.n_qtiles_pad = 0; sp.n_slices = 0; sp.kp = L.kp; sp.cap = L.cap; sp.k = k;
    sp.cap_a = lz_next_pow2(2 * L.kp + 2 * kSelThreads);
    sp.cap_b = k <= 16 ? lz_next_pow2(k + 128 + 32) : lz_next_pow2(k + 2 * kSelThreads);
    sp.id_base = id_base; sp.sub = kSub;
    sp.out_ids = out_ids; sp.out_scores = out_scores; sp.out_counts = out_counts; sp.certified = cert;
    sp.fb_list = fb_list; sp.fb_count = fb_count; sp.gthr = gthr; sp.metric = metric;
    sp.probe = probe; sp.pair_row = g.pair_row; sp.item_base = g.item_base; sp.qt_base = g.qt_base; sp.nsl = g.nsl;
    sp.shard_off = shard_offsets; sp.nprobe = nprobe; sp.n_shards = n_shards;
    const size_t ssmem = (size_t)(sp.cap_a + sp.cap_b) * 12;
    LZ_CUDA(cudaFuncSetAttribute(select_rescore_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssmem));
    select_rescore_kernel<<<n_queries, kSelThreads, ssmem, st>>>(sp);
    LZ_CUDA(cudaGetLastError());

    // queries that could not be certified: exact float64 scan of exactly their probed shards (a no-op launch when none)
    return lz_search_exact_impl(rows, ld, dim, norms, nullptr, 0, n_rows_total, g.range_lo, g.range_hi, nullptr, queries, q_ld,
                                n_queries, 1, k, metric, nprobe * L.fb_cps, id_base, s + L.fallback, L.fb_bytes, out_ids, out_scores,
                                out_counts, fb_list, fb_count, stream, nprobe);
}
