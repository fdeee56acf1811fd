// K1: exact flat search — one streaming pass over float32 rows with float64 accumulation and a fused
// per-CTA top-k, then a per-query merge of the per-CTA lists.  HBM-bound: every row is read once per
// query group with 128-bit non-allocating loads; the arithmetic (24 cvt + 24*QB DFMA per lane and
// row at D=768) stays well under B200's fp64 rate.
//
// Stands behind LanceDBStore.search_nodes (core/vector_store.py:132-140) and the cosine of
// MemorySystem._cosine_similarity (core/memory_system.py:197-203).
#include "lz_common.cuh"
#include "lz_topk.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kR = 4;                       // rows per warp per iteration
constexpr int kIters = 8;                   // iterations between two compaction checks
constexpr int kChunkRows = kWarps * kR * kIters;   // 256 rows -> at most 256 pushes per selector

// Sum over one vector of f(x_i) in the canonical order used by every kernel in this file: lane l
// owns elements seg*128 + 4l .. 4l+3, accumulates them seg by seg with fma, then a butterfly sum.
// Using one order everywhere makes bit-identical rows produce bit-identical norms and scores.
__device__ __forceinline__ double warp_sumsq(const float* __restrict__ v, int dim, int lane) {
    double acc = 0.0;
    for (int e = lane * 4; e < dim; e += 128) {
        float4 x = *reinterpret_cast<const float4*>(v + e);
        acc = fma((double)x.x, (double)x.x, acc);
        acc = fma((double)x.y, (double)x.y, acc);
        acc = fma((double)x.z, (double)x.z, acc);
        acc = fma((double)x.w, (double)x.w, acc);
    }
    return lz_warp_sum(acc);
}

__global__ void __launch_bounds__(kThreads) row_norms_kernel(const float* __restrict__ rows, int64_t n,
                                                             int dim, int64_t ld, double* __restrict__ norms) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * kWarps;
    for (int64_t r = warp; r < n; r += nwarps) {
        double s = warp_sumsq(rows + r * ld, dim, lane);
        if (lane == 0) norms[r] = sqrt(s);
    }
}

struct ScanParams {
    const float* rows; int64_t ld; int dim; const double* norms; const int32_t* tags;
    int64_t row_lo, row_hi; const int64_t* group_lo; const int64_t* group_hi; const int32_t* group_tag;
    const float* queries; int64_t q_ld; int n_queries;
    int k; int n_chunks; int cap;
    u64* part_key; u32* part_id; int32_t* part_cnt;
    const int32_t* work_list;   // optional: indices of the queries to answer (group_size 1); others untouched
    const int32_t* work_count;  // device count of work_list entries
    // fused single-query form (lz_search_one): at most 32 chunks; the LAST CTA to finish merges the chunk lists with one
    // warp and writes the answer straight to `fused_out` (host-mapped memory), then the sequence number behind it
    u32* fused_ticket;          // device counter, left at zero again by the merging CTA
    unsigned char* fused_out;   // [k int64 ids | k float64 scores | int32 count | int32 flag]
    int fused_seq; int metric; int64_t id_base;
    int want_tag_all;           // tag wanted by every group when group_tag is null (-1: any live row)
    int n_ranges;               // > 1: group g scans the n_ranges ranges [group_lo[g*n_ranges + j], group_hi[...]) (the probed
                                // shards of a two-level search); n_chunks is a multiple of n_ranges, chunk c works on range c / per
};

template <int QB, int METRIC, bool FUSED = false>
__global__ void __launch_bounds__(kThreads, 2) scan_kernel(ScanParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nseg = (p.dim + 127) >> 7;
    // Normally one (group, chunk) item per CTA.  With a work list (the batched search's fallback) a fixed
    // grid loops over (listed query, chunk) items, so an empty list costs one tiny launch.
    const int total = p.work_list ? (*p.work_count) * p.n_chunks : (int)gridDim.x;
    for (int item = blockIdx.x; item < total; item += gridDim.x) {
    const int g = p.work_list ? p.work_list[item / p.n_chunks] : item / p.n_chunks;
    const int c = item % p.n_chunks;
    const int q0 = g * QB;
    const int nq = min(QB, p.n_queries - q0);

    // shared memory carve-up
    double2* qs = reinterpret_cast<double2*>(smem);                 // [QB][nseg][2][32]
    u64* keys = reinterpret_cast<u64*>(qs + (size_t)QB * nseg * 64); // [QB][cap]
    u32* ids = reinterpret_cast<u32*>(keys + (size_t)QB * p.cap);    // [QB][cap]
    __shared__ LzSelector sel[QB];
    __shared__ double qnorm[QB];

    // queries -> shared memory as float64, zero padded to nseg*128
    for (int i = tid; i < QB * nseg * 128; i += kThreads) {
        int b = i / (nseg * 128), e = i - b * nseg * 128;
        double v = (b < nq && e < p.dim) ? (double)p.queries[(int64_t)(q0 + b) * p.q_ld + e] : 0.0;
        int seg = e >> 7, l = (e & 127) >> 2, comp = e & 3;
        double* dst = reinterpret_cast<double*>(&qs[((b * nseg + seg) * 2 + (comp >> 1)) * 32 + l]);
        dst[comp & 1] = v;
    }
    if (warp < QB) {
        double s = (warp < nq) ? warp_sumsq(p.queries + (int64_t)(q0 + warp) * p.q_ld, p.dim, lane) : 0.0;
        if (lane == 0) {
            qnorm[warp] = sqrt(s);
            lz_sel_init(&sel[warp], keys + (size_t)warp * p.cap, ids + (size_t)warp * p.cap, p.cap, p.k);
        }
    }
    __syncthreads();

    const int per_range = p.n_ranges > 1 ? p.n_chunks / p.n_ranges : p.n_chunks;       // chunks that share one range
    const size_t ri = p.n_ranges > 1 ? (size_t)g * p.n_ranges + c / per_range : (size_t)g;
    int64_t lo = p.group_lo ? p.group_lo[ri] : p.row_lo;
    int64_t hi = p.group_hi ? p.group_hi[ri] : p.row_hi;
    if (hi < lo) hi = lo;
    const int want_tag = p.group_tag ? p.group_tag[g] : p.want_tag_all;
    const int64_t len = (hi - lo + per_range - 1) / per_range;
    const int64_t cta_lo = lo + (int64_t)(c % per_range) * len;
    const int64_t cta_hi = min(hi, cta_lo + len);

    for (int64_t base = cta_lo; base < cta_hi; base += kChunkRows) {
#pragma unroll 1
        for (int it = 0; it < kIters; ++it) {
            const int64_t row0 = base + (int64_t)(it * kWarps + warp) * kR;
            if (row0 >= cta_hi) break;
            const float* xp[kR];
#pragma unroll
            for (int r = 0; r < kR; ++r) xp[r] = p.rows + min(row0 + r, cta_hi - 1) * p.ld + lane * 4;
            double my_norm = 0.0;
            int my_tag = 0;
            if (lane < kR) {
                int64_t row = min(row0 + lane, cta_hi - 1);
                if (METRIC == LZ_COSINE) my_norm = p.norms[row];
                if (p.tags) my_tag = p.tags[row];
            }
            double acc[kR][QB];
#pragma unroll
            for (int r = 0; r < kR; ++r)
#pragma unroll
                for (int b = 0; b < QB; ++b) acc[r][b] = 0.0;

#pragma unroll 2
            for (int seg = 0; seg < nseg; ++seg) {
                const bool act = seg * 128 + lane * 4 < p.dim;
                float4 x[kR];
#pragma unroll
                for (int r = 0; r < kR; ++r)
                    x[r] = act ? lz_ldg_stream(xp[r] + seg * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int b = 0; b < QB; ++b) {
                    const double2 qa = qs[((b * nseg + seg) * 2 + 0) * 32 + lane];
                    const double2 qb = qs[((b * nseg + seg) * 2 + 1) * 32 + lane];
#pragma unroll
                    for (int r = 0; r < kR; ++r) {
                        if (METRIC == LZ_COSINE) {
                            acc[r][b] = fma((double)x[r].x, qa.x, acc[r][b]);
                            acc[r][b] = fma((double)x[r].y, qa.y, acc[r][b]);
                            acc[r][b] = fma((double)x[r].z, qb.x, acc[r][b]);
                            acc[r][b] = fma((double)x[r].w, qb.y, acc[r][b]);
                        } else {
                            double d0 = (double)x[r].x - qa.x, d1 = (double)x[r].y - qa.y;
                            double d2 = (double)x[r].z - qb.x, d3 = (double)x[r].w - qb.y;
                            acc[r][b] = fma(d0, d0, acc[r][b]);
                            acc[r][b] = fma(d1, d1, acc[r][b]);
                            acc[r][b] = fma(d2, d2, acc[r][b]);
                            acc[r][b] = fma(d3, d3, acc[r][b]);
                        }
                    }
                }
            }
            // butterfly: every lane ends with every total; lane l then finishes pair (r,b) = (l/QB, l%QB)
            double mine = 0.0;
#pragma unroll
            for (int r = 0; r < kR; ++r)
#pragma unroll
                for (int b = 0; b < QB; ++b) {
                    double t = lz_warp_sum(acc[r][b]);
                    if (lane == r * QB + b) mine = t;
                }
            const int r = lane / QB, b = lane - r * QB;
            const double nx = __shfl_sync(0xffffffffu, my_norm, r & 31);
            const int tg = __shfl_sync(0xffffffffu, my_tag, r & 31);
            const int64_t row = row0 + r;
            bool ok = lane < kR * QB && row < cta_hi && b < nq;
            if (p.tags) ok = ok && tg >= 0 && (want_tag < 0 || tg == want_tag);
            if (ok) {
                double s;
                if (METRIC == LZ_COSINE) {
                    const double den = nx * qnorm[b];
                    s = den > 0.0 ? mine / den : 0.0;
                    s += 0.0;
                } else {
                    s = 0.0 - mine;
                }
                if (s == s) lz_sel_push(&sel[b], lz_ord(s), (u32)row);
            }
        }
#pragma unroll
        for (int b = 0; b < QB; ++b) lz_sel_maybe_compact<kThreads>(&sel[b], kChunkRows);
    }
    __syncthreads();
    for (int b = 0; b < nq; ++b) {
        lz_sel_compact<kThreads>(&sel[b]);
        const int n = min(sel[b].count, p.k);
        const size_t slot = (size_t)(q0 + b) * p.n_chunks + c;
        for (int j = tid; j < n; j += kThreads) {
            p.part_key[slot * p.k + j] = sel[b].key[j];
            p.part_id[slot * p.k + j] = sel[b].id[j];
        }
        if (tid == 0) p.part_cnt[slot] = n;
    }
    __syncthreads();            // shared state is re-initialised by the next item
    if (FUSED) {
        // single query, <= 32 chunks: no second launch — the last CTA merges.  Lane l walks chunk l's sorted list; k rounds
        // of a warp argmax under the (key desc, id asc) order give the global top-k.
        __shared__ bool s_last;
        if (tid == 0) {
            __threadfence();
            s_last = atomicAdd(p.fused_ticket, 1u) == (u32)p.n_chunks - 1u;
        }
        __syncthreads();
        if (s_last && warp == 0) {
            __threadfence();
            const int cnt = lane < p.n_chunks ? __ldcg(p.part_cnt + lane) : 0;
            int pos = 0;
            int64_t* o_ids = reinterpret_cast<int64_t*>(p.fused_out);
            double* o_sc = reinterpret_cast<double*>(p.fused_out + (size_t)p.k * 8);
            int32_t* o_cnt = reinterpret_cast<int32_t*>(p.fused_out + (size_t)p.k * 16);
            int found = 0;
            for (int r = 0; r < p.k; ++r) {
                u64 hk = pos < cnt ? __ldcg(p.part_key + (size_t)lane * p.k + pos) : 0ull;
                u32 hi = pos < cnt ? __ldcg(p.part_id + (size_t)lane * p.k + pos) : 0xffffffffu;
                u64 bk = hk; u32 bi = hi; int bl = lane;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const u64 ok = __shfl_xor_sync(0xffffffffu, bk, o);
                    const u32 oi = __shfl_xor_sync(0xffffffffu, bi, o);
                    const int ol = __shfl_xor_sync(0xffffffffu, bl, o);
                    if (lz_better(ok, oi, bk, bi)) { bk = ok; bi = oi; bl = ol; }
                }
                if (bk == 0ull) break;                                        // every list exhausted (warp-uniform)
                if (lane == bl) ++pos;
                if (lane == 0) {
                    const double v = lz_unord(bk);
                    o_ids[r] = (int64_t)bi + p.id_base;
                    o_sc[r] = p.metric == LZ_COSINE ? v : 0.0 - v;
                }
                ++found;
            }
            if (lane == 0) {
                for (int r = found; r < p.k; ++r) { o_ids[r] = -1; o_sc[r] = __longlong_as_double(0x7ff8000000000000ll); }
                *o_cnt = found;
                *p.fused_ticket = 0u;                                         // ready for the next call
                __threadfence_system();                                       // the answer is in host memory before the flag
                *reinterpret_cast<volatile int32_t*>(p.fused_out + (size_t)p.k * 16 + 4) = p.fused_seq;
            }
        }
    }
    }
}

// One CTA per query: stream the per-chunk lists through a selector.
constexpr int kMergePer = 8;
constexpr int kMergeMaxLists = 8192;
struct MergeParams {
    const u64* part_key; const u32* part_id; const int32_t* part_cnt;   // form A: scan partials
    const int64_t* in_ids; const double* in_scores; int n_queries;      // form B: (id, score) lists
    int n_lists; int k; int cap; int metric; int64_t id_base;
    int64_t list_stride;                                                // form B: elements between two lists
    int max_lists;                                                      // lists whose prefix offsets fit behind the selector (0: flat merge)
    const int32_t* work_list; const int32_t* work_count;
    int64_t* out_ids; double* out_scores; int32_t* out_counts;
};

__global__ void __launch_bounds__(kThreads) merge_kernel(MergeParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    u64* keys = reinterpret_cast<u64*>(smem);
    u32* ids = reinterpret_cast<u32*>(keys + p.cap);
    __shared__ LzSelector sel;
    const int tid = threadIdx.x;
    const int n_items = p.work_list ? *p.work_count : (int)gridDim.x;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int q = p.work_list ? p.work_list[item] : item;
    if (tid == 0) lz_sel_init(&sel, keys, ids, p.cap, p.k);
    __syncthreads();
    // Every list arrives sorted best-first.  Take each list's r-th best entry m_c; the ceil(k/r)-th largest m_c is
    // a value T with at least ceil(k/r) lists holding r entries >= T, i.e. at least k entries >= T overall, so T
    // is a lower bound of the merged k-th best: start from it (ties admitted).  With lists that are slices of
    // one population (per-CTA partials) a small r makes T tight, and only about k entries get past it — a full
    // list's own k-th entry (r = k) is a far weaker bound, and admitting n_lists * k entries through the
    // selector was most of this kernel's time.
    {
        int r = (4 * p.k + p.n_lists - 1) / p.n_lists;
        r = r < 1 ? 1 : (r > p.k ? p.k : r);
        const int need = (p.k + r - 1) / r;
        if (tid == 0) sel.k = need;
        __syncthreads();
        for (int base = 0; base < p.n_lists; base += kThreads) {
            const int c = base + tid;
            if (c < p.n_lists) {
                u64 m = 0ull;
                if (p.part_key) {
                    const size_t slot = (size_t)q * p.n_lists + c;
                    if (p.part_cnt[slot] >= r) m = p.part_key[slot * p.k + r - 1];
                } else {
                    const size_t at = (size_t)c * p.list_stride + (size_t)q * p.k + r - 1;
                    const double sc = p.in_scores[at];
                    if (p.in_ids[at] >= 0 && sc == sc) m = lz_ord((p.metric == LZ_COSINE) ? sc + 0.0 : 0.0 - sc);
                }
                if (m != 0ull) lz_sel_push(&sel, m, (u32)c);
            }
            lz_sel_maybe_compact<kThreads>(&sel, kThreads);
        }
        __syncthreads();
        lz_sel_compact<kThreads>(&sel);
        if (tid == 0) {
            const u64 T = sel.count >= need ? sel.key[need - 1] : 0ull;
            lz_sel_init(&sel, keys, ids, p.cap, p.k);
            if (T != 0ull) { sel.thr_key = T - 1; sel.thr_id = 0u; }       // beat (T-1, 0): any entry with key >= T
        }
        __syncthreads();
    }
    if (p.n_lists <= p.max_lists) {
        // The lists are sorted best-first (valid entries first), so the entries that can matter are a PREFIX of
        // each list: those at or above the seed, or simply the valid ones when no list was full.  One binary
        // search per list finds the prefix length, a scan lays the prefixes end to end, and only those entries
        // (about k in all when seeded) are fetched — not n_lists * k slots.
        int* s_off = reinterpret_cast<int*>(ids + p.cap);                     // [n_lists + 1], behind the selector
        const u64 seed = sel.thr_key != 0ull ? sel.thr_key + 1 : 1ull;        // key 0 = invalid slot
        const int per = (p.n_lists + kThreads - 1) / kThreads;
        const int c0 = min(p.n_lists, tid * per), c1 = min(p.n_lists, c0 + per);
        int mine = 0;
        for (int c = c0; c < c1; ++c) {
            const size_t slot = (size_t)q * p.n_lists + c;
            int lo = 0, hi = p.part_key ? p.part_cnt[slot] : p.k;              // first j in [lo, hi] with key_j < seed
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                u64 key;
                if (p.part_key) {
                    key = p.part_key[slot * p.k + mid];
                } else {
                    const size_t at = (size_t)c * p.list_stride + (size_t)q * p.k + mid;
                    const double sc = p.in_scores[at];
                    key = (p.in_ids[at] >= 0 && sc == sc) ? lz_ord((p.metric == LZ_COSINE) ? sc + 0.0 : 0.0 - sc) : 0ull;
                }
                if (key >= seed) lo = mid + 1; else hi = mid;
            }
            s_off[c + 1] = lo;                                                  // prefix length, scanned below
            mine += lo;
        }
        __shared__ int s_wsum[kThreads / 32];
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((tid & 31) >= o) incl += t;
        }
        if ((tid & 31) == 31) s_wsum[tid >> 5] = incl;
        __syncthreads();
        int run = incl - mine;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) if (w < (tid >> 5)) run += s_wsum[w];
        for (int c = c0; c < c1; ++c) { const int len = s_off[c + 1]; s_off[c + 1] = run + len; run += len; }
        if (tid == 0) s_off[0] = 0;
        __syncthreads();
        const int Q = s_off[p.n_lists];
        for (int base = 0; base < Q; base += kMergePer * kThreads) {
#pragma unroll
            for (int u = 0; u < kMergePer; ++u) {
                const int t = base + u * kThreads + tid;
                if (t < Q) {
                    int lo = 0, hi = p.n_lists - 1;                             // the list c with s_off[c] <= t < s_off[c+1]
                    while (lo < hi) {
                        const int mid = (lo + hi + 1) >> 1;
                        if (s_off[mid] <= t) lo = mid; else hi = mid - 1;
                    }
                    const int c = lo, j = t - s_off[c];
                    if (p.part_key) {
                        const size_t slot = (size_t)q * p.n_lists + c;
                        lz_sel_push(&sel, p.part_key[slot * p.k + j], p.part_id[slot * p.k + j]);
                    } else {
                        const size_t at = (size_t)c * p.list_stride + (size_t)q * p.k + j;
                        const double sc = p.in_scores[at];
                        lz_sel_push(&sel, lz_ord((p.metric == LZ_COSINE) ? sc + 0.0 : 0.0 - sc), (u32)p.in_ids[at]);
                    }
                }
            }
            lz_sel_maybe_compact<kThreads>(&sel, kMergePer * kThreads);
        }
    } else {
    // kMergePer entries per thread and round, loads issued together: one memory round trip per 2048 entries
    const int64_t total = (int64_t)p.n_lists * p.k;
    for (int64_t base = 0; base < total; base += kMergePer * kThreads) {
        u64 mk[kMergePer];
        u32 mi[kMergePer];
#pragma unroll
        for (int u = 0; u < kMergePer; ++u) {
            const int64_t i = base + (int64_t)u * kThreads + tid;
            mk[u] = 0ull; mi[u] = 0u;
            if (i < total) {
                const int c = (int)(i / p.k), j = (int)(i - (int64_t)c * p.k);
                if (p.part_key) {
                    const size_t slot = (size_t)q * p.n_lists + c;
                    if (j < p.part_cnt[slot]) { mk[u] = p.part_key[slot * p.k + j]; mi[u] = p.part_id[slot * p.k + j]; }
                } else {
                    const size_t at = (size_t)c * p.list_stride + (size_t)q * p.k + j;
                    const int64_t id = p.in_ids[at];
                    const double s = p.in_scores[at];
                    if (id >= 0 && s == s) { mk[u] = lz_ord((p.metric == LZ_COSINE) ? s + 0.0 : 0.0 - s); mi[u] = (u32)id; }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kMergePer; ++u)
            if (mk[u] != 0ull) lz_sel_push(&sel, mk[u], mi[u]);
        lz_sel_maybe_compact<kThreads>(&sel, kMergePer * kThreads);
    }
    }
    __syncthreads();
    lz_sel_compact<kThreads>(&sel);
    const int n = min(sel.count, p.k);
    for (int j = tid; j < p.k; j += kThreads) {
        if (j < n) {
            const double v = lz_unord(sel.key[j]);
            p.out_ids[(size_t)q * p.k + j] = (int64_t)sel.id[j] + p.id_base;
            p.out_scores[(size_t)q * p.k + j] = (p.metric == LZ_COSINE) ? v : 0.0 - v;
        } else {
            p.out_ids[(size_t)q * p.k + j] = -1;
            p.out_scores[(size_t)q * p.k + j] = __longlong_as_double(0x7ff8000000000000ll);
        }
    }
    if (tid == 0) p.out_counts[q] = n;
    __syncthreads();
    }
}

// One warp per (query, row) pair, same summation order as the scan.
template <int METRIC>
__global__ void __launch_bounds__(kThreads) score_pairs_kernel(const float* __restrict__ rows, int64_t ld, int dim,
                                                               const double* __restrict__ norms,
                                                               const float* __restrict__ queries, int64_t q_ld,
                                                               const int32_t* __restrict__ pair_q,
                                                               const int64_t* __restrict__ pair_row, int64_t n_pairs,
                                                               double* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * kWarps;
    for (int64_t i = warp; i < n_pairs; i += nwarps) {
        const int64_t row = pair_row[i];
        if (row < 0) {
            if (lane == 0) out[i] = __longlong_as_double(0x7ff8000000000000ll);
            continue;
        }
        const float* x = rows + row * ld;
        const float* q = queries + (int64_t)pair_q[i] * q_ld;
        double acc = 0.0;
        for (int e = lane * 4; e < dim; e += 128) {
            const float4 xv = lz_ldg_stream(x + e);
            const float4 qv = *reinterpret_cast<const float4*>(q + e);
            if (METRIC == LZ_COSINE) {
                acc = fma((double)xv.x, (double)qv.x, acc);
                acc = fma((double)xv.y, (double)qv.y, acc);
                acc = fma((double)xv.z, (double)qv.z, acc);
                acc = fma((double)xv.w, (double)qv.w, acc);
            } else {
                double d0 = (double)xv.x - (double)qv.x, d1 = (double)xv.y - (double)qv.y;
                double d2 = (double)xv.z - (double)qv.z, d3 = (double)xv.w - (double)qv.w;
                acc = fma(d0, d0, acc);
                acc = fma(d1, d1, acc);
                acc = fma(d2, d2, acc);
                acc = fma(d3, d3, acc);
            }
        }
        acc = lz_warp_sum(acc);
        double s;
        if (METRIC == LZ_COSINE) {
            const double qn = sqrt(warp_sumsq(q, dim, lane));
            const double den = norms[row] * qn;
            s = den > 0.0 ? acc / den : 0.0;
            s += 0.0;
        } else {
            s = acc;
        }
        if (lane == 0) out[i] = s;
    }
}

size_t scan_smem_bytes(int qb, int dim, int cap) {
    const int nseg = (dim + 127) / 128;
    return (size_t)qb * nseg * 1024 + (size_t)qb * cap * 12;
}

template <int QB, bool FUSED = false>
int launch_scan(const ScanParams& p, int metric, int n_groups, cudaStream_t st) {
    const size_t smem = scan_smem_bytes(QB, p.dim, p.cap);
    auto kern = metric == LZ_COSINE ? scan_kernel<QB, LZ_COSINE, FUSED> : scan_kernel<QB, LZ_L2, FUSED>;
    LZ_REQUIRE(smem <= 220 * 1024, "lz_search_exact: k=%d dim=%d needs %zu B of shared memory", p.k, p.dim, smem);
    LZ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t grid = (int64_t)n_groups * p.n_chunks;
    if (p.work_list) {                       // fallback mode: a fixed grid walks the device-side work list
        const int64_t cap = (int64_t)lz_sm_count() * 4;
        if (grid > cap) grid = cap;
    }
    kern<<<(unsigned)grid, kThreads, smem, st>>>(p);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

int check_rows(const char* who, const float* rows, int64_t ld, int dim) {
    LZ_REQUIRE(rows != nullptr, "%s: rows is NULL", who);
    LZ_REQUIRE(dim > 0 && dim % 4 == 0, "%s: dim=%d must be a positive multiple of 4", who, dim);
    LZ_REQUIRE(ld >= dim && ld % 4 == 0, "%s: ld=%lld must be >= dim and a multiple of 4", who, (long long)ld);
    LZ_REQUIRE(((uintptr_t)rows & 15) == 0, "%s: rows must be 16-byte aligned", who);
    return LZ_OK;
}

}  // namespace

extern "C" int lz_row_norms(const float* rows, int64_t n, int dim, int64_t ld, double* norms, void* stream) {
    if (n <= 0) return LZ_OK;
    if (int rc = check_rows("lz_row_norms", rows, ld, dim)) return rc;
    LZ_REQUIRE(norms != nullptr, "lz_row_norms: norms is NULL");
    const int sms = lz_sm_count();
    if (sms <= 0) return LZ_ERR_DEVICE;
    int64_t blocks = (n + kWarps - 1) / kWarps;
    if (blocks > (int64_t)sms * 16) blocks = (int64_t)sms * 16;
    row_norms_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(rows, n, dim, ld, norms);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

extern "C" size_t lz_search_exact_scratch_bytes(int n_queries, int k, int n_chunks) {
    if (n_queries <= 0 || k <= 0 || n_chunks <= 0) return 0;
    const size_t slots = (size_t)n_queries * n_chunks;
    size_t b = slots * k * sizeof(u64);
    b += slots * k * sizeof(u32);
    b = (b + 15) & ~(size_t)15;
    b += slots * sizeof(int32_t);
    return (b + 255) & ~(size_t)255;
}

extern "C" int lz_search_exact_default_chunks(int64_t n_rows, int n_groups) {
    const int sms = lz_sm_count();
    if (sms <= 0) return 1;
    if (n_groups < 1) n_groups = 1;
    int64_t by_rows = (n_rows + kChunkRows - 1) / kChunkRows;
    int64_t want = ((int64_t)2 * sms + n_groups - 1) / n_groups;
    int64_t c = by_rows < want ? by_rows : want;
    return (int)(c < 1 ? 1 : c);
}

int lz_search_exact_impl(const float* rows, int64_t ld, int dim, const double* norms, const int32_t* tags,
                         int64_t row_lo, int64_t row_hi, const int64_t* group_lo, const int64_t* group_hi,
                         const int32_t* group_tag, const float* queries, int64_t q_ld, int n_queries,
                         int group_size, int k, int metric, int n_chunks, int64_t id_base, void* scratch,
                         size_t scratch_bytes, int64_t* out_ids, double* out_scores, int32_t* out_counts,
                         const int32_t* work_list, const int32_t* work_count, void* stream, int n_ranges) {
    if (n_queries <= 0) return LZ_OK;
    if (int rc = check_rows("lz_search_exact", rows, ld, dim)) return rc;
    LZ_REQUIRE(n_ranges >= 1 && (n_ranges == 1 || (group_lo != nullptr && n_chunks % n_ranges == 0)),
               "lz_search_exact: n_ranges=%d needs per-group ranges and n_chunks (%d) a multiple of it", n_ranges, n_chunks);
    LZ_REQUIRE(metric == LZ_COSINE || metric == LZ_L2, "lz_search_exact: unknown metric %d", metric);
    LZ_REQUIRE(metric == LZ_L2 || norms != nullptr, "lz_search_exact: cosine needs norms");
    LZ_REQUIRE(k >= 1 && k <= LZ_MAX_K, "lz_search_exact: k=%d outside 1..%d", k, LZ_MAX_K);
    LZ_REQUIRE(group_size == 1 || group_size == 2 || group_size == 4,
               "lz_search_exact: group_size=%d must be 1, 2 or 4", group_size);
    LZ_REQUIRE(n_chunks >= 1, "lz_search_exact: n_chunks=%d", n_chunks);
    LZ_REQUIRE(work_list == nullptr || (group_size == 1 && work_count != nullptr),
               "lz_search_exact: a work list needs group_size 1 and a count");
    LZ_REQUIRE(queries && q_ld >= dim && q_ld % 4 == 0 && ((uintptr_t)queries & 15) == 0,
               "lz_search_exact: queries must be 16-byte aligned with q_ld >= dim, q_ld %% 4 == 0");
    LZ_REQUIRE((group_lo == nullptr) == (group_hi == nullptr), "lz_search_exact: group_lo/group_hi must come together");
    LZ_REQUIRE(out_ids && out_scores && out_counts && scratch, "lz_search_exact: NULL output or scratch");
    LZ_REQUIRE(row_hi <= 0xfffffffeLL, "lz_search_exact: row index %lld exceeds 32 bits", (long long)row_hi);
    const size_t need = lz_search_exact_scratch_bytes(n_queries, k, n_chunks);
    if (scratch_bytes < need) {
        lz_set_error("lz_search_exact: scratch has %zu bytes, needs %zu", scratch_bytes, need);
        return LZ_ERR_SCRATCH;
    }
    const int n_groups = (n_queries + group_size - 1) / group_size;
    const size_t slots = (size_t)n_queries * n_chunks;
    ScanParams p;
    p.rows = rows; p.ld = ld; p.dim = dim; p.norms = norms; p.tags = tags;
    p.row_lo = row_lo; p.row_hi = row_hi; p.group_lo = group_lo; p.group_hi = group_hi; p.group_tag = group_tag;
    p.queries = queries; p.q_ld = q_ld; p.n_queries = n_queries;
    p.k = k; p.n_chunks = n_chunks; p.cap = lz_next_pow2(k + kChunkRows);
    p.work_list = work_list; p.work_count = work_count; p.n_ranges = n_ranges;
    p.fused_ticket = nullptr; p.fused_out = nullptr; p.fused_seq = 0; p.metric = metric; p.id_base = id_base;
    p.want_tag_all = -1;
    unsigned char* s = (unsigned char*)scratch;
    p.part_key = (u64*)s;
    p.part_id = (u32*)(s + slots * k * sizeof(u64));
    p.part_cnt = (int32_t*)(s + (((slots * k * (sizeof(u64) + sizeof(u32))) + 15) & ~(size_t)15));
    cudaStream_t st = (cudaStream_t)stream;
    int rc = group_size == 1 ? launch_scan<1>(p, metric, n_groups, st)
           : group_size == 2 ? launch_scan<2>(p, metric, n_groups, st)
                             : launch_scan<4>(p, metric, n_groups, st);
    if (rc) return rc;
    MergeParams m;
    m.part_key = p.part_key; m.part_id = p.part_id; m.part_cnt = p.part_cnt;
    m.in_ids = nullptr; m.in_scores = nullptr; m.n_queries = n_queries; m.list_stride = 0;
    m.n_lists = n_chunks; m.k = k; m.cap = lz_next_pow2(k + kMergePer * kThreads); m.metric = metric; m.id_base = id_base;
    m.work_list = work_list; m.work_count = work_count;
    m.out_ids = out_ids; m.out_scores = out_scores; m.out_counts = out_counts;
    m.max_lists = m.n_lists <= kMergeMaxLists ? kMergeMaxLists : 0;
    const size_t msmem = (size_t)m.cap * 12 + (m.max_lists ? ((size_t)m.n_lists + 1) * 4 : 0);
    LZ_CUDA(cudaFuncSetAttribute(merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
    int mgrid = n_queries;
    if (work_list && mgrid > lz_sm_count() * 8) mgrid = lz_sm_count() * 8;
    merge_kernel<<<mgrid, kThreads, msmem, st>>>(m);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

extern "C" int lz_search_exact(const float* rows, int64_t ld, int dim, const double* norms, const int32_t* tags,
                               int64_t row_lo, int64_t row_hi, const int64_t* group_lo, const int64_t* group_hi,
                               const int32_t* group_tag, const float* queries, int64_t q_ld, int n_queries,
                               int group_size, int k, int metric, int n_chunks, int64_t id_base, void* scratch,
                               size_t scratch_bytes, int64_t* out_ids, double* out_scores, int32_t* out_counts,
                               void* stream) {
    return lz_search_exact_impl(rows, ld, dim, norms, tags, row_lo, row_hi, group_lo, group_hi, group_tag, queries,
                                q_ld, n_queries, group_size, k, metric, n_chunks, id_base, scratch, scratch_bytes,
                                out_ids, out_scores, out_counts, nullptr, nullptr, stream, 1);
}

extern "C" size_t lz_search_one_scratch_bytes(int k) {
    return lz_search_exact_scratch_bytes(1, k, 32) + 256;
}

extern "C" int lz_search_one(const float* rows, int64_t ld, int dim, const double* norms, const int32_t* tags, int want_tag,
                             int64_t row_lo, int64_t row_hi, const float* query, int k, int metric, int64_t id_base,
                             void* scratch, size_t scratch_bytes, void* out_block, int seq, void* stream) {
    if (int rc = check_rows("lz_search_one", rows, ld, dim)) return rc;
    LZ_REQUIRE(metric == LZ_COSINE || metric == LZ_L2, "lz_search_one: unknown metric %d", metric);
    LZ_REQUIRE(metric == LZ_L2 || norms != nullptr, "lz_search_one: cosine needs norms");
    LZ_REQUIRE(k >= 1 && k <= LZ_MAX_K, "lz_search_one: k=%d outside 1..%d", k, LZ_MAX_K);
    LZ_REQUIRE(query && ((uintptr_t)query & 15) == 0 && out_block && scratch, "lz_search_one: NULL or misaligned pointer");
    LZ_REQUIRE(0 <= row_lo && row_lo <= row_hi && row_hi <= 0xfffffffeLL, "lz_search_one: bad row range");
    const size_t need = lz_search_one_scratch_bytes(k);
    if (scratch_bytes < need) {
        lz_set_error("lz_search_one: scratch has %zu bytes, needs %zu", scratch_bytes, need);
        return LZ_ERR_SCRATCH;
    }
    const int sms = lz_sm_count();
    if (sms <= 0) return LZ_ERR_DEVICE;
    int64_t chunks = (row_hi - row_lo + 255) / 256;                       // at least ~256 rows per CTA, at most 32 CTAs
    const int n_chunks = (int)(chunks < 1 ? 1 : (chunks > 32 ? 32 : chunks));
    ScanParams p = {};
    p.rows = rows; p.ld = ld; p.dim = dim; p.norms = norms; p.tags = tags;
    p.row_lo = row_lo; p.row_hi = row_hi;
    p.queries = query; p.q_ld = dim; p.n_queries = 1;
    p.k = k; p.n_chunks = n_chunks; p.cap = lz_next_pow2(k + kChunkRows);
    p.n_ranges = 1; p.metric = metric; p.id_base = id_base;
    unsigned char* s = (unsigned char*)scratch;
    p.fused_ticket = (u32*)s;                                              // zero before the first call (the caller clears the scratch once)
    s += 256;
    const size_t slots = 32;
    p.part_key = (u64*)s;
    p.part_id = (u32*)(s + slots * k * sizeof(u64));
    p.part_cnt = (int32_t*)(s + (((slots * k * (sizeof(u64) + sizeof(u32))) + 15) & ~(size_t)15));
    p.fused_out = (unsigned char*)out_block; p.fused_seq = seq;
    p.group_tag = nullptr;
    p.want_tag_all = tags != nullptr ? want_tag : -1;
    return launch_scan<1, true>(p, metric, 1, (cudaStream_t)stream);
}

extern "C" int lz_search_exact_ranges(const float* rows, int64_t ld, int dim, const double* norms, const int32_t* tags,
                                      const int64_t* range_lo, const int64_t* range_hi, int n_ranges, int64_t n_rows_total,
                                      const float* queries, int64_t q_ld, int n_queries, int k, int metric,
                                      int chunks_per_range, int64_t id_base, void* scratch, size_t scratch_bytes,
                                      int64_t* out_ids, double* out_scores, int32_t* out_counts, void* stream) {
    LZ_REQUIRE(n_ranges >= 1 && chunks_per_range >= 1 && range_lo && range_hi, "lz_search_exact_ranges: bad ranges");
    return lz_search_exact_impl(rows, ld, dim, norms, tags, 0, n_rows_total, range_lo, range_hi, nullptr, queries,
                                q_ld, n_queries, 1, k, metric, n_ranges * chunks_per_range, id_base, scratch, scratch_bytes,
                                out_ids, out_scores, out_counts, nullptr, nullptr, stream, n_ranges);
}

extern "C" int lz_merge_topk(const int64_t* in_ids, const double* in_scores, int n_lists, int n_queries, int k,
                             int metric, int64_t* out_ids, double* out_scores, int32_t* out_counts, void* stream) {
    return lz_merge_topk_strided(in_ids, in_scores, (int64_t)n_queries * k, n_lists, n_queries, k, metric, out_ids,
                                 out_scores, out_counts, stream);
}

extern "C" int lz_merge_topk_strided(const int64_t* in_ids, const double* in_scores, int64_t list_stride, int n_lists,
                                     int n_queries, int k, int metric, int64_t* out_ids, double* out_scores,
                                     int32_t* out_counts, void* stream) {
    if (n_queries <= 0) return LZ_OK;
    LZ_REQUIRE(in_ids && in_scores && out_ids && out_scores && out_counts, "lz_merge_topk: NULL pointer");
    LZ_REQUIRE(list_stride >= (int64_t)n_queries * k, "lz_merge_topk: list_stride=%lld below n_queries*k", (long long)list_stride);
    LZ_REQUIRE(n_lists >= 1 && k >= 1 && k <= LZ_MAX_K, "lz_merge_topk: n_lists=%d k=%d", n_lists, k);
    LZ_REQUIRE(metric == LZ_COSINE || metric == LZ_L2, "lz_merge_topk: unknown metric %d", metric);
    MergeParams m;
    m.part_key = nullptr; m.part_id = nullptr; m.part_cnt = nullptr;
    m.in_ids = in_ids; m.in_scores = in_scores; m.n_queries = n_queries; m.list_stride = list_stride;
    m.n_lists = n_lists; m.k = k; m.cap = lz_next_pow2(k + kMergePer * kThreads); m.metric = metric; m.id_base = 0;
    m.work_list = nullptr; m.work_count = nullptr;
    m.out_ids = out_ids; m.out_scores = out_scores; m.out_counts = out_counts;
    m.max_lists = m.n_lists <= kMergeMaxLists ? kMergeMaxLists : 0;
    const size_t msmem = (size_t)m.cap * 12 + (m.max_lists ? ((size_t)m.n_lists + 1) * 4 : 0);
    LZ_CUDA(cudaFuncSetAttribute(merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
    merge_kernel<<<n_queries, kThreads, msmem, (cudaStream_t)stream>>>(m);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

extern "C" int lz_score_pairs(const float* rows, int64_t ld, int dim, const double* norms, const float* queries,
                              int64_t q_ld, const int32_t* pair_q, const int64_t* pair_row, int64_t n_pairs,
                              int metric, double* out_scores, void* stream) {
    if (n_pairs <= 0) return LZ_OK;
    if (int rc = check_rows("lz_score_pairs", rows, ld, dim)) return rc;
    LZ_REQUIRE(metric == LZ_COSINE || metric == LZ_L2, "lz_score_pairs: unknown metric %d", metric);
    LZ_REQUIRE(metric == LZ_L2 || norms != nullptr, "lz_score_pairs: cosine needs norms");
    LZ_REQUIRE(queries && q_ld >= dim && q_ld % 4 == 0 && ((uintptr_t)queries & 15) == 0,
               "lz_score_pairs: queries must be 16-byte aligned with q_ld >= dim, q_ld %% 4 == 0");
    LZ_REQUIRE(pair_q && pair_row && out_scores, "lz_score_pairs: NULL pointer");
    const int sms = lz_sm_count();
    if (sms <= 0) return LZ_ERR_DEVICE;
    int64_t blocks = (n_pairs + kWarps - 1) / kWarps;
    if (blocks > (int64_t)sms * 16) blocks = (int64_t)sms * 16;
    if (metric == LZ_COSINE)
        score_pairs_kernel<LZ_COSINE><<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(
            rows, ld, dim, norms, queries, q_ld, pair_q, pair_row, n_pairs, out_scores);
    else
        score_pairs_kernel<LZ_L2><<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(
            rows, ld, dim, norms, queries, q_ld, pair_q, pair_row, n_pairs, out_scores);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

// ---- K6: segmented mean (centroids) ----------------------------------------------------------------------
// Stands behind _create_super_nodes_for_shard (core/memory_system.py:916-917: centroid = np.mean of the
// member embeddings) and the centroid update of shard reassignment.  One CTA per (segment, 128-column
// chunk); warp w adds rows w, w+8, ... in float64 in a fixed order and the eight partial sums are combined
// in warp order, so the result is deterministic.  HBM-bound: dim*4 bytes per member row.
namespace {
__global__ void __launch_bounds__(kThreads) segment_mean_kernel(const float* __restrict__ rows, int64_t ld, int dim,
                                                                const int64_t* __restrict__ order,
                                                                const int64_t* __restrict__ seg, double* __restrict__ out) {
    __shared__ double part[kWarps][128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int s = blockIdx.x, col = blockIdx.y * 128 + lane * 4;
    const int64_t lo = seg[s], hi = seg[s + 1];
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    if (col < dim) {
        for (int64_t i = lo + warp; i < hi; i += kWarps) {
            const int64_t r = order ? order[i] : i;
            const float4 v = lz_ldg_stream(rows + r * ld + col);
            a0 += (double)v.x; a1 += (double)v.y; a2 += (double)v.z; a3 += (double)v.w;
        }
    }
    part[warp][lane * 4 + 0] = a0; part[warp][lane * 4 + 1] = a1;
    part[warp][lane * 4 + 2] = a2; part[warp][lane * 4 + 3] = a3;
    __syncthreads();
    if (threadIdx.x < 128 && blockIdx.y * 128 + threadIdx.x < dim) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) t += part[w][threadIdx.x];
        const int64_t n = hi - lo;
        out[(int64_t)s * dim + blockIdx.y * 128 + threadIdx.x] = n > 0 ? t / (double)n : 0.0;
    }
}
}  // namespace

extern "C" int lz_segment_mean(const float* rows, int64_t ld, int dim, const int64_t* order, const int64_t* seg_offsets,
                               int n_segments, double* out_means, void* stream) {
    if (n_segments <= 0) return LZ_OK;
    if (int rc = check_rows("lz_segment_mean", rows, ld, dim)) return rc;
    LZ_REQUIRE(seg_offsets && out_means, "lz_segment_mean: NULL pointer");
    dim3 grid((unsigned)n_segments, (unsigned)((dim + 127) / 128));
    LZ_REQUIRE(grid.y <= 65535, "lz_segment_mean: dim too large");
    segment_mean_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(rows, ld, dim, order, seg_offsets, out_means);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

// ---- grouping rows by label: the member lists of a k-means style (re)assignment ------------------------------------
// A STABLE counting sort of the row indices by label (labels < n_labels <= kGroupMaxLabels), so that lz_segment_mean
// sums every centroid's members in ascending row order — a fixed order, hence the same bits on every run.  Three
// launches, no CTA waits on another: per-tile label histograms; per label an exclusive scan over the tiles (+ the label
// offsets); per tile the scatter, ranking equal labels in row order (rounds of 256 rows, warps admitted in order,
// __match_any inside a warp).
namespace {
constexpr int kGroupTile = 2048;
constexpr int kGroupMaxLabels = 16384;

template <bool SCATTER>
__global__ void __launch_bounds__(kThreads) group_by_label_kernel(const int64_t* __restrict__ labels, int64_t n, int n_labels,
                                                                  int32_t* __restrict__ tile_hist,     // [n_tiles][n_labels]
                                                                  int64_t* __restrict__ order) {
    extern __shared__ int s_cnt[];                                            // [n_labels] running count / base
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t base = (int64_t)blockIdx.x * kGroupTile;
    int32_t* hist = tile_hist + (size_t)blockIdx.x * n_labels;
    for (int l = tid; l < n_labels; l += kThreads) s_cnt[l] = SCATTER ? hist[l] : 0;     // scatter: where the tile's first row of label l goes
    __syncthreads();
    for (int r = 0; r < kGroupTile / kThreads; ++r) {
        const int64_t i = base + (int64_t)r * kThreads + tid;
        const int64_t lab64 = i < n ? labels[i] : -1;
        const int lab = (lab64 >= 0 && lab64 < n_labels) ? (int)lab64 : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, lab);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        const bool leader = rank == 0;
        int at = 0;
        for (int w = 0; w < kThreads / 32; ++w) {                             // warps in order: rows keep their order within a label
            if (warp == w && lab >= 0 && leader) {
                at = s_cnt[lab];
                s_cnt[lab] = at + __popc(peers);
            }
            __syncthreads();
        }
        if (SCATTER && lab >= 0) {
            at = __shfl_sync(peers, at, __ffs(peers) - 1);
            order[at + rank] = i;
        }
    }
    if (!SCATTER) {
        __syncthreads();
        for (int l = tid; l < n_labels; l += kThreads) hist[l] = s_cnt[l];
    }
}

// per label: counts of the tiles -> exclusive positions (label offset + tiles before), in place; seg[l] = label offsets
__global__ void __launch_bounds__(kThreads) group_scan_kernel(int32_t* __restrict__ tile_hist, int n_tiles, int n_labels,
                                                              int64_t* __restrict__ seg) {
    __shared__ long long s_warp[kThreads / 32];
    __shared__ long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    // pass 1: total per label, block-wide exclusive scan over the labels in chunks of 256
    for (int l0 = 0; l0 < n_labels; l0 += kThreads) {
        const int l = l0 + tid;
        long long tot = 0;
        if (l < n_labels)
            for (int t = 0; t < n_tiles; ++t) tot += tile_hist[(size_t)t * n_labels + l];
        long long incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        long long off = s_carry + incl - tot;
        for (int w = 0; w < warp; ++w) off += s_warp[w];
        if (l < n_labels) {
            seg[l] = off;
            long long run = off;
            for (int t = 0; t < n_tiles; ++t) {
                const int c = tile_hist[(size_t)t * n_labels + l];
                tile_hist[(size_t)t * n_labels + l] = (int32_t)run;
                run += c;
            }
        }
        __syncthreads();
        if (tid == kThreads - 1) s_carry = off + tot;
        __syncthreads();
    }
    if (tid == 0) seg[n_labels] = s_carry;
}
}  // namespace

extern "C" size_t lz_group_by_label_scratch_bytes(int64_t n, int n_labels) {
    if (n <= 0 || n_labels <= 0) return 256;
    const int64_t tiles = (n + kGroupTile - 1) / kGroupTile;
    return (size_t)(((size_t)tiles * n_labels * 4 + 255) & ~(size_t)255);
}

extern "C" int lz_group_by_label(const int64_t* labels, int64_t n, int n_labels, void* scratch, size_t scratch_bytes,
                                 int64_t* out_order, int64_t* out_offsets, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    LZ_REQUIRE(n_labels >= 1 && n_labels <= kGroupMaxLabels, "lz_group_by_label: n_labels=%d outside 1..%d", n_labels, kGroupMaxLabels);
    LZ_REQUIRE(out_offsets != nullptr, "lz_group_by_label: NULL output");
    if (n <= 0) {
        LZ_CUDA(cudaMemsetAsync(out_offsets, 0, (size_t)(n_labels + 1) * 8, st));
        return LZ_OK;
    }
    LZ_REQUIRE(labels && out_order && scratch, "lz_group_by_label: NULL pointer");
    LZ_REQUIRE(n < (1ll << 31), "lz_group_by_label: n too large");
    const size_t need = lz_group_by_label_scratch_bytes(n, n_labels);
    if (scratch_bytes < need) {
        lz_set_error("lz_group_by_label: scratch has %zu bytes, needs %zu", scratch_bytes, need);
        return LZ_ERR_SCRATCH;
    }
    const int tiles = (int)((n + kGroupTile - 1) / kGroupTile);
    const size_t smem = (size_t)n_labels * 4;
    LZ_CUDA(cudaFuncSetAttribute(group_by_label_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    LZ_CUDA(cudaFuncSetAttribute(group_by_label_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int32_t* hist = (int32_t*)scratch;
    group_by_label_kernel<false><<<tiles, kThreads, smem, st>>>(labels, n, n_labels, hist, nullptr);
    group_scan_kernel<<<1, kThreads, 0, st>>>(hist, tiles, n_labels, out_offsets);
    group_by_label_kernel<true><<<tiles, kThreads, smem, st>>>(labels, n, n_labels, hist, out_order);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

// ---- importance scoring + lowest-k selection (buffer-limit eviction) ---------------------------------------
// Stands behind MemorySystem._enforce_buffer_limit (core/memory_system.py:535-578):
//   days_old   = (now - last_accessed) / 86400
//   importance = salience * 0.5 + min(1.0, access_count / 10) * 0.3 + (1.0 / (1.0 + days_old)) * 0.2
// evaluated in float64 with the reference's operation order (no fused multiply-adds: every product and sum is
// rounded as Python rounds it), then the `k` LEAST important eligible nodes in (importance asc, index asc)
// order — the reference's stable ascending sort over its iteration order, which the caller's index order must
// follow.  One streaming pass: 17 bytes per node (float32 salience) or 21 (float64), per-CTA selector, then the
// per-list merge the scan uses.  HBM-bound.
namespace {
constexpr int kEvictPerThread = 4;       // 1024 nodes per tile: a 24 KB selector, 8 CTAs per SM to hide the division latency
constexpr int kEvictChunk = kThreads * kEvictPerThread;

struct EvictParams {
    const float* sal32; const double* sal64; const int32_t* access; const double* last; const uint8_t* skip;
    int64_t n; double now; int k, cap, n_chunks;
    double* out_all;                                       // optional: importance of every node
    u64* part_key; u32* part_id; int32_t* part_cnt;
    u64* gthr;                                             // best k-th key any CTA has reached (0 = none yet)
    int64_t stride;                                        // 1, or > 1 in the sampling pass (CTA b reads tile b*stride)
    int sample;                                            // sampling pass: one tile per CTA, its sel_k-th best goes to sample_keys
    int sel_k;                                             // selector size in the sampling pass (the rank r)
    u64* sample_keys;                                      // [gridDim.x] in the sampling pass
};

__device__ __forceinline__ double importance_of(double sal, int access, double last, double now, const double* tenth) {
    const double days = __ddiv_rn(__dsub_rn(now, last), 86400.0);
    const double t1 = __dmul_rn(sal, 0.5);
    // access / 10 for 0 <= access < 10 is the correctly rounded quotient = the double nearest to the decimal
    // tenth, which is what the table holds; >= 10 saturates at 1.0 (saves one float64 division per node).  A table
    // in shared memory, not a switch: a switch compiles to a jump table and a warp with ten different counts runs
    // its ten arms one after the other.
    const double frac = access >= 10 ? 1.0 : (access >= 0 ? tenth[access] : fmin(1.0, __ddiv_rn((double)access, 10.0)));
    const double t2 = __dmul_rn(frac, 0.3);
    const double t3 = __dmul_rn(__drcp_rn(__dadd_rn(1.0, days)), 0.2);       // 1.0 / x, correctly rounded
    return __dadd_rn(__dadd_rn(t1, t2), t3);
}

__global__ void __launch_bounds__(kThreads) importance_select_kernel(EvictParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    u64* keys = reinterpret_cast<u64*>(smem);
    u32* ids = reinterpret_cast<u32*>(keys + p.cap);
    __shared__ LzSelector sel;
    __shared__ double s_tenth[10];
    const int tid = threadIdx.x;
    if (tid == 0) lz_sel_init(&sel, keys, ids, p.cap, p.sample ? p.sel_k : p.k);
    if (tid < 10) {
        const double t[10] = {0.0, 0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9};
        s_tenth[tid] = t[tid];
    }
    __syncthreads();
    const int64_t n_all = (p.n + kEvictChunk - 1) / kEvictChunk;
    const int64_t n_tiles = p.sample ? min(n_all, (int64_t)blockIdx.x * p.stride + 1) : n_all;     // sampling: one tile
    for (int64_t t = (int64_t)blockIdx.x * p.stride; t < n_tiles; t += (int64_t)gridDim.x * p.stride) {
        const int64_t base = t * kEvictChunk;
        // any CTA's k-th best key bounds the global k-th best from below: every thread filters against the best
        // one published so far (ties admitted), so after the first tiles almost nothing is pushed any more
        u64 g;
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(g) : "l"(p.gthr) : "memory");
        // all of the tile's loads are issued before the first push (a push is a warp vote + shared-memory
        // traffic the compiler will not move loads across): one memory round trip per tile, not sixteen
        u64 key[kEvictPerThread];
#pragma unroll
        for (int j = 0; j < kEvictPerThread; ++j) {
            const int64_t i = base + (int64_t)j * kThreads + tid;                // coalesced across the CTA
            key[j] = 0ull;                                                        // 0 = nothing to push
            if (i < p.n) {
                const double s = p.sal64 ? p.sal64[i] : (double)p.sal32[i];
                const double imp = importance_of(s, p.access[i], p.last[i], p.now, s_tenth);
                if (p.out_all && !p.sample) p.out_all[i] = imp;
                // larger key = better = LESS important; NaN importance is never evicted
                const u64 kk = lz_ord(0.0 - imp);
                if (imp == imp && kk >= g && !(p.skip && p.skip[i])) key[j] = kk;
            }
        }
#pragma unroll
        for (int j = 0; j < kEvictPerThread; ++j)
            lz_sel_push_warp(&sel, key[j], (u32)(base + (int64_t)j * kThreads + tid), key[j] != 0ull);
        lz_sel_maybe_compact<kThreads>(&sel, kEvictChunk);
        if (tid == 0 && !p.sample && sel.thr_key != 0ull) atomicMax(p.gthr, sel.thr_key);
    }
    __syncthreads();
    lz_sel_compact<kThreads>(&sel);
    if (p.sample) {                                        // this tile's r-th best, for evict_seed_kernel
        if (tid == 0) p.sample_keys[blockIdx.x] = sel.count >= p.sel_k ? sel.key[p.sel_k - 1] : 0ull;
        return;
    }
    const int n = min(sel.count, p.k);
    const size_t slot = blockIdx.x;
    for (int j = tid; j < n; j += kThreads) {
        p.part_key[slot * p.k + j] = sel.key[j];
        p.part_id[slot * p.k + j] = sel.id[j];
    }
    if (tid == 0) p.part_cnt[slot] = n;
}

// The `need`-th largest of the sampled tiles' r-th best keys: at least `need` tiles hold r nodes at or above it,
// so at least need * r >= k nodes are — a lower bound of the k-th best key, found before the full pass starts,
// so that the full pass filters from its first node instead of filling and sorting every CTA's buffer.
__global__ void __launch_bounds__(kThreads) evict_seed_kernel(const u64* __restrict__ sample_keys, int n_keys, int need,
                                                              int cap, u64* __restrict__ gthr) {
    extern __shared__ __align__(16) unsigned char smem[];
    u64* keys = reinterpret_cast<u64*>(smem);
    u32* ids = reinterpret_cast<u32*>(keys + cap);
    __shared__ LzSelector sel;
    const int tid = threadIdx.x;
    if (tid == 0) lz_sel_init(&sel, keys, ids, cap, need);
    __syncthreads();
    for (int base = 0; base < n_keys; base += kThreads) {
        const int i = base + tid;
        const u64 k = i < n_keys ? sample_keys[i] : 0ull;
        if (k != 0ull) lz_sel_push(&sel, k, (u32)i);
        lz_sel_maybe_compact<kThreads>(&sel, kThreads);
    }
    __syncthreads();
    lz_sel_compact<kThreads>(&sel);
    if (tid == 0 && sel.count >= need) *gthr = sel.key[need - 1];
}

int evict_chunks(int64_t n, int k) {
    const int sms = lz_sm_count();
    int64_t tiles = (n + kEvictChunk - 1) / kEvictChunk;
    int64_t want = (int64_t)(sms > 0 ? sms : 1) * (k <= 256 ? 8 : 2);      // selector: 12 B x next_pow2(k + 1024) of shared memory
    return (int)(tiles < want ? (tiles < 1 ? 1 : tiles) : want);
}
}  // namespace

extern "C" size_t lz_evict_lowest_scratch_bytes(int64_t n, int k) {
    if (n <= 0 || k <= 0) return 0;
    return lz_search_exact_scratch_bytes(1, k, evict_chunks(n, k)) + 256 + (size_t)evict_chunks(n, k) * 8;   // + threshold word + sample keys
}

extern "C" int lz_evict_lowest(const void* salience, int salience_is_f64, const int32_t* access_count,
                               const double* last_accessed, const uint8_t* skip, int64_t n, double now, int k,
                               void* scratch, size_t scratch_bytes, int64_t* out_idx, double* out_importance,
                               int32_t* out_count, double* out_all_importance, void* stream) {
    LZ_REQUIRE(out_count != nullptr, "lz_evict_lowest: out_count is NULL");
    cudaStream_t st = (cudaStream_t)stream;
    if (n <= 0 || k <= 0) {
        LZ_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int32_t), st));
        return LZ_OK;
    }
    LZ_REQUIRE(salience && access_count && last_accessed && out_idx && out_importance && scratch,
               "lz_evict_lowest: NULL pointer");
    LZ_REQUIRE(k <= LZ_MAX_K, "lz_evict_lowest: k=%d exceeds %d (evict in rounds)", k, LZ_MAX_K);
    LZ_REQUIRE(n <= 0xfffffffeLL, "lz_evict_lowest: n=%lld exceeds 32-bit indices", (long long)n);
    const int chunks = evict_chunks(n, k);
    const size_t part_bytes = lz_search_exact_scratch_bytes(1, k, chunks);
    const size_t need = part_bytes + 256 + (size_t)chunks * 8;
    if (scratch_bytes < need) {
        lz_set_error("lz_evict_lowest: scratch has %zu bytes, needs %zu", scratch_bytes, need);
        return LZ_ERR_SCRATCH;
    }
    EvictParams p;
    p.sal32 = salience_is_f64 ? nullptr : (const float*)salience;
    p.sal64 = salience_is_f64 ? (const double*)salience : nullptr;
    p.access = access_count; p.last = last_accessed; p.skip = skip; p.n = n; p.now = now;
    p.k = k; p.cap = lz_next_pow2(k + kEvictChunk); p.n_chunks = chunks; p.out_all = out_all_importance;
    unsigned char* s = (unsigned char*)scratch;
    const size_t slots = (size_t)chunks;
    p.part_key = (u64*)s;
    p.part_id = (u32*)(s + slots * k * sizeof(u64));
    p.part_cnt = (int32_t*)(s + (((slots * k * (sizeof(u64) + sizeof(u32))) + 15) & ~(size_t)15));
    p.gthr = (u64*)(s + part_bytes);
    LZ_CUDA(cudaMemsetAsync(p.gthr, 0, 8, st));
    const size_t smem = (size_t)p.cap * 12;
    LZ_CUDA(cudaFuncSetAttribute(importance_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    p.stride = 1; p.sample = 0; p.sel_k = k; p.sample_keys = (u64*)(s + part_bytes + 256);
    {
        // sampling pass: G tiles spread evenly over the nodes, one per CTA; see evict_seed_kernel
        const int64_t n_tiles = (n + kEvictChunk - 1) / kEvictChunk;
        const int g_max = chunks < 2 * lz_sm_count() ? chunks : 2 * lz_sm_count();   // a few hundred tiles rank k <= 1024 well
        const int G = (int)(n_tiles / 8 < g_max ? n_tiles / 8 : g_max);
        int r = G > 0 ? (4 * k + G - 1) / G : 0;
        if (r < 1) r = 1;
        const int need_tiles = (k + r - 1) / r;
        if (G >= 16 && r <= kEvictChunk / 4 && need_tiles <= G) {
            EvictParams ps = p;
            ps.sample = 1; ps.sel_k = r; ps.stride = n_tiles / G; ps.out_all = nullptr;
            importance_select_kernel<<<G, kThreads, smem, st>>>(ps);
            const int scap = lz_next_pow2(need_tiles + kThreads);
            const size_t ssmem = (size_t)scap * 12;
            LZ_CUDA(cudaFuncSetAttribute(evict_seed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssmem));
            evict_seed_kernel<<<1, kThreads, ssmem, st>>>(p.sample_keys, G, need_tiles, scap, p.gthr);
        }
    }
    importance_select_kernel<<<chunks, kThreads, smem, st>>>(p);
    LZ_CUDA(cudaGetLastError());
    MergeParams m;
    m.part_key = p.part_key; m.part_id = p.part_id; m.part_cnt = p.part_cnt;
    m.in_ids = nullptr; m.in_scores = nullptr; m.n_queries = 1; m.list_stride = 0;
    m.n_lists = chunks; m.k = k; m.cap = lz_next_pow2(k + kMergePer * kThreads); m.metric = LZ_L2;   // L2 form returns -key = importance
    m.id_base = 0; m.work_list = nullptr; m.work_count = nullptr;
    m.out_ids = out_idx; m.out_scores = out_importance; m.out_counts = out_count;
    m.max_lists = m.n_lists <= kMergeMaxLists ? kMergeMaxLists : 0;
    const size_t msmem = (size_t)m.cap * 12 + (m.max_lists ? ((size_t)m.n_lists + 1) * 4 : 0);
    LZ_CUDA(cudaFuncSetAttribute(merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
    merge_kernel<<<1, kThreads, msmem, st>>>(m);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}
