// Adjacency on the device: a CSR over the edge table's (src, dst) columns and the neighbour lookup on it.
//
// Stands behind MemoryShard.get_neighbors (core/memory_shard.py:54-62; fan-out core/buffer_graph.py:72-77):
// the reference walks EVERY edge of the shard per call and returns the other endpoint of each edge that
// touches the node with weight >= min_weight, in edge-dict (= insertion = edge index) order.  Here the edges
// are indexed once per edge-table version:
//   entries   two per edge, (key = src, val = 2e) and (key = dst, val = 2e + 1); a self-loop keeps only the
//             first (the reference's `if src == node ... elif tgt == node` yields it once); edges with an
//             endpoint outside [0, n_nodes) get the sentinel key n_nodes and sort to the end
//   sort      stable LSD radix sort by key, 8 bits per pass, ceil(bits(n_nodes + 1) / 8) passes — stable, so
//             each node's entries stay in ascending edge order, which IS the reference's list order
//   offsets   degree histogram (one atomic per entry) + exclusive scan
// and a lookup is one warp per queried node over its own entries (weights are read live from the edge
// table, so decay does not invalidate the index; prune / insert do).  All passes are HBM-bound streaming
// kernels; nothing here waits on another CTA.
#include "lz_common.cuh"

namespace {

constexpr int kT = 256;                       // threads per CTA everywhere in this file
constexpr int kScanPer = 8;
constexpr int kScanTile = kT * kScanPer;      // 2048 elements per scan tile

// ---- exclusive scan int32 -> int64, three launches ------------------------------------------------------
__device__ __forceinline__ long long block_exclusive(long long v, long long* s_warp, long long* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    long long incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const long long t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    long long base = 0, all = 0;
#pragma unroll
    for (int w = 0; w < kT / 32; ++w) {
        const long long t = s_warp[w];
        if (w < warp) base += t;
        all += t;
    }
    if (total) *total = all;
    __syncthreads();
    return base + incl - v;
}

__global__ void __launch_bounds__(kT) scan_sums_kernel(const int32_t* __restrict__ in, int64_t n, int64_t* __restrict__ sums) {
    __shared__ long long s_warp[kT / 32];
    const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanPer;
    long long v = 0;
#pragma unroll
    for (int j = 0; j < kScanPer; ++j) if (base + j < n) v += in[base + j];
    long long total;
    block_exclusive(v, s_warp, &total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// one CTA: exclusive scan of the tile sums in place; the grand total goes to *total (may be NULL)
__global__ void __launch_bounds__(1024) scan_single_kernel(int64_t* __restrict__ sums, int n_tiles, int64_t* __restrict__ total) {
    __shared__ long long s_warp[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (n_tiles + 1023) / 1024;
    const int lo = tid * per, hi = min(n_tiles, lo + per);
    long long sum = 0;
    for (int i = lo; i < hi; ++i) sum += sums[i];
    long long incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const long long t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        long long w = s_warp[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        s_warp[lane] = wi - w;
        if (lane == 31 && total) *total = wi;
    }
    __syncthreads();
    long long run = s_warp[warp] + incl - sum;
    for (int i = lo; i < hi; ++i) { const long long c = sums[i]; sums[i] = run; run += c; }
}

__global__ void __launch_bounds__(kT) scan_apply_kernel(const int32_t* __restrict__ in, int64_t n,
                                                        const int64_t* __restrict__ tile_off, int64_t* __restrict__ out) {
    __shared__ long long s_warp[kT / 32];
    const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanPer;
    int32_t x[kScanPer];
    long long v = 0;
#pragma unroll
    for (int j = 0; j < kScanPer; ++j) { x[j] = base + j < n ? in[base + j] : 0; v += x[j]; }
    long long run = block_exclusive(v, s_warp, nullptr) + tile_off[blockIdx.x];
#pragma unroll
    for (int j = 0; j < kScanPer; ++j) {
        if (base + j < n) out[base + j] = run;
        run += x[j];
    }
}

int scan_tiles(int64_t n) { return (int)((n + kScanTile - 1) / kScanTile); }

// out[i] = sum of in[0..i), i < n; *total = sum of all (device pointer, may be NULL). tmp: scan_tiles(n) int64.
int exclusive_scan(const int32_t* in, int64_t n, int64_t* out, int64_t* tmp, int64_t* total, cudaStream_t st) {
    if (n <= 0) return LZ_OK;
    const int tiles = scan_tiles(n);
    scan_sums_kernel<<<tiles, kT, 0, st>>>(in, n, tmp);
    scan_single_kernel<<<1, 1024, 0, st>>>(tmp, tiles, total);
    scan_apply_kernel<<<tiles, kT, 0, st>>>(in, n, tmp, out);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

// ---- stable LSD radix sort of the edge endpoints, 8 bits per pass ------------------------------------------
// An entry is a packed pair (key << 32 | val): key = the node an edge end belongs to, val = 2*edge + side.
// Entry i of the unsorted sequence is edge i/2, side i%2 — pass 0 generates it from (src, dst) on the fly
// (GEN), so the unsorted sequence is never written.  Item order inside a tile is index order: warp w owns
// items [w*512, (w+1)*512) and walks them 32 at a time, so (warp, round, lane) order = index order and the
// ranks below are stable.  A tile's items are first placed in shared memory in digit order and then written
// out in runs (all items of one digit go to consecutive addresses), so global writes are coalesced.
constexpr int kRWarps = kT / 32;
constexpr int kRRounds = 16;
constexpr int kRWarpItems = 32 * kRRounds;            // 512 consecutive items per warp
constexpr int kRTile = kRWarps * kRWarpItems;         // 4096 items per CTA

__device__ __forceinline__ u64 gen_entry(const int32_t* __restrict__ src, const int32_t* __restrict__ dst,
                                         int32_t n_nodes, int64_t i) {
    const int64_t e = i >> 1;
    const int32_t a = src[e], b = dst[e];
    const bool ok = a >= 0 && a < n_nodes && b >= 0 && b < n_nodes;
    u32 key = (u32)n_nodes;                                                    // sentinel: sorts behind every node
    if (ok) {
        if ((i & 1) == 0) key = (u32)a;
        else if (b != a) key = (u32)b;                                         // a self-loop keeps only its src end
    }
    return ((u64)key << 32) | (u64)(u32)i;
}

template <bool GEN>
__global__ void __launch_bounds__(kT) radix_hist_kernel(const u64* __restrict__ pairs, const int32_t* __restrict__ src,
                                                        const int32_t* __restrict__ dst, int32_t n_nodes, int64_t n,
                                                        int shift, int32_t* __restrict__ hist, int n_tiles,
                                                        int32_t* __restrict__ deg) {
    __shared__ int h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kRTile;
    // all loads of the tile first (atomics are ordering points the compiler keeps loads behind)
    u32 key[kRTile / kT];
#pragma unroll
    for (int r = 0; r < kRTile / kT; ++r) {
        const int64_t i = base + (int64_t)r * kT + threadIdx.x;
        key[r] = 0xffffffffu;
        if (i < n) key[r] = (u32)((GEN ? gen_entry(src, dst, n_nodes, i) : pairs[i]) >> 32);
    }
#pragma unroll
    for (int r = 0; r < kRTile / kT; ++r) {
        const int64_t i = base + (int64_t)r * kT + threadIdx.x;
        if (i < n) {
            atomicAdd(&h[(key[r] >> shift) & 255u], 1);
            if (GEN && key[r] != (u32)n_nodes) atomicAdd(deg + key[r], 1);    // degrees, once, while generating
        }
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];       // digit-major: a flat scan gives the bases
}

template <bool GEN, bool LAST>
__global__ void __launch_bounds__(kT) radix_scatter_kernel(const u64* __restrict__ pairs_in, const int32_t* __restrict__ src,
                                                           const int32_t* __restrict__ dst, int32_t n_nodes,
                                                           u64* __restrict__ pairs_out, u32* __restrict__ vals_out,
                                                           int64_t n, int shift, const int64_t* __restrict__ base,
                                                           int n_tiles) {
    __shared__ u64 stage[kRTile];                                             // 32 KB: the tile in digit order
    __shared__ int wcnt[kRWarps][256];                                        // per (warp, digit): count, then local start
    __shared__ long long gdelta[256];                                         // global position - local position, per digit
    __shared__ int s_warp_tot[kRWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < kRWarps * 256; i += kT) (&wcnt[0][0])[i] = 0;
    __syncthreads();
    const int64_t start = (int64_t)blockIdx.x * kRTile + (int64_t)warp * kRWarpItems;
    u64 pr[kRRounds];
    int rank[kRRounds];
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kRRounds; ++r) {
        const int64_t i = start + r * 32 + lane;
        const bool valid = i < n;
        pr[r] = valid ? (GEN ? gen_entry(src, dst, n_nodes, i) : pairs_in[i]) : 0ull;
        const unsigned live = __ballot_sync(0xffffffffu, valid);
        rank[r] = 0;
        if (valid) {
            const int d = (int)((pr[r] >> (32 + shift)) & 255ull);
            // lanes holding the same digit: eight votes, one per digit bit (measured: MATCH.ANY's result latency
            // made this loop the kernel's bound — 28 short-scoreboard stall cycles per issue)
            unsigned peers = live;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const unsigned m = __ballot_sync(live, (d >> b) & 1);
                peers &= ((d >> b) & 1) ? m : ~m;
            }
            const int leader = __ffs(peers) - 1;
            int old = 0;
            if (lane == leader) { old = wcnt[warp][d]; wcnt[warp][d] = old + __popc(peers); }
            old = __shfl_sync(peers, old, leader);
            rank[r] = old + __popc(peers & lt);
        }
        __syncwarp();
    }
    __syncthreads();
    {
        // thread d: the digit's count over the warps, an exclusive scan over digits for its start in the tile,
        // then every warp's start inside the digit's run
        const int d = threadIdx.x;
        int tot = 0;
#pragma unroll
        for (int w = 0; w < kRWarps; ++w) tot += wcnt[w][d];
        int incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp_tot[warp] = incl;
        __syncthreads();
        int before = 0;
#pragma unroll
        for (int w = 0; w < kRWarps; ++w) if (w < warp) before += s_warp_tot[w];
        int run = before + incl - tot;                                        // first tile-local slot of digit d
        gdelta[d] = base[(size_t)d * n_tiles + blockIdx.x] - run;
#pragma unroll
        for (int w = 0; w < kRWarps; ++w) { const int c = wcnt[w][d]; wcnt[w][d] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kRRounds; ++r) {
        const int64_t i = start + r * 32 + lane;
        if (i < n) {
            const int d = (int)((pr[r] >> (32 + shift)) & 255ull);
            stage[wcnt[warp][d] + rank[r]] = pr[r];
        }
    }
    __syncthreads();
    const int n_here = (int)min((int64_t)kRTile, n - (int64_t)blockIdx.x * kRTile);
    for (int j = threadIdx.x; j < n_here; j += kT) {
        const u64 v = stage[j];
        const long long pos = gdelta[(int)((v >> (32 + shift)) & 255ull)] + j;
        if (LAST) vals_out[pos] = (u32)v;
        else pairs_out[pos] = v;
    }
}

int radix_tiles(int64_t n) { return (int)((n + kRTile - 1) / kRTile); }

int key_passes(int32_t n_nodes) {
    int bits = 1;
    while (bits < 32 && (1ll << bits) <= (long long)n_nodes) ++bits;           // keys go up to n_nodes (the sentinel)
    return (bits + 7) / 8;
}

struct CsrLayout { size_t pairs0, pairs1, deg, hist, hist_off, tmp, total; };

CsrLayout csr_layout(int64_t n_edges, int32_t n_nodes) {
    CsrLayout L;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t at = off; off = (off + bytes + 255) & ~(size_t)255; return at; };
    const size_t n_ent = (size_t)2 * n_edges;
    const size_t rt = (size_t)radix_tiles((int64_t)n_ent);
    const int passes = key_passes(n_nodes);
    L.pairs0 = take(passes >= 2 ? n_ent * 8 : 0);
    L.pairs1 = take(passes >= 3 ? n_ent * 8 : 0);
    L.deg = take(((size_t)n_nodes + 1) * 4);
    L.hist = take(256 * rt * 4);
    L.hist_off = take(256 * rt * 8);
    const size_t st_hist = (size_t)scan_tiles((int64_t)(256 * rt)), st_deg = (size_t)scan_tiles((int64_t)n_nodes + 1);
    L.tmp = take((st_hist > st_deg ? st_hist : st_deg) * 8 + 8);
    L.total = off;
    return L;
}

__global__ void __launch_bounds__(kT) neighbors_kernel(const int64_t* __restrict__ offsets, const u32* __restrict__ adj,
                                                       const int32_t* __restrict__ src, const int32_t* __restrict__ dst,
                                                       const float* __restrict__ weight, int32_t n_nodes,
                                                       const int32_t* __restrict__ nodes, int n_queries, double min_weight,
                                                       int max_out, int32_t* __restrict__ out_nbr, int32_t* __restrict__ out_edge,
                                                       int32_t* __restrict__ out_count) {
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * (kT / 32) + (threadIdx.x >> 5);
    if (q >= n_queries) return;
    const int32_t v = nodes[q];
    int cnt = 0;
    if (v >= 0 && v < n_nodes) {
        const int64_t lo = offsets[v], hi = offsets[v + 1];
        const unsigned lt = (1u << lane) - 1u;
        for (int64_t base = lo; base < hi; base += 32) {
            const int64_t i = base + lane;
            bool ok = false;
            int32_t nb = -1, e = -1;
            if (i < hi) {
                const u32 ent = adj[i];
                e = (int32_t)(ent >> 1);
                ok = (double)weight[e] >= min_weight;                         // the reference compares float64 >= float64
                nb = (ent & 1u) ? src[e] : dst[e];
            }
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                const int pos = cnt + __popc(m & lt);
                if (pos < max_out) {
                    out_nbr[(size_t)q * max_out + pos] = nb;
                    if (out_edge) out_edge[(size_t)q * max_out + pos] = e;
                }
            }
            cnt += __popc(m);
        }
    }
    if (lane == 0) out_count[q] = cnt;
}

}  // namespace

extern "C" size_t lz_csr_build_scratch_bytes(int64_t n_edges, int32_t n_nodes) {
    if (n_edges <= 0 || n_nodes <= 0) return 0;
    return csr_layout(n_edges, n_nodes).total;
}

extern "C" int lz_csr_build(const int32_t* src, const int32_t* dst, int64_t n_edges, int32_t n_nodes, void* scratch,
                            size_t scratch_bytes, int64_t* offsets, uint32_t* adj, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    LZ_REQUIRE(n_nodes >= 0 && n_edges >= 0, "lz_csr_build: negative size");
    LZ_REQUIRE(offsets != nullptr, "lz_csr_build: offsets is NULL");
    if (n_edges == 0 || n_nodes == 0) {
        LZ_CUDA(cudaMemsetAsync(offsets, 0, ((size_t)n_nodes + 1) * 8, st));
        return LZ_OK;
    }
    LZ_REQUIRE(src && dst && adj && scratch, "lz_csr_build: NULL pointer");
    LZ_REQUIRE(n_edges <= 0x3fffffffLL, "lz_csr_build: n_edges=%lld exceeds 2^30 - 1", (long long)n_edges);
    const CsrLayout L = csr_layout(n_edges, n_nodes);
    if (scratch_bytes < L.total) {
        lz_set_error("lz_csr_build: scratch has %zu bytes, needs %zu", scratch_bytes, L.total);
        return LZ_ERR_SCRATCH;
    }
    unsigned char* s = (unsigned char*)scratch;
    const int64_t n_ent = 2 * n_edges;
    const int passes = key_passes(n_nodes);
    u64* pbuf[2] = {(u64*)(s + L.pairs0), (u64*)(s + L.pairs1)};
    int32_t* deg = (int32_t*)(s + L.deg);
    int32_t* hist = (int32_t*)(s + L.hist);
    int64_t* hist_off = (int64_t*)(s + L.hist_off);
    int64_t* tmp = (int64_t*)(s + L.tmp);
    LZ_CUDA(cudaMemsetAsync(deg, 0, ((size_t)n_nodes + 1) * 4, st));
    const int rt = radix_tiles(n_ent);
    for (int pass = 0; pass < passes; ++pass) {
        const bool first = pass == 0, last = pass == passes - 1;
        const u64* in = first ? nullptr : pbuf[(pass - 1) & 1];
        u64* out = last ? nullptr : pbuf[pass & 1];
        const int shift = pass * 8;
        if (first) radix_hist_kernel<true><<<rt, kT, 0, st>>>(in, src, dst, n_nodes, n_ent, shift, hist, rt, deg);
        else radix_hist_kernel<false><<<rt, kT, 0, st>>>(in, src, dst, n_nodes, n_ent, shift, hist, rt, deg);
        if (int rc = exclusive_scan(hist, (int64_t)256 * rt, hist_off, tmp, nullptr, st)) return rc;
        if (first && last) radix_scatter_kernel<true, true><<<rt, kT, 0, st>>>(in, src, dst, n_nodes, out, adj, n_ent, shift, hist_off, rt);
        else if (first) radix_scatter_kernel<true, false><<<rt, kT, 0, st>>>(in, src, dst, n_nodes, out, adj, n_ent, shift, hist_off, rt);
        else if (last) radix_scatter_kernel<false, true><<<rt, kT, 0, st>>>(in, src, dst, n_nodes, out, adj, n_ent, shift, hist_off, rt);
        else radix_scatter_kernel<false, false><<<rt, kT, 0, st>>>(in, src, dst, n_nodes, out, adj, n_ent, shift, hist_off, rt);
        LZ_CUDA(cudaGetLastError());
    }
    // offsets[v] = first entry of node v; offsets[n_nodes] = number of real entries (deg[n_nodes] stays 0)
    return exclusive_scan(deg, (int64_t)n_nodes + 1, offsets, tmp, nullptr, st);
}

extern "C" int lz_neighbors(const int64_t* offsets, const uint32_t* adj, const int32_t* src, const int32_t* dst,
                            const float* weight, int32_t n_nodes, const int32_t* nodes, int n_queries, double min_weight,
                            int max_out, int32_t* out_nbr, int32_t* out_edge, int32_t* out_count, void* stream) {
    if (n_queries <= 0) return LZ_OK;
    LZ_REQUIRE(offsets && nodes && out_count, "lz_neighbors: NULL pointer");
    LZ_REQUIRE(max_out >= 0 && (max_out == 0 || out_nbr != nullptr), "lz_neighbors: max_out=%d without an output buffer", max_out);
    LZ_REQUIRE(n_nodes >= 0, "lz_neighbors: n_nodes=%d", n_nodes);
    neighbors_kernel<<<(n_queries + kT / 32 - 1) / (kT / 32), kT, 0, (cudaStream_t)stream>>>(
        offsets, adj, src, dst, weight, n_nodes, nodes, n_queries, min_weight, max_out, out_nbr, out_edge, out_count);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

// ---- edge lookup and reinforcement (stands behind MemoryShard.add_edge, core/memory_shard.py:42-52) ----------
// add_edge keys its dict by (source, target): an existing key is reinforced (weight = min(1, weight + 0.1),
// co_occurrence += 1), a new one is inserted.  The lookup runs on the CSR: a warp walks the source node's
// entries and reports the edge whose source end it is and whose target matches (the lowest such edge index).
namespace {
__global__ void __launch_bounds__(kT) edge_find_kernel(const int64_t* __restrict__ offsets, const u32* __restrict__ adj,
                                                       const int32_t* __restrict__ dst, int32_t n_nodes,
                                                       const int32_t* __restrict__ q_src, const int32_t* __restrict__ q_dst,
                                                       int n_queries, int32_t* __restrict__ out_edge) {
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * (kT / 32) + (threadIdx.x >> 5);
    if (q >= n_queries) return;
    const int32_t a = q_src[q], b = q_dst[q];
    int32_t found = -1;
    if (a >= 0 && a < n_nodes) {
        const int64_t lo = offsets[a], hi = offsets[a + 1];
        for (int64_t base = lo; base < hi && found < 0; base += 32) {
            const int64_t i = base + lane;
            bool hit = false;
            u32 ent = 0u;
            if (i < hi) {
                ent = adj[i];
                hit = (ent & 1u) == 0u && dst[ent >> 1] == b;                 // a's source end of an edge to b
            }
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (m) found = (int32_t)(__shfl_sync(0xffffffffu, ent, __ffs(m) - 1) >> 1);   // entries ascend by edge
        }
    }
    if (lane == 0) out_edge[q] = found;
}

__global__ void __launch_bounds__(kT) edge_reinforce_kernel(float* __restrict__ weight, int32_t* __restrict__ cooc,
                                                            const int32_t* __restrict__ edge_idx,
                                                            const int32_t* __restrict__ times, int64_t m, int64_t n_edges) {
    const int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    if (i >= m) return;
    const int32_t e = edge_idx[i];
    const int t = times ? times[i] : 1;
    if (e < 0 || e >= n_edges || t <= 0) return;
    float w = weight[e];
    for (int r = 0; r < t; ++r) w = (float)fmin(1.0, __dadd_rn((double)w, 0.1));   // float64 add, stored float32
    weight[e] = w;
    cooc[e] += t;
}
}  // namespace

extern "C" int lz_edge_find(const int64_t* offsets, const uint32_t* adj, const int32_t* dst, int32_t n_nodes,
                            const int32_t* q_src, const int32_t* q_dst, int n_queries, int32_t* out_edge, void* stream) {
    if (n_queries <= 0) return LZ_OK;
    LZ_REQUIRE(offsets && q_src && q_dst && out_edge, "lz_edge_find: NULL pointer");
    LZ_REQUIRE(n_nodes >= 0, "lz_edge_find: n_nodes=%d", n_nodes);
    edge_find_kernel<<<(n_queries + kT / 32 - 1) / (kT / 32), kT, 0, (cudaStream_t)stream>>>(
        offsets, adj, dst, n_nodes, q_src, q_dst, n_queries, out_edge);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

extern "C" int lz_edge_reinforce(float* weight, int32_t* co_occurrence, int64_t n_edges, const int32_t* edge_idx,
                                 const int32_t* times, int64_t m, void* stream) {
    if (m <= 0) return LZ_OK;
    LZ_REQUIRE(weight && co_occurrence && edge_idx, "lz_edge_reinforce: NULL pointer");
    edge_reinforce_kernel<<<(unsigned)((m + kT - 1) / kT), kT, 0, (cudaStream_t)stream>>>(weight, co_occurrence, edge_idx,
                                                                                         times, m, n_edges);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}
