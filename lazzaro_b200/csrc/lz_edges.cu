// K7: biological decay + threshold prune over the edge table (struct of arrays), one pass.
//
// Stands behind MemoryShard.apply_temporal_decay / prune_weak_edges (core/memory_shard.py:64-84).
// HBM-bound integer/float streaming: 16 B read per edge (src, dst, weight, co_occurrence) + 16 B written per
// surviving edge.  Survivors are compacted stably in a single pass by the streamed kernel below (bulk copies into
// a shared-memory ring, survivor counts published ahead of the data path, coalesced write-out); the earlier forms
// (chained look-back, count + scatter in two passes) stay selectable with LZ_PRUNE_FORM for comparison.
#include "lz_common.cuh"
#include "lz_sm100.cuh"

#include <cmath>
#include <mutex>
#include <stdlib.h>

namespace {

constexpr int kThreads = 256;
constexpr int kItems = 8;                       // edges per thread (two 128-bit loads per array)
constexpr int kTile = kThreads * kItems;        // 2048 edges per tile

constexpr int kGroup = 64;                      // tiles per group of the two-pass form's offset sums

constexpr u64 kFlagAgg = 1ull << 62;            // tile aggregate published
constexpr u64 kFlagPrefix = 2ull << 62;         // inclusive prefix published
constexpr u64 kValueMask = (1ull << 62) - 1;

__device__ __forceinline__ u64 ld_state(const u64* p) {
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_state(u64* p, u64 v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Coherent (not .nc) streaming load: outputs may alias inputs, so the read-only path is off limits.
__device__ __forceinline__ int4 ld_stream_i4(const void* p) {
    int4 r;
    asm volatile("ld.global.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}

struct PruneParams {
    const int32_t* src; const int32_t* dst; const float* w; const int32_t* cooc;
    int32_t* o_src; int32_t* o_dst; float* o_w; int32_t* o_cooc;
    int64_t n; double factor; double thr;
    u64* state;        // [n_tiles] look-back words, zeroed before launch
    u32* ticket;       // zeroed before launch
    int64_t* out_kept;
    int n_tiles;
    const int64_t* offsets;   // two-pass mode with the scan kernel: exclusive prefix of the per-tile survivor counts
    const int32_t* counts;    // two-pass mode: survivors per tile ...
    const u64* group_sum;     // ... and per group of kGroup tiles; a CTA sums what lies before it (offsets == nullptr)
    float wstar; int use_wstar;   // factor > 0: keep <=> not (w < wstar), see prune_bound()
    int32_t* loaded;              // streamed form, outputs aliasing inputs: [n_tiles] set once a tile's columns are on chip
    int check_loaded;
};

// CHAINED = true: single pass, tile order from an atomic ticket, offsets by decoupled look-back (safe in
// place).  CHAINED = false: second pass of the two-pass form, offsets precomputed (no inter-CTA waiting).
template <bool CHAINED>
__global__ void __launch_bounds__(kThreads) decay_prune_kernel(PruneParams p) {
    __shared__ int s_src[kTile];
    __shared__ int s_dst[kTile];
    __shared__ float s_w[kTile];
    __shared__ int s_cooc[kTile];
    __shared__ int s_warp[kThreads / 32];
    __shared__ long long s_pre[kThreads / 32];
    __shared__ u32 s_tile;
    __shared__ long long s_prefix;
    __shared__ int s_total;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (CHAINED) {
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
    }
    const int tile = CHAINED ? (int)s_tile : (int)blockIdx.x;
    const int64_t base = (int64_t)tile * kTile + (int64_t)tid * kItems;

    int v_src[kItems], v_dst[kItems], v_cooc[kItems];
    float v_w[kItems];
    if (base + kItems <= p.n) {
#pragma unroll
        for (int h = 0; h < kItems / 4; ++h) {
            int4 a = ld_stream_i4(p.src + base + 4 * h);
            int4 b = ld_stream_i4(p.dst + base + 4 * h);
            int4 c = ld_stream_i4(p.w + base + 4 * h);
            int4 d = ld_stream_i4(p.cooc + base + 4 * h);
            v_src[4 * h] = a.x; v_src[4 * h + 1] = a.y; v_src[4 * h + 2] = a.z; v_src[4 * h + 3] = a.w;
            v_dst[4 * h] = b.x; v_dst[4 * h + 1] = b.y; v_dst[4 * h + 2] = b.z; v_dst[4 * h + 3] = b.w;
            v_w[4 * h] = __int_as_float(c.x); v_w[4 * h + 1] = __int_as_float(c.y);
            v_w[4 * h + 2] = __int_as_float(c.z); v_w[4 * h + 3] = __int_as_float(c.w);
            v_cooc[4 * h] = d.x; v_cooc[4 * h + 1] = d.y; v_cooc[4 * h + 2] = d.z; v_cooc[4 * h + 3] = d.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
            const bool in = base + j < p.n;
            v_src[j] = in ? p.src[base + j] : 0;
            v_dst[j] = in ? p.dst[base + j] : 0;
            v_w[j] = in ? p.w[base + j] : 0.f;
            v_cooc[j] = in ? p.cooc[base + j] : 0;
        }
    }
    // two-pass form without the scan kernel: what lies before this tile = the sums of the groups before its group +
    // the counts of the tiles before it in its group (a handful of L2 loads per thread, in flight beside the columns)
    long long before = 0;
    if (!CHAINED && p.offsets == nullptr) {
        const int g = tile / kGroup;
        for (int i = tid; i < g; i += kThreads) before += (long long)p.group_sum[i];
        if (tid < tile - g * kGroup) before += p.counts[g * kGroup + tid];
    }
    // decay in float64, strict '<' prune in float64, store the float32 rounding
    unsigned keep = 0;
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        const double w64 = (double)v_w[j] * p.factor;
        const bool k = (base + j < p.n) && !(w64 < p.thr);
        v_w[j] = (float)w64;
        keep |= (unsigned)k << j;
        cnt += k;
    }
    // block-wide exclusive scan of cnt
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    if (!CHAINED && p.offsets == nullptr) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
        if (lane == 0) s_pre[warp] = before;
    }
    __syncthreads();
    if (warp == 0) {
        int ws = lane < kThreads / 32 ? s_warp[lane] : 0;
        int wi = ws;
#pragma unroll
        for (int o = 1; o < kThreads / 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        if (lane < kThreads / 32) s_warp[lane] = wi - ws;       // exclusive warp offsets
        if (lane == kThreads / 32 - 1) s_total = wi;
    }
    __syncthreads();
    // Tile-local compaction into shared memory.  This consumes every loaded value, so once the
    // barrier below is passed all global loads of this tile have completed — which is what makes
    // in-place operation safe: a later tile may overwrite this tile's input range only after it has
    // seen this tile's published aggregate, and that is published after the barrier.
    int off = s_warp[warp] + incl - cnt;
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        if (keep >> j & 1) {
            s_src[off] = v_src[j]; s_dst[off] = v_dst[j]; s_w[off] = v_w[j]; s_cooc[off] = v_cooc[j];
            ++off;
        }
    }
    __syncthreads();
    if (!CHAINED) {
        if (tid == 0) {
            if (p.offsets != nullptr) {
                s_prefix = p.offsets[tile];
            } else {
                long long b = 0;
#pragma unroll
                for (int i = 0; i < kThreads / 32; ++i) b += s_pre[i];
                s_prefix = b;
                if (tile == p.n_tiles - 1) *p.out_kept = b + s_total;
            }
        }
    } else if (warp == 0) {
        const int total = s_total;
        // decoupled look-back: publish the aggregate, then walk predecessors 32 at a time
        long long excl = 0;
        if (tile == 0) {
            if (lane == 0) st_state(p.state, kFlagPrefix | (u64)total);
        } else {
            if (lane == 0) st_state(p.state + tile, kFlagAgg | (u64)total);
            int look = tile - 1;
            while (true) {
                const int idx = look - lane;
                u64 st = idx >= 0 ? ld_state(p.state + idx) : kFlagPrefix;     // before tile 0: prefix 0
                const unsigned empty = __ballot_sync(0xffffffffu, (st >> 62) == 0);
                if (empty) continue;                                           // someone not published yet
                const unsigned pref = __ballot_sync(0xffffffffu, (st >> 62) == 2);
                const int first = pref ? __ffs(pref) - 1 : 32;                 // nearest tile with a full prefix
                long long v = lane <= first ? (long long)(st & kValueMask) : 0;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                excl += v;
                if (pref) break;
                look -= 32;
            }
            if (lane == 0) st_state(p.state + tile, kFlagPrefix | (u64)(excl + total));
        }
        if (lane == 0) {
            s_prefix = excl;
            if (tile == p.n_tiles - 1) *p.out_kept = excl + total;
        }
    }
    __syncthreads();
    const int total = s_total;
    const int64_t out0 = s_prefix;
    for (int i = tid; i < total; i += kThreads) {
        p.o_src[out0 + i] = s_src[i];
        p.o_dst[out0 + i] = s_dst[i];
        p.o_w[out0 + i] = s_w[i];
        p.o_cooc[out0 + i] = s_cooc[i];
    }
}

// ---- streamed single-pass form -------------------------------------------------------------------
// One pass, every column read from HBM once, and no warp on the data path ever waits for another CTA.
// A CTA is 16 compute warps plus three service warps and works through tiles whose numbers it draws from the
// ticket, with a ring of `stages` tile buffers in shared memory and a small queue of tile records:
//   count warp    runs `ahead` tiles in front of the data path: draws the next ticket, fetches that tile's weights
//                 (8 KB bulk copy, L2 evict-last), counts the survivors and publishes the tile's aggregate
//   offset warp   turns aggregates into output offsets.  Tiles form waves of kWave consecutive tickets;
//                 offset(t) = what lies before t's wave + the aggregates of the tiles before t in its wave (at most
//                 kWave/32 loads per lane, all issued at once).  The CTA holding a wave's last tile publishes the
//                 wave's total as soon as it knows it and the inclusive prefix once it has its own offset; "what lies
//                 before the wave" is a 32-wide look-back over those wave words, so no offset waits on a chain of
//                 CTAs — only on counts, and those were published `ahead` tiles early
//   loader warp   issues each tile's four column segments as plain bulk copies (cp.async.bulk global -> shared,
//                 completion on the stage's mbarrier) as soon as a stage is free; the weights come out of the L2
//   compute warps flags from shared memory, block scan, compaction in place in the stage buffer, coalesced
//                 write-out at the tile's offset, stage handed back
// Ticket order keeps it deadlock-free: a tile waits only for tiles with smaller tickets, and those are held by
// CTAs that are running.  When outputs alias inputs, a tile's survivors land in the input ranges of at most two tiles
// (itself or earlier ones); the compute warps flag a tile once its columns are in shared memory and the offset warp
// releases an offset only when those two tiles are flagged (the count warp has read a tile's weights before its
// aggregate is public, and nobody writes before they have seen every earlier aggregate).  AHEAD = false is the
// earlier in-place form: the compute warps publish the aggregate after reading the stage, the count warp only draws
// tickets (LZ_PRUNE_INPLACE_OLD=1).
constexpr int kMaxStages = 8;
constexpr int kStreamThreads = 512;             // compute threads per CTA of the streamed form (4 edges each)
constexpr int kStreamWarps = kStreamThreads / 32 + 3;
constexpr int kWave = 256;                      // tiles per wave of the offset scheme
constexpr int kQueue = 16;                      // tile records per CTA (>= ahead + 2)

struct TileRec { int tile; int total; long long prefix; };

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(sm100::smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(sm100::smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void named_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ int ld_vol(const int* p) { return *reinterpret_cast<const volatile int*>(p); }
__device__ __forceinline__ void st_vol(int* p, int v) { *reinterpret_cast<volatile int*>(p) = v; }
__device__ __forceinline__ void wait_above(const int* counter, int n) {        // until *counter > n
    while (ld_vol(counter) <= n) __nanosleep(20);
    __threadfence_block();
}

// keep = not (float64(w) * factor < thr).  For factor > 0 that is monotone in w, so it equals not (w < wstar) for a
// float32 bound found on the host; the float64 form stays for every other factor.
__device__ __forceinline__ bool keep_edge(float w, const PruneParams& p) {
    return p.use_wstar ? !(w < p.wstar) : !((double)w * p.factor < p.thr);
}

template <int THREADS, int ITEMS, bool AHEAD>
__global__ void __launch_bounds__(THREADS + 96, 2) prune_stream_kernel(PruneParams p, int stages, int ahead) {
    constexpr int TILE = THREADS * ITEMS;
    constexpr int NW = THREADS / 32;
    static_assert(TILE == kTile && NW <= 32 && ITEMS % 4 == 0, "tile geometry");
    extern __shared__ __align__(128) unsigned char s_ring[];          // stages x [4 columns][TILE] words + 2 x TILE weights of the count warp
    __shared__ uint64_t full[kMaxStages], empty[kMaxStages], cbar[2];
    __shared__ TileRec queue[kQueue];
    __shared__ int q_drawn, q_counted, q_offset, q_scattered;          // how many records each role has finished
    __shared__ int s_tile[kMaxStages];
    __shared__ int s_warp[2][NW];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        for (int s = 0; s < stages; ++s) { sm100::mbar_init(&full[s], 1); sm100::mbar_init(&empty[s], NW); }
        sm100::mbar_init(&cbar[0], 1); sm100::mbar_init(&cbar[1], 1);
        sm100::fence_mbar_init();
        q_drawn = q_counted = q_offset = q_scattered = 0;
    }
    __syncthreads();

    if (warp == NW) {
        // ---------------- count warp ----------------
        float* cw = reinterpret_cast<float*>(s_ring + (size_t)stages * 16 * TILE);      // two buffers of TILE weights
        const uint64_t keep_in_l2 = policy_evict_last();
        // draws the ticket of record n and starts the copy of that tile's weights; -1 past the last tile
        auto fetch = [&](int n) -> int {
            while (n - ld_vol(&q_scattered) >= ahead) __nanosleep(20);
            int tile = 0;
            if (lane == 0) tile = (int)atomicAdd(p.ticket, 1u);
            tile = __shfl_sync(0xffffffffu, tile, 0);
            if (tile >= p.n_tiles) return -1;
            if (AHEAD && lane == 0 && (int64_t)(tile + 1) * TILE <= p.n) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                sm100::mbar_expect_tx(&cbar[n & 1], 4u * TILE);
                bulk_load(cw + (n & 1) * TILE, p.w + (int64_t)tile * TILE, 4u * TILE, &cbar[n & 1], keep_in_l2);
            }
            return tile;
        };
        int tile = fetch(0);
        for (int n = 0;; ++n) {
            TileRec* e = &queue[n % kQueue];
            if (lane == 0) {
                e->tile = tile;
                __threadfence_block();
                st_vol(&q_drawn, n + 1);                                   // the loader may fetch the tile from here on
                if (tile < 0 && AHEAD) st_vol(&q_counted, n + 1);          // aliased call: the compute warps pass the end mark on
            }
            if (tile < 0) break;
            const int next = fetch(n + 1);                                 // in flight while this tile is counted
            if (AHEAD) {
                const int64_t base = (int64_t)tile * TILE;
                int cnt = 0;
                if (base + TILE <= p.n) {
                    sm100::mbar_wait(&cbar[n & 1], (n >> 1) & 1);
                    const float4* src = reinterpret_cast<const float4*>(cw + (n & 1) * TILE);
#pragma unroll 4
                    for (int r = 0; r < TILE / 128; ++r) {
                        const float4 v = src[r * 32 + lane];
                        cnt += keep_edge(v.x, p) + keep_edge(v.y, p) + keep_edge(v.z, p) + keep_edge(v.w, p);
                    }
                } else {
                    for (int i = lane; base + i < p.n; i += 32) cnt += keep_edge(p.w[base + i], p);
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
                if (lane == 0) {
                    e->total = cnt;
                    st_state(p.state + tile, kFlagAgg | (u64)cnt);
                    __threadfence_block();
                    st_vol(&q_counted, n + 1);
                }
            }
            __syncwarp();
            tile = next;
        }
        return;
    }
    if (warp == NW + 1) {
        // ---------------- loader warp ----------------
        const uint64_t stream_through = policy_evict_first();
        for (int n = 0;; ++n) {
            wait_above(&q_drawn, n);
            const int tile = queue[n % kQueue].tile;
            const int s = n % stages;
            if (n >= stages) sm100::mbar_wait(&empty[s], ((n / stages) - 1) & 1);
            int* buf = reinterpret_cast<int*>(s_ring) + (size_t)s * 4 * TILE;
            const int64_t base = (int64_t)tile * TILE;
            if (tile < 0) {
                if (lane == 0) { s_tile[s] = -1; sm100::mbar_arrive(&full[s]); }
                break;
            }
            if (base + TILE <= p.n) {
                if (lane == 0) {
                    s_tile[s] = tile;
                    sm100::mbar_expect_tx(&full[s], 16u * TILE);
                    bulk_load(buf + 2 * TILE, p.w + base, 4u * TILE, &full[s], stream_through);
                    bulk_load(buf, p.src + base, 4u * TILE, &full[s], stream_through);
                    bulk_load(buf + TILE, p.dst + base, 4u * TILE, &full[s], stream_through);
                    bulk_load(buf + 3 * TILE, p.cooc + base, 4u * TILE, &full[s], stream_through);
                }
            } else {                                   // the ragged last tile: guarded loads by this warp
                for (int i = lane; i < TILE; i += 32) {
                    const bool in = base + i < p.n;
                    buf[i] = in ? p.src[base + i] : 0;
                    buf[TILE + i] = in ? p.dst[base + i] : 0;
                    buf[2 * TILE + i] = in ? __float_as_int(p.w[base + i]) : 0;
                    buf[3 * TILE + i] = in ? p.cooc[base + i] : 0;
                }
                __syncwarp();
                if (lane == 0) { s_tile[s] = tile; sm100::mbar_arrive(&full[s]); }
            }
            __syncwarp();
        }
        return;
    }
    if (warp == NW + 2) {
        // ---------------- offset warp ----------------
        u64* wave_word = p.state + p.n_tiles;      // [n_waves]: a wave's total, later its inclusive prefix (flagged)
        for (int n = 0;; ++n) {
            wait_above(&q_counted, n);
            TileRec* e = &queue[n % kQueue];
            const int tile = e->tile;
            if (tile < 0) break;
            const int total = e->total;
            long long sum;
            if (p.offsets != nullptr) {                               // measurement mode: offsets precomputed
                sum = p.offsets[tile];
            } else {
                const int w = tile / kWave, j = tile % kWave;
                const u64* agg = p.state + (size_t)w * kWave;
                u64 v[kWave / 32];
#pragma unroll
                for (int r = 0; r < kWave / 32; ++r) {
                    const int idx = j - 1 - lane - 32 * r;
                    v[r] = idx >= 0 ? ld_state(agg + idx) : kFlagAgg;
                }
                u64 ws = w - 1 - lane >= 0 ? ld_state(wave_word + (w - 1 - lane)) : kFlagPrefix;   // in flight with the above
                while (true) {
                    bool pending = false;
#pragma unroll
                    for (int r = 0; r < kWave / 32; ++r) pending = pending || (v[r] >> 62) == 0;
                    if (!__any_sync(0xffffffffu, pending)) break;
                    __nanosleep(40);
#pragma unroll
                    for (int r = 0; r < kWave / 32; ++r)
                        if ((v[r] >> 62) == 0) v[r] = ld_state(agg + (j - 1 - lane - 32 * r));
                }
                sum = 0;
#pragma unroll
                for (int r = 0; r < kWave / 32; ++r) sum += (long long)(v[r] & kValueMask);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                // the wave's last tile knows the wave's total now: publish it before its own offset is known, so that
                // later waves never wait on a chain of CTAs, only on counts (which run ahead)
                const bool closes_wave = j == kWave - 1;
                if (closes_wave && lane == 0) st_state(wave_word + w, kFlagAgg | (u64)(sum + total));
                // look-back over the wave words: totals back to the nearest published inclusive prefix
                long long before = 0;
                for (int look = w - 1;; look -= 32) {
                    unsigned pref, empty;
                    int first;
                    while (true) {
                        pref = __ballot_sync(0xffffffffu, (ws >> 62) == 2);
                        first = pref ? __ffs(pref) - 1 : 31;
                        empty = __ballot_sync(0xffffffffu, (ws >> 62) == 0) & (0xffffffffu >> (31 - first));
                        if (!empty) break;
                        __nanosleep(40);
                        if ((ws >> 62) == 0) ws = ld_state(wave_word + (look - lane));
                    }
                    long long t = lane <= first ? (long long)(ws & kValueMask) : 0;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
                    before += t;
                    if (pref) break;
                    ws = look - 32 - lane >= 0 ? ld_state(wave_word + (look - 32 - lane)) : kFlagPrefix;
                }
                if (closes_wave && lane == 0) st_state(wave_word + w, kFlagPrefix | (u64)(before + sum + total));
                sum += before;
            }
            if (AHEAD && p.check_loaded && total > 0 && lane < 2) {
                // in place: the survivors land in the input ranges of at most two tiles (both <= this one); those must
                // be on chip in whichever CTA holds them before anything is written over them
                const long long t_in = (lane == 0 ? sum : sum + total - 1) / TILE;
                unsigned f = 0;
                while (true) {
                    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(f) : "l"(p.loaded + t_in) : "memory");
                    if (f != 0u) break;
                    __nanosleep(40);
                }
            }
            __syncwarp();
            if (lane == 0) {
                e->prefix = sum;
                if (tile == p.n_tiles - 1) *p.out_kept = sum + total;
                __threadfence_block();
                st_vol(&q_offset, n + 1);
            }
            __syncwarp();
        }
        return;
    }

    // ---------------- compute warps ----------------
    for (int it = 0;; ++it) {
        const int s = it % stages, par = it & 1;
        sm100::mbar_wait(&full[s], (it / stages) & 1);
        const int tile = s_tile[s];
        if (tile < 0) {
            if (!AHEAD && tid == 0) st_vol(&q_counted, it + 1);
            break;
        }
        int* buf = reinterpret_cast<int*>(s_ring) + (size_t)s * 4 * TILE;
        // in place: from here on nobody needs this tile's columns in global memory any more (the count warp read the
        // weights before it published the aggregate, and nobody writes before they have seen that)
        if (AHEAD && p.check_loaded && tid == 0)
            asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p.loaded + tile), "r"(1u) : "memory");
        const int64_t base = (int64_t)tile * TILE + (int64_t)tid * ITEMS;
        int v_src[ITEMS], v_dst[ITEMS], v_cooc[ITEMS];
        float v_w[ITEMS];
#pragma unroll
        for (int h = 0; h < ITEMS / 4; ++h) {
            const int4 a = reinterpret_cast<const int4*>(buf)[tid * (ITEMS / 4) + h];
            const int4 b = reinterpret_cast<const int4*>(buf + TILE)[tid * (ITEMS / 4) + h];
            const int4 c = reinterpret_cast<const int4*>(buf + 2 * TILE)[tid * (ITEMS / 4) + h];
            const int4 d = reinterpret_cast<const int4*>(buf + 3 * TILE)[tid * (ITEMS / 4) + h];
            v_src[4 * h] = a.x; v_src[4 * h + 1] = a.y; v_src[4 * h + 2] = a.z; v_src[4 * h + 3] = a.w;
            v_dst[4 * h] = b.x; v_dst[4 * h + 1] = b.y; v_dst[4 * h + 2] = b.z; v_dst[4 * h + 3] = b.w;
            v_w[4 * h] = __int_as_float(c.x); v_w[4 * h + 1] = __int_as_float(c.y);
            v_w[4 * h + 2] = __int_as_float(c.z); v_w[4 * h + 3] = __int_as_float(c.w);
            v_cooc[4 * h] = d.x; v_cooc[4 * h + 1] = d.y; v_cooc[4 * h + 2] = d.z; v_cooc[4 * h + 3] = d.w;
        }
        unsigned keep = 0;
        int cnt = 0;
        // only the table's last tile can be ragged
        const int n_valid = (int)min((int64_t)ITEMS, p.n - base);
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const bool k = j < n_valid && keep_edge(v_w[j], p);
            v_w[j] = (float)((double)v_w[j] * p.factor);
            keep |= (unsigned)k << j;
            cnt += k;
        }
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[par][warp] = incl;
        named_bar(1, THREADS);                         // every thread has read its items out of the stage by now
        // every warp scans the warp totals for itself
        const int ws = lane < NW ? s_warp[par][lane] : 0;
        int wi = ws;
#pragma unroll
        for (int o = 1; o < NW; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        const int total = __shfl_sync(0xffffffffu, wi, NW - 1);
        if (!AHEAD && warp == 0 && lane == 0) {        // aliased call: the aggregate goes out only now, with the columns on chip
            TileRec* e = &queue[it % kQueue];
            e->total = total;
            st_state(p.state + tile, kFlagAgg | (u64)total);
            __threadfence_block();
            st_vol(&q_counted, it + 1);
        }
        int off = __shfl_sync(0xffffffffu, wi - ws, warp) + incl - cnt;
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            if (keep >> j & 1) {
                buf[off] = v_src[j]; buf[TILE + off] = v_dst[j];
                buf[2 * TILE + off] = __float_as_int(v_w[j]); buf[3 * TILE + off] = v_cooc[j];
                ++off;
            }
        }
        named_bar(1, THREADS);                         // compaction complete
        wait_above(&q_offset, it);
        const int64_t out0 = queue[it % kQueue].prefix;
        for (int i = tid; i < total; i += THREADS) {
            p.o_src[out0 + i] = buf[i];
            p.o_dst[out0 + i] = buf[TILE + i];
            p.o_w[out0 + i] = __int_as_float(buf[2 * TILE + i]);
            p.o_cooc[out0 + i] = buf[3 * TILE + i];
        }
        __syncwarp();
        if (lane == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the next writer of this stage is a bulk copy
            sm100::mbar_arrive(&empty[s]);
            if (warp == 0) st_vol(&q_scattered, it + 1);
        }
    }
}

// First pass of the two-pass form: survivors per tile (reads the weights only, 4 B per edge).
__global__ void __launch_bounds__(kThreads) count_kernel(const float* __restrict__ w, int64_t n, double factor, double thr,
                                                         int32_t* __restrict__ counts, u64* __restrict__ group_sum) {
    __shared__ int s_warp[kThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t base = (int64_t)blockIdx.x * kTile + (int64_t)tid * kItems;
    int cnt = 0;
    if (base + kItems <= n) {
#pragma unroll
        for (int h = 0; h < kItems / 4; ++h) {
            const float4 v = lz_ldg_stream(w + base + 4 * h);
            cnt += !((double)v.x * factor < thr);
            cnt += !((double)v.y * factor < thr);
            cnt += !((double)v.z * factor < thr);
            cnt += !((double)v.w * factor < thr);
        }
    } else {
        for (int j = 0; j < kItems; ++j)
            if (base + j < n) cnt += !((double)w[base + j] * factor < thr);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) s_warp[warp] = cnt;
    __syncthreads();
    if (tid == 0) {
        int t = 0;
#pragma unroll
        for (int i = 0; i < kThreads / 32; ++i) t += s_warp[i];
        counts[blockIdx.x] = t;
        if (group_sum != nullptr) atomicAdd((unsigned long long*)(group_sum + blockIdx.x / kGroup), (unsigned long long)t);
    }
}

// Exclusive scan of the tile counts by one CTA (tens of thousands of tiles at most): per-thread serial
// sums, warp shuffle scans, one shared-memory hop for the 32 warp totals.
__global__ void __launch_bounds__(1024) tile_scan_kernel(const int32_t* __restrict__ counts, int n_tiles,
                                                         int64_t* __restrict__ offsets, int64_t* __restrict__ total) {
    __shared__ long long s_warp[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (n_tiles + 1023) / 1024;
    const int lo = tid * per, hi = min(n_tiles, lo + per);
    long long sum = 0;
    for (int i = lo; i < hi; ++i) sum += counts[i];
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
        s_warp[lane] = wi - w;                           // exclusive warp offsets
        if (lane == 31) *total = wi;
    }
    __syncthreads();
    long long run = s_warp[warp] + incl - sum;
    for (int i = lo; i < hi; ++i) { offsets[i] = run; run += counts[i]; }
}

__global__ void __launch_bounds__(kThreads) decay_kernel(float* __restrict__ w, int64_t n, double factor) {
    const int64_t stride = (int64_t)gridDim.x * kThreads * 4;
    for (int64_t i = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 4 <= n) {
            float4 v = *reinterpret_cast<float4*>(w + i);
            v.x = (float)((double)v.x * factor); v.y = (float)((double)v.y * factor);
            v.z = (float)((double)v.z * factor); v.w = (float)((double)v.w * factor);
            *reinterpret_cast<float4*>(w + i) = v;
        } else {
            for (int64_t j = i; j < n; ++j) w[j] = (float)((double)w[j] * factor);
        }
    }
}

__device__ __forceinline__ float salience_step(float s32, double keep, double floor) {
    const double s = (double)s32;
    return (float)(s > floor ? floor + (s - floor) * keep : floor);
}

__global__ void __launch_bounds__(kThreads) salience_kernel(float* __restrict__ s, int64_t n, double keep, double floor) {
    const int64_t stride = (int64_t)gridDim.x * kThreads * 4;
    for (int64_t i = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 4 <= n) {
            float4 v = *reinterpret_cast<float4*>(s + i);
            v.x = salience_step(v.x, keep, floor); v.y = salience_step(v.y, keep, floor);
            v.z = salience_step(v.z, keep, floor); v.w = salience_step(v.w, keep, floor);
            *reinterpret_cast<float4*>(s + i) = v;
        } else {
            for (int64_t j = i; j < n; ++j) s[j] = salience_step(s[j], keep, floor);
        }
    }
}

int elementwise_blocks(int64_t n) {
    const int sms = lz_sm_count();
    int64_t blocks = (n + kThreads * 4 - 1) / (kThreads * 4);
    const int64_t cap = (int64_t)(sms > 0 ? sms : 148) * 8;
    return (int)(blocks < cap ? (blocks < 1 ? 1 : blocks) : cap);
}

// Streamed form: ring depth, run-ahead and CTAs per SM fixed at first use (LZ_PRUNE_STAGES / LZ_PRUNE_AHEAD / LZ_PRUNE_CTAS
// override them for measurements).  Scratch: ticket word, one aggregate word per tile, one prefix word per wave; this
// call zeroes them.
struct StreamCfg { int stages = 3, ahead = 6, ctas_per_sm = 0; bool ready = false; };
StreamCfg g_stream_cfg;
std::mutex g_stream_mu;

size_t stream_state_words(int64_t tiles) { return (size_t)(1 + tiles + tiles / kWave + 2); }

// The smallest float32 w with not (float64(w) * factor < thr), for factor > 0 (the product is monotone in w then).
bool prune_bound(double factor, double thr, float* out) {
    if (!(factor > 0.0) || !std::isfinite(factor) || !std::isfinite(thr)) return false;
    auto pruned = [&](float w) { return (double)w * factor < thr; };
    float c = (float)(thr / factor);
    if (!std::isfinite(c)) return false;
    for (int i = 0; i < 8 && pruned(c); ++i) c = nextafterf(c, INFINITY);
    for (int i = 0; i < 8; ++i) {
        const float d = nextafterf(c, -INFINITY);
        if (pruned(d)) break;
        c = d;
    }
    if (pruned(c) || !pruned(nextafterf(c, -INFINITY))) return false;
    *out = c;
    return true;
}

// Do any output columns overlap any input columns?
bool columns_alias(const PruneParams& p) {
    const uintptr_t in[4] = {(uintptr_t)p.src, (uintptr_t)p.dst, (uintptr_t)p.w, (uintptr_t)p.cooc};
    const uintptr_t out[4] = {(uintptr_t)p.o_src, (uintptr_t)p.o_dst, (uintptr_t)p.o_w, (uintptr_t)p.o_cooc};
    bool aliased = false;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            aliased = aliased || (out[j] < in[i] + (uintptr_t)p.n * 4 && in[i] < out[j] + (uintptr_t)p.n * 4);
    return aliased;
}

int launch_stream(PruneParams p, void* scratch, bool aliased, cudaStream_t st) {
    StreamCfg c;
    constexpr int kItemsS = kTile / kStreamThreads;
    {
        std::lock_guard<std::mutex> lock(g_stream_mu);
        if (!g_stream_cfg.ready) {
            if (const char* e = getenv("LZ_PRUNE_STAGES")) g_stream_cfg.stages = max(2, min(kMaxStages, atoi(e)));
            if (const char* e = getenv("LZ_PRUNE_AHEAD")) g_stream_cfg.ahead = max(2, min(kQueue - 2, atoi(e)));
            const size_t smem = (size_t)g_stream_cfg.stages * 16 * kTile + 8 * kTile;
            LZ_CUDA(cudaFuncSetAttribute(prune_stream_kernel<kStreamThreads, kItemsS, true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            LZ_CUDA(cudaFuncSetAttribute(prune_stream_kernel<kStreamThreads, kItemsS, false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            LZ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_stream_cfg.ctas_per_sm,
                                                                  prune_stream_kernel<kStreamThreads, kItemsS, true>,
                                                                  kStreamThreads + 96, smem));
            LZ_REQUIRE(g_stream_cfg.ctas_per_sm >= 1, "lz_edge_decay_prune: streamed kernel does not fit (%d stages)",
                       g_stream_cfg.stages);
            if (const char* e = getenv("LZ_PRUNE_CTAS")) g_stream_cfg.ctas_per_sm = max(1, min(g_stream_cfg.ctas_per_sm, atoi(e)));
            if (getenv("LZ_PRUNE_VERBOSE"))
                fprintf(stderr, "[lz] streamed prune: %d stages, ahead %d, %d CTAs per SM\n", g_stream_cfg.stages, g_stream_cfg.ahead,
                        g_stream_cfg.ctas_per_sm);
            g_stream_cfg.ready = true;
        }
        c = g_stream_cfg;
    }
    const int64_t tiles = (p.n + kTile - 1) / kTile;
    static const bool old_in_place = [] { const char* e = getenv("LZ_PRUNE_INPLACE_OLD"); return e && atoi(e) != 0; }();
    const bool ahead_in_place = aliased && !old_in_place;
    LZ_CUDA(cudaMemsetAsync(scratch, 0, stream_state_words(tiles) * 8 + (ahead_in_place ? (size_t)tiles * 4 : 0), st));
    p.ticket = (u32*)scratch;
    p.state = (u64*)scratch + 1;
    p.loaded = (int32_t*)((u64*)scratch + stream_state_words(tiles));
    p.check_loaded = ahead_in_place ? 1 : 0;
    p.n_tiles = (int)tiles;
    p.use_wstar = prune_bound(p.factor, p.thr, &p.wstar) ? 1 : 0;
    const int sms = lz_sm_count() > 0 ? lz_sm_count() : 148;
    const unsigned grid = (unsigned)min(tiles, (int64_t)sms * c.ctas_per_sm);
    const size_t smem = (size_t)c.stages * 16 * kTile + 8 * kTile;
    // in place the counts run ahead too; what a tile must wait for before it writes is that the (at most two) tiles whose
    // input ranges its survivors land in are on chip (LZ_PRUNE_INPLACE_OLD=1: the earlier form, counts published by the
    // compute warps once the columns are on chip)
    if (aliased && old_in_place)
        prune_stream_kernel<kStreamThreads, kItemsS, false><<<grid, kStreamThreads + 96, smem, st>>>(p, c.stages, c.stages);
    else
        prune_stream_kernel<kStreamThreads, kItemsS, true><<<grid, kStreamThreads + 96, smem, st>>>(p, c.stages, c.ahead);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

}  // namespace

extern "C" size_t lz_edge_decay_prune_scratch_bytes(int64_t n) {
    if (n <= 0) return 256;
    const int64_t tiles = (n + kTile - 1) / kTile;
    // single-pass forms: ticket + one word per tile (+ one per wave of 256 tiles in the streamed form); two-pass form:
    // int64 offsets + int32 counts
    return (size_t)(((tiles + tiles / 256 + 4) * 8 + tiles * 4 + 255) & ~255ll);
}

extern "C" int lz_edge_decay_prune(const int32_t* src, const int32_t* dst, const float* weight, const int32_t* cooc,
                                   int64_t n, double factor, double threshold, int32_t* out_src, int32_t* out_dst,
                                   float* out_weight, int32_t* out_cooc, int64_t* out_kept, void* scratch,
                                   size_t scratch_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    LZ_REQUIRE(out_kept != nullptr, "lz_edge_decay_prune: out_kept is NULL");
    LZ_REQUIRE(n >= 0, "lz_edge_decay_prune: n=%lld", (long long)n);
    if (n == 0) {
        LZ_CUDA(cudaMemsetAsync(out_kept, 0, sizeof(int64_t), st));
        return LZ_OK;
    }
    LZ_REQUIRE(src && dst && weight && cooc && out_src && out_dst && out_weight && out_cooc && scratch,
               "lz_edge_decay_prune: NULL pointer");
    LZ_REQUIRE((((uintptr_t)src | (uintptr_t)dst | (uintptr_t)weight | (uintptr_t)cooc) & 15) == 0,
               "lz_edge_decay_prune: edge arrays must be 16-byte aligned");
    LZ_REQUIRE(n < (1ll << 40), "lz_edge_decay_prune: n too large");
    const size_t need = lz_edge_decay_prune_scratch_bytes(n);
    if (scratch_bytes < need) {
        lz_set_error("lz_edge_decay_prune: scratch has %zu bytes, needs %zu", scratch_bytes, need);
        return LZ_ERR_SCRATCH;
    }
    const int64_t tiles = (n + kTile - 1) / kTile;
    PruneParams p;
    p.src = src; p.dst = dst; p.w = weight; p.cooc = cooc;
    p.o_src = out_src; p.o_dst = out_dst; p.o_w = out_weight; p.o_cooc = out_cooc;
    p.n = n; p.factor = factor; p.thr = threshold;
    p.ticket = (u32*)scratch;
    p.state = (u64*)scratch + 1;
    p.out_kept = out_kept; p.n_tiles = (int)tiles; p.offsets = nullptr; p.counts = nullptr; p.group_sum = nullptr; p.wstar = 0.f; p.use_wstar = 0;
    // Do any outputs overlap any inputs?  In place, a tile may only be overwritten after it has been read: the
    // streamed kernel then publishes a tile's aggregate only once its columns are in shared memory; out of place the
    // aggregates are counted ahead of the data path.
    const bool aliased = columns_alias(p);
    static const int form = [] { const char* e = getenv("LZ_PRUNE_FORM"); return e ? atoi(e) : 0; }();
    // default: the streamed single-pass kernel.  For comparisons: 2 chained look-back, 3 two-pass with the scan kernel,
    // 5 two-pass with group sums, 4 streamed data path alone (offsets from a first call; scripts/run_edges_form4.py)
    if (form == 4) {                                      // measurement only: streamed data path with offsets from a first call
        static bool primed = false;
        const size_t words = stream_state_words(tiles);
        int64_t* offsets = (int64_t*)scratch + words;
        int32_t* counts = (int32_t*)(offsets + tiles + 1);
        if (!primed) {
            count_kernel<<<(unsigned)tiles, kThreads, 0, st>>>(weight, n, factor, threshold, counts, nullptr);
            tile_scan_kernel<<<1, 1024, 0, st>>>(counts, (int)tiles, offsets, out_kept);
            primed = true;
        }
        p.offsets = offsets;
        if (int rc = launch_stream(p, scratch, aliased, st)) return rc;
    } else if (form <= 1) {
        if (int rc = launch_stream(p, scratch, aliased, st)) return rc;
    } else if (aliased || form == 2) {
        LZ_CUDA(cudaMemsetAsync(scratch, 0, (size_t)(tiles + 1) * 8, st));
        decay_prune_kernel<true><<<(unsigned)tiles, kThreads, 0, st>>>(p);
    } else {
        int32_t* counts = (int32_t*)((unsigned char*)scratch + (size_t)(tiles + 1) * 8);
        p.counts = counts;
        if (form == 3 || tiles > (1ll << 20)) {          // with the scan kernel (very large tables: too many groups to sum per CTA)
            int64_t* offsets = (int64_t*)scratch;
            count_kernel<<<(unsigned)tiles, kThreads, 0, st>>>(weight, n, factor, threshold, counts, nullptr);
            tile_scan_kernel<<<1, 1024, 0, st>>>(counts, (int)tiles, offsets, out_kept);
            p.offsets = offsets;
        } else {
            u64* group_sum = (u64*)scratch;
            LZ_CUDA(cudaMemsetAsync(group_sum, 0, (size_t)(tiles / kGroup + 1) * 8, st));
            count_kernel<<<(unsigned)tiles, kThreads, 0, st>>>(weight, n, factor, threshold, counts, group_sum);
            p.group_sum = group_sum;
        }
        decay_prune_kernel<false><<<(unsigned)tiles, kThreads, 0, st>>>(p);
    }
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

extern "C" int lz_edge_decay(float* weight, int64_t n, double factor, void* stream) {
    if (n <= 0) return LZ_OK;
    LZ_REQUIRE(weight && ((uintptr_t)weight & 15) == 0, "lz_edge_decay: weight must be non-NULL and 16-byte aligned");
    decay_kernel<<<elementwise_blocks(n), kThreads, 0, (cudaStream_t)stream>>>(weight, n, factor);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

extern "C" int lz_salience_decay(float* salience, int64_t n, double rate, double floor, void* stream) {
    if (n <= 0) return LZ_OK;
    LZ_REQUIRE(salience && ((uintptr_t)salience & 15) == 0,
               "lz_salience_decay: salience must be non-NULL and 16-byte aligned");
    salience_kernel<<<elementwise_blocks(n), kThreads, 0, (cudaStream_t)stream>>>(salience, n, 1.0 - rate, floor);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

// ---- allocation-free variant of the decay + prune sweep --------------------------------------------
//
// lz_edge_decay_prune needs a small device buffer (one ticket word + one look-back word per tile of
// 2048 edges) that its caller must allocate.  Some callers of this library cannot do that at the
// point where the sweep runs:
//   * a service that sized the row store to fill HBM and pre-allocated everything at start-up — a
//     late cudaMalloc can fail there, and cudaMalloc/cudaFree also synchronise the device, which
//     would stall the searches running on other streams while a conversation is being closed;
//   * a binding written against the bare C-ABI (C, Go/cgo, a JNI shim) that has device pointers for
//     its edge columns but no device allocator of its own to ask for scratch.
// For them the buffer below lives in the library image itself: a file-scope __device__ array, one
// copy per GPU that loads the library, held until the process exits.  Its size fixes the largest
// table the entry point accepts (kNoAllocMaxEdges); larger tables must use lz_edge_decay_prune.
namespace {
constexpr int64_t kNoAllocMaxEdges = LZ_NOALLOC_MAX_EDGES;                 // 2^26 edges
constexpr int64_t kNoAllocMaxTiles = kNoAllocMaxEdges / kTile;             // 32768 tiles
}  // namespace

static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNoAllocMaxTiles + kNoAllocMaxTiles / kWave + 4 + kNoAllocMaxTiles / 2];   // ~385 KiB

namespace {
std::mutex g_noalloc_mu;
void* g_noalloc_addr[64] = {nullptr};      // device address of g_noalloc_scratch, per device ordinal

// In host code `g_noalloc_scratch` is only a handle for the symbol, not a device address, so it must
// never be handed to a kernel as is.  The real address is looked up once per device (with that
// device current) and cached.
int noalloc_scratch(void** out) {
    int dev = 0;
    LZ_CUDA(cudaGetDevice(&dev));
    LZ_REQUIRE(dev >= 0 && dev < 64, "lz_edge_decay_prune_noalloc: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_addr[dev] == nullptr) {
        void* addr = nullptr;
        LZ_CUDA(cudaGetSymbolAddress(&addr, g_noalloc_scratch));
        LZ_REQUIRE(addr != nullptr, "lz_edge_decay_prune_noalloc: cudaGetSymbolAddress returned NULL");
        g_noalloc_addr[dev] = addr;
    }
    *out = g_noalloc_addr[dev];
    return LZ_OK;
}
}  // namespace

extern "C" int lz_edge_decay_prune_noalloc(const int32_t* src, const int32_t* dst, const float* weight,
                                           const int32_t* cooc, int64_t n, double factor, double threshold,
                                           int32_t* out_src, int32_t* out_dst, float* out_weight,
                                           int32_t* out_cooc, int64_t* out_kept, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    LZ_REQUIRE(out_kept != nullptr, "lz_edge_decay_prune_noalloc: out_kept is NULL");
    LZ_REQUIRE(n >= 0, "lz_edge_decay_prune_noalloc: n=%lld", (long long)n);
    LZ_REQUIRE(n <= kNoAllocMaxEdges,
               "lz_edge_decay_prune_noalloc: n=%lld exceeds the built-in scratch limit of %lld edges; "
               "use lz_edge_decay_prune with caller-provided scratch",
               (long long)n, (long long)kNoAllocMaxEdges);
    if (n == 0) {
        LZ_CUDA(cudaMemsetAsync(out_kept, 0, sizeof(int64_t), st));
        return LZ_OK;
    }
    LZ_REQUIRE(src && dst && weight && cooc && out_src && out_dst && out_weight && out_cooc,
               "lz_edge_decay_prune_noalloc: NULL pointer");
    LZ_REQUIRE((((uintptr_t)src | (uintptr_t)dst | (uintptr_t)weight | (uintptr_t)cooc) & 15) == 0,
               "lz_edge_decay_prune_noalloc: edge arrays must be 16-byte aligned");
    void* scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch)) return rc;
    // the array still holds the previous call's ticket and aggregate words: launch_stream starts them from zero
    PruneParams p;
    p.src = src; p.dst = dst; p.w = weight; p.cooc = cooc;
    p.o_src = out_src; p.o_dst = out_dst; p.o_w = out_weight; p.o_cooc = out_cooc;
    p.n = n; p.factor = factor; p.thr = threshold;
    p.out_kept = out_kept; p.offsets = nullptr; p.counts = nullptr; p.group_sum = nullptr; p.wstar = 0.f; p.use_wstar = 0;
    return launch_stream(p, scratch, columns_alias(p), st);
}

// ---- connected components over the edge table --------------------------------------
// Stands behind BufferGraph.get_connected_components (core/buffer_graph.py:99-120), which walks the graph
// with a recursive DFS that rescans every edge of a shard per visited node.  Here: union-find over the edge
// struct-of-arrays — hook every edge with a compare-and-swap that always attaches the larger root under the
// smaller one, then flatten — so labels[v] is the smallest node index of v's component whatever order the
// edges were processed in (deterministic).  Edges are treated as undirected.
namespace {
__device__ __forceinline__ int cc_find(int* parent, int v) {
    int p = parent[v];
    while (p != v) {                    // path halving
        const int gp = parent[p];
        if (gp != p) parent[v] = gp;
        v = p;
        p = gp;
    }
    return v;
}

__global__ void __launch_bounds__(kThreads) cc_init_kernel(int* parent, int n) {
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) parent[i] = i;
}

__global__ void __launch_bounds__(kThreads) cc_hook_kernel(const int32_t* __restrict__ src, const int32_t* __restrict__ dst,
                                                           int64_t n_edges, int n_nodes, int* parent) {
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < n_edges; e += (int64_t)gridDim.x * kThreads) {
        int u = src[e], v = dst[e];
        if (u < 0 || v < 0 || u >= n_nodes || v >= n_nodes || u == v) continue;
        while (true) {
            u = cc_find(parent, u);
            v = cc_find(parent, v);
            if (u == v) break;
            const int hi = u > v ? u : v, lo = u > v ? v : u;
            if (atomicCAS(&parent[hi], hi, lo) == hi) break;      // hi was still a root: attached
        }
    }
}

__global__ void __launch_bounds__(kThreads) cc_flatten_kernel(int* parent, int n) {
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) {
        int r = i;
        while (parent[r] != r) r = parent[r];
        parent[i] = r;
    }
}
}  // namespace

extern "C" int lz_connected_components(const int32_t* src, const int32_t* dst, int64_t n_edges, int32_t n_nodes,
                                       int32_t* labels, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n_nodes <= 0) return LZ_OK;
    LZ_REQUIRE(labels != nullptr && (n_edges == 0 || (src && dst)), "lz_connected_components: NULL pointer");
    LZ_REQUIRE(n_edges >= 0, "lz_connected_components: n_edges=%lld", (long long)n_edges);
    const int sms = lz_sm_count();
    if (sms <= 0) return LZ_ERR_DEVICE;
    auto blocks = [&](int64_t n) { int64_t b = (n + kThreads - 1) / kThreads; return (unsigned)(b < 1 ? 1 : (b > sms * 16 ? sms * 16 : b)); };
    cc_init_kernel<<<blocks(n_nodes), kThreads, 0, st>>>(labels, n_nodes);
    if (n_edges > 0) cc_hook_kernel<<<blocks(n_edges), kThreads, 0, st>>>(src, dst, n_edges, n_nodes, labels);
    cc_flatten_kernel<<<blocks(n_nodes), kThreads, 0, st>>>(labels, n_nodes);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

// ---- components the way the reference's traversal forms them ----------------------------------
// BufferGraph.get_connected_components (core/buffer_graph.py:99-120) is NOT plain undirected connectivity: it visits
// the nodes in dict order, and from a node it only follows the edges of the shard that node lives in
// (buffer_graph.py:72-77 -> memory_shard.py:54-62).  An edge whose other endpoint lives elsewhere (a cross-shard
// link, an id that is not resident) can therefore be walked from one end only, and the component such a target
// lands in depends on the visiting order.  What the DFS computes is:  label(v) = the SMALLEST index among the
// nodes that can reach v  (indices = visiting order).  Proof sketch: an unvisited node reachable from the current
// root is reachable through unvisited nodes only (a visited node on the path would have pulled it in earlier),
// so a root takes exactly the unvisited part of its reachable set, and the least ancestor of v is always a root.
//
// Device form: union-find over the two-way edges (every node of such a component reaches every other), then
// min-label propagation along the one-way edges between those components until nothing changes.
namespace {
__global__ void __launch_bounds__(kThreads) edge_global_kernel(const int32_t* __restrict__ src, const int32_t* __restrict__ dst,
                                                               const float* __restrict__ w, int64_t n, const int32_t* __restrict__ slot_global,
                                                               const uint8_t* __restrict__ slot_resident, int n_slots, double min_weight,
                                                               int32_t* __restrict__ o_src, int32_t* __restrict__ o_dst,
                                                               uint8_t* __restrict__ o_walk) {
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < n; e += (int64_t)gridDim.x * kThreads) {
        const int a = src[e], b = dst[e];
        const bool ok = a >= 0 && b >= 0 && a < n_slots && b < n_slots && (w == nullptr || (double)w[e] >= min_weight);
        o_src[e] = ok ? slot_global[a] : -1;
        o_dst[e] = ok ? slot_global[b] : -1;
        o_walk[e] = ok ? (uint8_t)((slot_resident[a] ? 1 : 0) | (slot_resident[b] ? 2 : 0)) : (uint8_t)0;
    }
}

__global__ void __launch_bounds__(kThreads) reach_hook_kernel(const int32_t* __restrict__ src, const int32_t* __restrict__ dst,
                                                              const uint8_t* __restrict__ walk, int64_t n_edges, int n_nodes,
                                                              int* parent, int2* __restrict__ list, int* list_n) {
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < n_edges; e += (int64_t)gridDim.x * kThreads) {
        int u = src[e], v = dst[e];
        const int wk = walk ? walk[e] : 3;
        if (wk == 0 || u < 0 || v < 0 || u >= n_nodes || v >= n_nodes || u == v) continue;
        if (wk != 3) {                                             // one-way: remember it as (from, to)
            list[atomicAdd(list_n, 1)] = wk == 1 ? make_int2(u, v) : make_int2(v, u);
            continue;
        }
        while (true) {
            u = cc_find(parent, u);
            v = cc_find(parent, v);
            if (u == v) break;
            const int hi = u > v ? u : v, lo = u > v ? v : u;
            if (atomicCAS(&parent[hi], hi, lo) == hi) break;
        }
    }
}

__device__ __forceinline__ int ld_cg(const int* p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// One CTA: L[root] = min(L[root], L[root of a one-way predecessor]) until a whole sweep changes nothing.  One-way
// edges are the cross-shard links (a handful per new memory), so a single CTA sweeping the list is ample.
__global__ void __launch_bounds__(1024) reach_propagate_kernel(const int2* __restrict__ list, const int* __restrict__ list_n,
                                                               const int* __restrict__ root, int* L) {
    const int n = *list_n;
    while (true) {
        int changed = 0;
        for (int i = threadIdx.x; i < n; i += 1024) {
            const int2 e = list[i];
            const int lu = ld_cg(L + root[e.x]), rv = root[e.y];
            if (lu < ld_cg(L + rv)) { atomicMin(L + rv, lu); changed = 1; }
        }
        if (!__syncthreads_or(changed)) break;
    }
}

__global__ void __launch_bounds__(kThreads) reach_apply_kernel(int* labels, const int* __restrict__ L, int n) {
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) labels[i] = L[labels[i]];
}
}  // namespace

extern "C" int lz_edge_endpoints_to_global(const int32_t* src, const int32_t* dst, const float* weight, int64_t n,
                                           const int32_t* slot_global, const uint8_t* slot_resident, int32_t n_slots,
                                           double min_weight, int32_t* out_src, int32_t* out_dst, uint8_t* out_walk,
                                           void* stream) {
    if (n <= 0) return LZ_OK;
    LZ_REQUIRE(src && dst && slot_global && slot_resident && out_src && out_dst && out_walk,
               "lz_edge_endpoints_to_global: NULL pointer");
    const int sms = lz_sm_count();
    if (sms <= 0) return LZ_ERR_DEVICE;
    int64_t b = (n + kThreads - 1) / kThreads;
    if (b > (int64_t)sms * 16) b = (int64_t)sms * 16;
    edge_global_kernel<<<(unsigned)b, kThreads, 0, (cudaStream_t)stream>>>(src, dst, weight, n, slot_global, slot_resident,
                                                                           n_slots, min_weight, out_src, out_dst, out_walk);
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}

extern "C" size_t lz_reach_components_scratch_bytes(int64_t n_edges, int32_t n_nodes) {
    const size_t nodes = ((size_t)(n_nodes > 0 ? n_nodes : 0) * 4 + 255) & ~(size_t)255;
    return nodes + 256 + (size_t)(n_edges > 0 ? n_edges : 0) * 8;
}

extern "C" int lz_reach_components(const int32_t* src, const int32_t* dst, const uint8_t* walk, int64_t n_edges,
                                   int32_t n_nodes, void* scratch, size_t scratch_bytes, int32_t* labels, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n_nodes <= 0) return LZ_OK;
    LZ_REQUIRE(labels != nullptr && scratch != nullptr && (n_edges == 0 || (src && dst)), "lz_reach_components: NULL pointer");
    LZ_REQUIRE(n_edges >= 0 && n_edges < (1ll << 31), "lz_reach_components: n_edges=%lld", (long long)n_edges);
    const size_t need = lz_reach_components_scratch_bytes(n_edges, n_nodes);
    if (scratch_bytes < need) {
        lz_set_error("lz_reach_components: scratch has %zu bytes, needs %zu", scratch_bytes, need);
        return LZ_ERR_SCRATCH;
    }
    const int sms = lz_sm_count();
    if (sms <= 0) return LZ_ERR_DEVICE;
    auto blocks = [&](int64_t n) { int64_t b = (n + kThreads - 1) / kThreads; return (unsigned)(b < 1 ? 1 : (b > sms * 16 ? sms * 16 : b)); };
    const size_t nodes_bytes = ((size_t)n_nodes * 4 + 255) & ~(size_t)255;
    int* L = (int*)scratch;
    int* list_n = (int*)((unsigned char*)scratch + nodes_bytes);
    int2* list = (int2*)((unsigned char*)scratch + nodes_bytes + 256);
    LZ_CUDA(cudaMemsetAsync(list_n, 0, 4, st));
    cc_init_kernel<<<blocks(n_nodes), kThreads, 0, st>>>(labels, n_nodes);
    cc_init_kernel<<<blocks(n_nodes), kThreads, 0, st>>>(L, n_nodes);
    if (n_edges > 0) reach_hook_kernel<<<blocks(n_edges), kThreads, 0, st>>>(src, dst, walk, n_edges, n_nodes, labels, list, list_n);
    cc_flatten_kernel<<<blocks(n_nodes), kThreads, 0, st>>>(labels, n_nodes);
    if (n_edges > 0 && walk != nullptr) {
        reach_propagate_kernel<<<1, 1024, 0, st>>>(list, list_n, labels, L);
        reach_apply_kernel<<<blocks(n_nodes), kThreads, 0, st>>>(labels, L, n_nodes);
    }
    LZ_CUDA(cudaGetLastError());
    return LZ_OK;
}
