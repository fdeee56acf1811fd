// CTA-level streaming top-k selector kept in shared memory.
//
// Entries are (key, id): key = lz_ord(score) with "larger = better", id = row index; the order is
// (key desc, id asc), a total order because ids are unique.  Threads push() candidates that beat the
// current threshold into an unsorted buffer; every so often the whole CTA compact()s the buffer with
// a bitonic sort, keeps the best k and raises the threshold to the k-th entry.  The caller bounds the
// number of pushes between two maybe_compact() calls by `chunk`, and capacity >= k + chunk (rounded
// up to a power of two), so a push can never overflow.
#pragma once
#include "lz_common.cuh"

struct LzSelector {
    u64* key;       // [cap] shared memory
    u32* id;        // [cap] shared memory
    int  cap;       // power of two
    int  k;
    int  count;     // number of live entries (buffer prefix)
    u32  thr_id;
    u64  thr_key;   // an entry must beat (thr_key, thr_id) to be pushed
};

__device__ __forceinline__ void lz_sel_init(LzSelector* s, u64* key, u32* id, int cap, int k) {
    s->key = key; s->id = id; s->cap = cap; s->k = k; s->count = 0;
    s->thr_key = 0ull; s->thr_id = 0xffffffffu;          // everything real beats this
}

// Any thread. `key` must come from a non-NaN score.
__device__ __forceinline__ void lz_sel_push(LzSelector* s, u64 key, u32 id) {
    if (lz_better(key, id, s->thr_key, s->thr_id)) {
        int pos = atomicAdd(&s->count, 1);
        s->key[pos] = key;
        s->id[pos] = id;
    }
}

// All 32 lanes of a warp, convergently; `want` says whether this lane has an entry.  One shared-memory atomic
// per warp instead of one per entry (hundreds of lanes hammering the one counter serialise).
__device__ __forceinline__ void lz_sel_push_warp(LzSelector* s, u64 key, u32 id, bool want) {
    const bool ok = want && lz_better(key, id, s->thr_key, s->thr_id);
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if (m == 0u) return;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(&s->count, __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (ok) {
        const int pos = base + __popc(m & ((1u << lane) - 1u));
        s->key[pos] = key;
        s->id[pos] = id;
    }
}

// All threads of the CTA (T threads). Sorts descending, keeps the best k, tightens the threshold.
template <int T>
__device__ void lz_sel_compact(LzSelector* s) {
    const int tid = threadIdx.x;
    const int n = s->count;
    const int P = lz_next_pow2(n < 2 ? 2 : n);
    u64* key = s->key;
    u32* id = s->id;
    for (int i = n + tid; i < P; i += T) { key[i] = 0ull; id[i] = 0xffffffffu; }
    __syncthreads();
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < (P >> 1); t += T) {
                int lo = 2 * t - (t & (stride - 1));
                int hi = lo + stride;
                bool desc = (lo & size) == 0;
                u64 ka = key[lo], kb = key[hi];
                u32 ia = id[lo], ib = id[hi];
                if (lz_better(ka, ia, kb, ib) != desc) {
                    key[lo] = kb; key[hi] = ka; id[lo] = ib; id[hi] = ia;
                }
            }
            __syncthreads();
        }
    }
    if (tid == 0) {
        if (n >= s->k) {
            s->count = s->k;
            s->thr_key = key[s->k - 1];
            s->thr_id = id[s->k - 1];
        }
    }
    __syncthreads();
}

// All threads. Compacts when the next `chunk` pushes might not fit. Condition is CTA-uniform
// (count is read after a barrier).
template <int T>
__device__ __forceinline__ void lz_sel_maybe_compact(LzSelector* s, int chunk) {
    __syncthreads();
    if (s->count + chunk > s->cap) lz_sel_compact<T>(s);
}
