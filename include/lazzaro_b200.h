/*
 * lazzaro_b200 — C-ABI of the B200-native memory hot path (liblazzaro_b200.so).
 *
 * The reference (thelaycon/lazzaro) is pure Python; the only native code on its hot path is the
 * third-party `lancedb` engine reached through pyo3.  This header is what a binding for that path
 * binds instead: plain C symbols, raw DEVICE pointers and sizes, an explicit CUDA stream, an `int`
 * status (0 = ok, < 0 = error; text from lz_last_error()).  No exceptions cross the boundary and no
 * call synchronises the device unless its comment says so.  Unless stated otherwise every pointer
 * is a device pointer owned by the caller, `stream` is a `cudaStream_t` passed as `void*`, and work
 * is enqueued on that stream.
 *
 * Citations are to /root/reference/src/lazzaro/ (the interface each entry point stands behind).
 * INTEGRATION.md shows the ctypes binding and how the Python `Store` mirror calls these.
 *
 * Conventions
 *   rows      float32 [n, ld] row-major, `dim` used floats per row, dim % 4 == 0, ld % 4 == 0, base 16-B aligned
 *   norms     float64 [n]  Euclidean norm of each row (lz_row_norms)
 *   tags      int32   [n]  owner tag of each row (the store's user index); tag < 0 = deleted row
 *   metric    LZ_COSINE: score = x.q/(|x||q|), 0 when |x||q| == 0, best = largest
 *             LZ_L2    : score = sum (x-q)^2, best = smallest
 *   ordering  best score first, ties by ascending row index; NaN scores are never returned
 */
#ifndef LAZZARO_B200_H
#define LAZZARO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LZ_OK            0
#define LZ_ERR_ARG      -1   /* bad argument (size, alignment, unsupported k/dim) */
#define LZ_ERR_CUDA     -2   /* a CUDA runtime call failed */
#define LZ_ERR_SCRATCH  -3   /* scratch buffer too small */
#define LZ_ERR_DEVICE   -4   /* not an sm_100 device */

#define LZ_COSINE 0
#define LZ_L2     1

#define LZ_MAX_K          1024  /* largest top-k a search call accepts */
#define LZ_MAX_GROUP      4     /* most queries that share one pass over the rows in lz_search_exact */

/* Library version (major*10000 + minor*100 + patch). Host only. */
int lz_version(void);

/* Text of the calling thread's last error ("" if none). Host only; the pointer stays valid until
 * the thread's next failing call. */
const char* lz_last_error(void);

/* Checks that the current device is sm_100 and returns its SM count in *sm_count. Host only.
 * The kernels are built for sm_100a only: every compute entry point fails with LZ_ERR_DEVICE /
 * LZ_ERR_CUDA on anything else — there is no CPU or generic-GPU fallback. */
int lz_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- row ingest (stands behind LanceDBStore.add_nodes, core/vector_store.py:89-112) ------------ */

/* norms[i] = sqrt(sum_j double(rows[i][j])^2), float64.  Precomputed once per added row so that a
 * search reads each row exactly once (memory_system.py:202 recomputes both norms per pair). */
int lz_row_norms(const float* rows, int64_t n, int dim, int64_t ld, double* norms, void* stream);

/* ---- exact flat search (stands behind LanceDBStore.search_nodes, core/vector_store.py:132-140,
 *      and MemorySystem._cosine_similarity, core/memory_system.py:197-203) ----------------------- */

/* Bytes of device scratch lz_search_exact needs for (n_queries, k, n_chunks). */
size_t lz_search_exact_scratch_bytes(int n_queries, int k, int n_chunks);

/* Suggested n_chunks (CTAs per query group) for a scan of `n_rows` rows by `n_groups` groups on
 * the current device: a multiple of the SM count when the rows allow it. Host only. */
int lz_search_exact_default_chunks(int64_t n_rows, int n_groups);

/*
 * Exhaustive top-k in float64 over float32 rows: one streaming pass over the rows per query group.
 *
 * Queries are taken in groups of `group_size` (1..LZ_MAX_GROUP) consecutive queries; a group scans
 * rows [lo, hi) once for all its queries.  If group_lo/group_hi are NULL every group scans
 * [row_lo, row_hi); otherwise group g scans [group_lo[g], group_hi[g]) (per-user or per-shard
 * ranges).  If tags != NULL a row is eligible only when tags[row] >= 0 and, when group_tag != NULL
 * and group_tag[g] >= 0, tags[row] == group_tag[g]  (the `.where(user_id = ...)` pre-filter of
 * vector_store.py:137, applied BEFORE the top-k).
 *
 *   queries    float32 [n_queries, q_ld]
 *   out_ids    int64   [n_queries, k]   row indices, -1 padded
 *   out_scores float64 [n_queries, k]   cosine similarity or squared L2 distance, NaN padded
 *   out_counts int32   [n_queries]      number of valid entries per query (<= k)
 *   scratch    >= lz_search_exact_scratch_bytes(n_queries, k, n_chunks) bytes; contents are
 *              overwritten; concurrent calls must not share it
 *   norms      may be NULL for LZ_L2
 *   id_base    added to every returned row index (a partition's first global row id, so that
 *              per-GPU results can be merged by lz_merge_topk without another pass)
 */
int lz_search_exact(const float* rows, int64_t ld, int dim, const double* norms, const int32_t* tags,
                    int64_t row_lo, int64_t row_hi,
                    const int64_t* group_lo, const int64_t* group_hi, const int32_t* group_tag,
                    const float* queries, int64_t q_ld, int n_queries, int group_size,
                    int k, int metric, int n_chunks, int64_t id_base,
                    void* scratch, size_t scratch_bytes,
                    int64_t* out_ids, double* out_scores, int32_t* out_counts, void* stream);

/* ---- batched search on the tensor cores (same reference call sites, for query batches; also the
 *      dedup / linking scans of core/memory_system.py:719-742, 838-891) ------------------------------ */

#define LZ_BATCHED_MAX_K 128   /* largest k lz_search_batched accepts (larger k: lz_search_exact) */

/* The fp16 copy of the rows ("shadow") that the coarse pass reads; refresh the rows you add.  dim % 8 == 0.
 * (fp16 rather than bf16: same tensor-core rate, and its 11 significant bits keep the coarse score's rounding
 * error 8x smaller — small enough for the deterministic certification margin below to be practical.)
 *   LZ_COSINE: shadow[i][j] = fp16(rows[i][j] / norms[i]) (zeros for a zero row), fp16 [n, dim] dense
 *   LZ_L2    : shadow[i] = [fp16(rows[i]), h, m, l, 0, 0, 0, 0], fp16 [n, dim + 8] dense, h + m + l = a three-way
 *              fp16 split of -|row|^2 / 2, so that one contraction yields x.q - |x|^2/2 (L2 ranking)
 *   row_err   optional float32 [n] (NULL to skip): an upper bound of the 2-norm of each row's rounding error,
 *             |shadow - row/|row|| (LZ_COSINE) or |fp16(row) - row| (LZ_L2); +inf for a row whose components or
 *             norm term leave fp16's range.  The largest value over the rows a search scans is what
 *             lz_search_batched's deterministic margin takes as det_row_err. */
int lz_rows_to_shadow(const float* rows, int64_t n, int dim, int64_t ld, const double* norms, int metric,
                      void* shadow, float* row_err, void* stream);

/* Bytes of device scratch lz_search_batched needs for a scan of n_rows = row_hi - row_lo rows. */
size_t lz_search_batched_scratch_bytes(int64_t n_rows, int n_queries, int k, int dim);

/*
 * Exact top-k (LZ_COSINE or LZ_L2) for a batch of queries, computed as: fp16 tcgen05 GEMM (queries x rows, never
 * materialised) with a fused per-query candidate selection -> float64 rescoring of the candidates from
 * the float32 rows -> top-k with the id tie-break -> certification (k-th exact score must exceed the
 * best dropped coarse score by the fp16 error margin) -> exact float64 scan of the queries that fail
 * it.  Results are those of lz_search_exact with the same metric whenever the margin holds.  Two margins:
 *   det_row_err < 0   STATISTICAL: margin_sigmas standard deviations of the fp16 rounding error of the coarse
 *                     score, assuming the operands' rounding errors are independent (true of real-valued
 *                     embeddings; NOT of quantised / fp16-derived / near-constant vectors, whose errors line
 *                     up: such a row can be dropped and the query still reported certified).
 *   det_row_err >= 0  DETERMINISTIC: |coarse - exact| <= det_row_err*|q^| + dq*|x^| + det_row_err*dq + the
 *                     fp32 accumulation bound (Cauchy-Schwarz; dq is measured per query), valid for any data.
 *                     det_row_err = max of lz_rows_to_shadow's row_err over [row_lo, row_hi).  Some 10x wider
 *                     than the statistical margin on Gaussian-like rows (5e-4 against 4e-5 at dim 768).
 *   shadow        fp16 [n_rows_total, dim (+ 8 for LZ_L2)] from lz_rows_to_shadow with the same metric.  Must be plain
 *                 device memory allocated on the CURRENT GPU (checked; LZ_ERR_ARG otherwise): the coarse pass loads
 *                 it by TMA multicast, and TMA multicast loads here read only this GPU's memory, since on B200 a
 *                 TMA multicast load from another GPU's memory hung the GPU.  A multi-GPU search keeps one shadow
 *                 per GPU and merges the per-GPU results (lz_merge_topk) — it never hands one GPU another's rows
 *   max_row_norm  LZ_L2 only: an upper bound of the row norms (scales the error margin)
 *   tags/want_tag as lz_search_exact with one tag for the whole batch (want_tag < 0: any live row)
 *   row_lo,row_hi the row range every query scans
 *   out_certified optional int32 [n_queries]: 1 = answered by the tensor-core pass, 0 = by the fallback
 */
int lz_search_batched(const void* shadow, const float* rows, int64_t ld, int dim, const double* norms,
                      const int32_t* tags, int want_tag, int64_t n_rows_total, int64_t row_lo, int64_t row_hi,
                      const float* queries, int64_t q_ld, int n_queries, int k, double margin_sigmas,
                      int metric, double max_row_norm, double det_row_err, int64_t id_base,
                      void* scratch, size_t scratch_bytes,
                      int64_t* out_ids, double* out_scores, int32_t* out_counts, int32_t* out_certified,
                      void* stream);

/* ONE query, ONE launch, no separate merge and no device->host copy: the call for a small store at the reference's own
 * scale (10k rows), where launches and copies cost more than the scan.  Same answer as lz_search_exact for one query.
 *   query      float32 [dim], 16-byte aligned, DEVICE-ACCESSIBLE — typically pinned host memory (cudaHostAlloc /
 *              torch pin_memory: under unified addressing the host pointer is valid on the device), read in place
 *   out_block  device-accessible, typically pinned host memory, k*16 + 8 bytes: [k int64 ids | k float64 scores |
 *              int32 count | int32 flag]; the flag is set to `seq` by the kernel AFTER the answer is in place, so the
 *              host may poll it instead of synchronising the stream (pass a different seq on every call)
 *   scratch    device, >= lz_search_one_scratch_bytes(k); its first 256 bytes must be ZERO before the first call
 *              (they hold the kernel's completion ticket, which it resets itself); one call at a time per scratch
 *   tags / want_tag  as lz_search_exact with one group (want_tag < 0: any live row) */
size_t lz_search_one_scratch_bytes(int k);
int lz_search_one(const float* rows, int64_t ld, int dim, const double* norms, const int32_t* tags, int want_tag,
                  int64_t row_lo, int64_t row_hi, const float* query, int k, int metric, int64_t id_base,
                  void* scratch, size_t scratch_bytes, void* out_block, int seq, void* stream);

/* Exhaustive float64 top-k of query q over the UNION of n_ranges row ranges [range_lo[q*n_ranges + j],
 * range_hi[q*n_ranges + j]) (device int64 arrays; an empty range contributes nothing; ranges of one query must not
 * overlap) — the member scan of a two-level search whose probed shards differ per query, one pass, no merge of
 * per-range results by the caller.  chunks_per_range CTAs share each range.  Scratch:
 * lz_search_exact_scratch_bytes(n_queries, k, n_ranges * chunks_per_range).  Other arguments as lz_search_exact. */
int lz_search_exact_ranges(const float* rows, int64_t ld, int dim, const double* norms, const int32_t* tags,
                           const int64_t* range_lo, const int64_t* range_hi, int n_ranges, int64_t n_rows_total,
                           const float* queries, int64_t q_ld, int n_queries, int k, int metric,
                           int chunks_per_range, int64_t id_base, void* scratch, size_t scratch_bytes,
                           int64_t* out_ids, double* out_scores, int32_t* out_counts, void* stream);

/* ---- two-level semantic-shard search on the tensor cores (north_star (1); the reference's level 1 is the cosine of
 *      the query against every super-node centroid, core/memory_system.py:464-470) ---------------------------------
 *
 * Rows are stored shard by shard: shard s = rows [shard_offsets[s], shard_offsets[s+1]).  Level 1 scores the query batch
 * against the shard centroids (lz_search_batched over the centroid table, k = nprobe: cosine, exact, ties to the lower
 * shard index).  Level 2 scores each query against the member rows of its nprobe shards only: the (query, shard) pairs
 * are grouped by shard on the device, the queries of a shard are copied next to each other into a permuted query
 * array, and the same tcgen05/TMA GEMM + fused candidate epilogue runs over a work list of (128-query tile, shard
 * slice) items; float64 rescoring, top-k with the id tie-break and certification as in lz_search_batched; queries
 * that fail certification are answered by the exact scan of exactly their probed shards (lz_search_exact_ranges).
 * The result is the exact top-k over the union of the probed shards (oracle/flat_search.py two_level_topk);
 * nprobe >= n_shards visits everything and equals lz_search_batched bit for bit.  nprobe < n_shards is an
 * APPROXIMATION of the flat search — report it with recall, never as parity.
 *   cen_shadow / centroids / cen_norms   centroid table [n_shards, dim]: lz_rows_to_shadow(LZ_COSINE) copy, float32 rows
 *                 (lz_segment_mean, cast to float32), lz_row_norms; cen_det_row_err: max row_err of that shadow, or < 0
 *                 for the statistical margin at level 1
 *   shard_offsets device int64 [n_shards + 1];  k <= LZ_BATCHED_MAX_K;  nprobe <= LZ_MAX_K (clamped to n_shards)
 *   out_probe     optional int64 [n_queries, nprobe]: the shards each query visited, best first
 * margin_sigmas, metric, max_row_norm, det_row_err, id_base, outputs: as lz_search_batched (for the member pass). */
size_t lz_search_two_level_scratch_bytes(int n_shards, int n_queries, int k, int nprobe, int dim);
int lz_search_two_level(const void* shadow, const float* rows, int64_t ld, int dim, const double* norms,
                        int64_t n_rows_total, const int64_t* shard_offsets, int n_shards,
                        const void* cen_shadow, const float* centroids, int64_t cen_ld, const double* cen_norms,
                        double cen_det_row_err, const float* queries, int64_t q_ld, int n_queries, int k, int nprobe,
                        double margin_sigmas, int metric, double max_row_norm, double det_row_err, int64_t id_base,
                        void* scratch, size_t scratch_bytes, int64_t* out_ids, double* out_scores, int32_t* out_counts,
                        int32_t* out_certified, int64_t* out_probe, void* stream);

/* ---- shard (re)assignment: nearest centroid of every row (north_star (2); the reference has no k-means — its only
 *      centroid arithmetic is the super-node mean of core/memory_system.py:916-917 — so the rule is fixed here:
 *      argmax over the table of cosine(row, table row), ties to the LOWEST table index) --------------------------------
 *
 * The rows are the query side of the contraction, the (tiny) centroid table the scanned side: the same GEMM in its
 * top-1 mode (candidates = single table rows within 3 eps of the running best), then every row is scored in float64
 * against its surviving candidates from a per-CTA shared-memory cache of the table rows its block needs.
 *   table_shadow / table / table_norms   centroid table [n_table, dim] (n_table <= 2048 for the cached kernel; larger tables
 *                 take the generic top-1 path): lz_rows_to_shadow(LZ_COSINE), float32 rows, lz_row_norms
 *   table_det_row_err   max row_err of table_shadow (deterministic margin) or < 0 (statistical, margin_sigmas)
 *   row_shadow, row_err optional: the rows' OWN fp16 cosine shadow and rounding-error norms (a resident store already
 *                 holds both) — the conversion pass over the rows is then skipped (needs table_det_row_err >= 0)
 *   out_idx int64 [n_rows], out_cos float64 [n_rows], out_counts int32 [n_rows] (0 only when the table is empty or all
 *   scores are NaN), out_certified optional int32 [n_rows].  Scratch: lz_assign_nearest_scratch_bytes(...). */
size_t lz_assign_nearest_scratch_bytes(int n_table, int64_t n_rows, int dim, int have_row_shadow);
int lz_assign_nearest(const void* table_shadow, const float* table, int64_t table_ld, int dim, const double* table_norms,
                      int n_table, double table_det_row_err, const float* rows, int64_t ld, int64_t n_rows,
                      const void* row_shadow, const float* row_err, double margin_sigmas, void* scratch,
                      size_t scratch_bytes, int64_t* out_idx, double* out_cos, int32_t* out_counts,
                      int32_t* out_certified, void* stream);

/* Merge `n_lists` per-partition top-k lists per query into one (the all-gather merge of the
 * multi-GPU search; ids are global row ids).  in_ids int64 [n_lists, n_queries, k] (-1 = empty
 * slot), in_scores float64 same shape; outputs as lz_search_exact.  Uses no scratch. */
int lz_merge_topk(const int64_t* in_ids, const double* in_scores, int n_lists, int n_queries, int k,
                  int metric, int64_t* out_ids, double* out_scores, int32_t* out_counts, void* stream);
/* Same, with `list_stride` elements (>= n_queries * k) between two consecutive lists in BOTH input arrays: lets
 * one all-gather of a packed [ids | scores] block per GPU feed the merge directly (ids of list c at
 * in_ids + c*list_stride, scores at in_scores + c*list_stride).  No reference counterpart. */
int lz_merge_topk_strided(const int64_t* in_ids, const double* in_scores, int64_t list_stride, int n_lists,
                          int n_queries, int k, int metric, int64_t* out_ids, double* out_scores,
                          int32_t* out_counts, void* stream);

/* The exchange step of the multi-GPU search in one call: ncclAllGather of this rank's packed block, then the
 * strided merge, both on `stream`.  No reference counterpart (the reference is a single process).
 *   nccl_comm    an ncclComm_t (passed as void*) of `world` ranks, one per GPU, created by the NCCL library that is
 *                loaded in this process (resolved with dlopen at first use; no NCCL is needed unless this is called)
 *   local_block  int64 [2, n_queries, k]: this rank's ids (global row ids, -1 = empty) then its float64 scores' bits —
 *                what lz_search_exact / lz_search_batched write when out_ids / out_scores point into one block
 *   gathered     int64 [world, 2, n_queries, k] receive buffer
 * Every rank ends with the same global top-k.  Each rank's kernels read that rank's own memory only. */
int lz_allgather_merge(void* nccl_comm, int world, const int64_t* local_block, int64_t* gathered, int n_queries, int k,
                       int metric, int64_t* out_ids, double* out_scores, int32_t* out_counts, void* stream);

/* Exact float64 scores of explicit (query, row) pairs — the rescoring step after a coarse pass and
 * the pairwise cosine of memory_system.py:197-203.  pair_q int32 [n_pairs], pair_row int64
 * [n_pairs] (row < 0 gives NaN); out float64 [n_pairs]. */
int lz_score_pairs(const float* rows, int64_t ld, int dim, const double* norms,
                   const float* queries, int64_t q_ld, const int32_t* pair_q, const int64_t* pair_row,
                   int64_t n_pairs, int metric, double* out_scores, void* stream);

/* ---- centroids (stands behind MemorySystem._create_super_nodes_for_shard, core/memory_system.py:916-917,
 *      and the centroid update of shard reassignment) ----------------------------------------------- */

/* out_means[s][j] = mean over i in [seg_offsets[s], seg_offsets[s+1]) of rows[order ? order[i] : i][j],
 * accumulated in float64 in a fixed order (deterministic); an empty segment gives zeros.
 *   order        optional int64 [seg_offsets[n_segments]] row indices (members grouped by segment);
 *                NULL = segments are contiguous row ranges (rows stored shard by shard)
 *   seg_offsets  int64 [n_segments + 1];  out_means float64 [n_segments, dim] */
int lz_segment_mean(const float* rows, int64_t ld, int dim, const int64_t* order, const int64_t* seg_offsets,
                    int n_segments, double* out_means, void* stream);

/* Member lists of an assignment: a STABLE counting sort of the row indices 0..n-1 by labels[i] (int64, entries outside
 * [0, n_labels) are left out), so that lz_segment_mean(order = out_order, seg_offsets = out_offsets) sums every centroid's
 * members in ascending row order (deterministic).  n_labels <= 16384.  out_order int64 [n] (first out_offsets[n_labels]
 * entries valid), out_offsets int64 [n_labels + 1].  No reference counterpart (the reference has no k-means). */
size_t lz_group_by_label_scratch_bytes(int64_t n, int n_labels);
int lz_group_by_label(const int64_t* labels, int64_t n, int n_labels, void* scratch, size_t scratch_bytes,
                      int64_t* out_order, int64_t* out_offsets, void* stream);

/* ---- biological decay and pruning (stands behind MemoryShard.apply_temporal_decay and
 *      MemoryShard.prune_weak_edges, core/memory_shard.py:64-84; fan-out core/buffer_graph.py:87-97) */

/* Bytes of device scratch lz_edge_decay_prune needs for n edges. */
size_t lz_edge_decay_prune_scratch_bytes(int64_t n);

/*
 * One fused sweep over the edge table (struct of arrays): w' = float32(double(w) * factor)
 * (memory_shard.py:69-70 with factor = (1-decay_rate)^sweeps), drop edges with double(w)*factor <
 * threshold (strict, memory_shard.py:81), and compact the survivors stably (their order is kept,
 * as dict order is in the reference).  Outputs may alias the inputs (in-place compaction).
 * One pass over the table: every column is read from HBM once (one kernel launch + one memset).
 *   src,dst,cooc int32 [n]; weight float32 [n]; out_* same types, capacity >= n
 *   out_kept     int64 [1] device: number of surviving edges (pruned = n - kept)
 *   scratch      >= lz_edge_decay_prune_scratch_bytes(n); cleared by the call itself
 */
int lz_edge_decay_prune(const int32_t* src, const int32_t* dst, const float* weight, const int32_t* cooc,
                        int64_t n, double factor, double threshold,
                        int32_t* out_src, int32_t* out_dst, float* out_weight, int32_t* out_cooc,
                        int64_t* out_kept, void* scratch, size_t scratch_bytes, void* stream);

/*
 * The same sweep for callers that cannot allocate device memory where it runs.  (No counterpart in
 * the reference: it exists because of how this library is deployed, not because of the algorithm.)
 * Such callers are, for example, a service that filled HBM with the row store at start-up, where a
 * late cudaMalloc may fail and would in any case synchronise the device under the searches in
 * flight, or a plain-C / cgo / JNI binding that owns the edge columns but has no device allocator
 * to ask for scratch.  The per-tile words live in an array inside the library image instead
 * (about 385 KiB of global memory on every GPU that loads the library, kept until the process exits),
 * so this call takes inputs and outputs only and never calls cudaMalloc, cudaFree or any allocating
 * routine.
 *   limit      n <= LZ_NOALLOC_MAX_EDGES (2^26 = 67,108,864 edges); a larger n is rejected with
 *              LZ_ERR_ARG and a message — use lz_edge_decay_prune with your own scratch for those.
 *   first call the first call on each device looks the array's address up (with that device
 *              current) and caches it, and the first sweep of a process sets the kernel's
 *              shared-memory attribute; make that first call outside of any stream capture.
 *   sharing    every call on a device uses the same array and finds the previous call's contents
 *              in it; the call zeroes the words it needs on `stream` before the kernel, so calls
 *              issued one after another on one stream are fine, but two calls that can overlap on
 *              the GPU (different streams, different host threads) race on it and corrupt each
 *              other — serialise them, or use lz_edge_decay_prune with separate scratch buffers.
 * Arguments, results and aliasing rules are otherwise those of lz_edge_decay_prune.
 */
#define LZ_NOALLOC_MAX_EDGES 67108864
int lz_edge_decay_prune_noalloc(const int32_t* src, const int32_t* dst, const float* weight, const int32_t* cooc,
                                int64_t n, double factor, double threshold,
                                int32_t* out_src, int32_t* out_dst, float* out_weight, int32_t* out_cooc,
                                int64_t* out_kept, void* stream);

/* Decay only: weight[i] = float32(double(weight[i]) * factor) in place (memory_shard.py:69-70). */
int lz_edge_decay(float* weight, int64_t n, double factor, void* stream);

/* Node salience decay in place (memory_shard.py:72-77):
 * s = s > floor ? floor + (s - floor) * (1 - rate) : floor, computed in float64, stored float32. */
int lz_salience_decay(float* salience, int64_t n, double rate, double floor, void* stream);

/* ---- connected components (stands behind BufferGraph.get_connected_components,
 *      core/buffer_graph.py:99-120; the reference's recursive DFS rescans a shard's edges per node) ------- */

/* labels[v] = smallest node index in v's connected component of the undirected graph with edges
 * (src[i], dst[i]), nodes 0..n_nodes-1 (edges with an endpoint outside that range are ignored).
 * Union-find with min-root hooking; deterministic.  src, dst int32 [n_edges]; labels int32 [n_nodes]. */
int lz_connected_components(const int32_t* src, const int32_t* dst, int64_t n_edges, int32_t n_nodes,
                            int32_t* labels, void* stream);

/* Components exactly as the reference's traversal forms them (core/buffer_graph.py:99-120): nodes are visited in
 * index order, and an edge is followed from an endpoint only if that endpoint may walk it — the reference looks a
 * node's neighbours up among the edges of the shard the node lives in (buffer_graph.py:72-77), so an edge whose
 * other endpoint lives in another shard (or is not resident) is one-way, and where such a target lands depends on
 * the visiting order.  Result: labels[v] = the smallest index among the nodes that can reach v (= the root the
 * reference's DFS assigns when indices are its dict order); components in ascending label order are the
 * reference's list.  walk uint8 [n_edges]: bit 0 = walkable src->dst, bit 1 = walkable dst->src, 0 = ignored
 * (NULL: every edge both ways, i.e. lz_connected_components).  Deterministic.
 *   scratch  >= lz_reach_components_scratch_bytes(n_edges, n_nodes); labels int32 [n_nodes] */
size_t lz_reach_components_scratch_bytes(int64_t n_edges, int32_t n_nodes);
int lz_reach_components(const int32_t* src, const int32_t* dst, const uint8_t* walk, int64_t n_edges, int32_t n_nodes,
                        void* scratch, size_t scratch_bytes, int32_t* labels, void* stream);

/* Prepares one shard's edge table for lz_reach_components: maps the shard-local endpoint slots to the caller's
 * global node indices and derives the walk flags from residency — out_walk[e] = (slot_resident[src[e]] ? 1 : 0) |
 * (slot_resident[dst[e]] ? 2 : 0), or 0 (edge ignored, endpoints -1) when (double)weight[e] < min_weight (the
 * reference passes min_weight = 0.0, buffer_graph.py:110) or a slot is out of range.  weight may be NULL.
 *   slot_global int32 [n_slots], slot_resident uint8 [n_slots]; outputs [n] */
int lz_edge_endpoints_to_global(const int32_t* src, const int32_t* dst, const float* weight, int64_t n,
                                const int32_t* slot_global, const uint8_t* slot_resident, int32_t n_slots,
                                double min_weight, int32_t* out_src, int32_t* out_dst, uint8_t* out_walk, void* stream);

/* ---- buffer-limit eviction (stands behind MemorySystem._enforce_buffer_limit,
 *      core/memory_system.py:535-578) ------------------------------------------------------------------ */

/* importance[i] = salience*0.5 + min(1, access_count/10)*0.3 + (1/(1 + (now - last_accessed)/86400))*0.2,
 * float64, rounded operation by operation as the reference's Python expression (memory_system.py:545-550);
 * the k LEAST important nodes with skip[i] == 0 come back in (importance ascending, index ascending) order —
 * the reference's stable ascending sort (memory_system.py:552-553) when the caller's index order is the
 * reference's iteration order (shards in dict order, nodes in dict order; super-nodes flagged in `skip`).
 * The reference reads the clock once per node; here `now` is one value for the whole sweep.
 * salience: float32 [n] (salience_is_f64 = 0, the stored column type, vector_store.py:42) or float64 [n];
 * access_count int32 [n]; last_accessed float64 [n] (epoch seconds); skip uint8 [n] or NULL.
 * k <= LZ_MAX_K (evict in rounds beyond that, flagging the evicted in `skip`).  NaN importances are never
 * evicted.  out_idx int64 [k], out_importance float64 [k], out_count int32 [1] (all device);
 * out_all_importance float64 [n] or NULL receives every node's importance. */
size_t lz_evict_lowest_scratch_bytes(int64_t n, int k);
int lz_evict_lowest(const void* salience, int salience_is_f64, const int32_t* access_count,
                    const double* last_accessed, const uint8_t* skip, int64_t n, double now, int k,
                    void* scratch, size_t scratch_bytes, int64_t* out_idx, double* out_importance,
                    int32_t* out_count, double* out_all_importance, void* stream);

/* ---- adjacency: CSR over the edge table + neighbour lookup (stands behind MemoryShard.get_neighbors,
 *      core/memory_shard.py:54-62, fan-out core/buffer_graph.py:72-77) ----------------------------------- */

/* Index the undirected adjacency of edges (src[e], dst[e]), e < n_edges, over nodes 0..n_nodes-1.
 * offsets int64 [n_nodes + 1]; adj uint32 [2 * n_edges]: node v's entries are adj[offsets[v] .. offsets[v+1])
 * in ASCENDING EDGE ORDER (the reference's edge-dict order), entry = 2*e + (1 if v is the edge's dst end, 0 if
 * its src end); a self-loop appears once (src end); edges with an endpoint outside [0, n_nodes) are left out.
 * offsets[n_nodes] = number of entries.  n_edges < 2^30.  Deterministic (stable radix sort). */
size_t lz_csr_build_scratch_bytes(int64_t n_edges, int32_t n_nodes);
int lz_csr_build(const int32_t* src, const int32_t* dst, int64_t n_edges, int32_t n_nodes, void* scratch,
                 size_t scratch_bytes, int64_t* offsets, uint32_t* adj, void* stream);

/* For each queried node: the other endpoint of every incident edge with (double)weight[e] >= min_weight, in
 * edge order (memory_shard.py:57-61: `>=`, source end tested first).  nodes int32 [n_queries] (ids outside
 * [0, n_nodes) give 0 neighbours).  out_count[q] = number of qualifying edges; the first min(count, max_out)
 * are written to out_nbr[q*max_out ..] (and their edge indices to out_edge if not NULL) — a count above
 * max_out means the list was truncated and the call should be repeated with a larger max_out. */
int lz_neighbors(const int64_t* offsets, const uint32_t* adj, const int32_t* src, const int32_t* dst,
                 const float* weight, int32_t n_nodes, const int32_t* nodes, int n_queries, double min_weight,
                 int max_out, int32_t* out_nbr, int32_t* out_edge, int32_t* out_count, void* stream);

/* ---- edge upsert (stands behind MemoryShard.add_edge, core/memory_shard.py:42-52) ------------------------ */

/* out_edge[q] = index of the edge with src == q_src[q] and dst == q_dst[q] (the reference's dict key
 * (source, target), memory_shard.py:44), or -1; found on the CSR of lz_csr_build (`dst` is the edge table's
 * column).  q_src, q_dst, out_edge: int32 [n_queries]. */
int lz_edge_find(const int64_t* offsets, const uint32_t* adj, const int32_t* dst, int32_t n_nodes,
                 const int32_t* q_src, const int32_t* q_dst, int n_queries, int32_t* out_edge, void* stream);

/* Reinforcement of existing edges (memory_shard.py:45-49): for each i < m, `times[i]` times (1 if times is NULL)
 * weight[e] = min(1.0, weight[e] + 0.1) — float64 add and min, stored float32 — and co_occurrence[e] += times[i],
 * e = edge_idx[i].  Indices outside [0, n_edges) are skipped; an index must not repeat within one call. */
int lz_edge_reinforce(float* weight, int32_t* co_occurrence, int64_t n_edges, const int32_t* edge_idx,
                      const int32_t* times, int64_t m, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LAZZARO_B200_H */
