"""Oracle: exhaustive (flat) k-nearest-neighbour search.  TEST INFRASTRUCTURE — see oracle/__init__.py.

Restates what `LanceDBStore.search_nodes` asks LanceDB for
(/root/reference/src/lazzaro/core/vector_store.py:132-140): an exhaustive scan of the float32
`vector` column (schema vector_store.py:37) restricted to rows of one user (the `.where()` at
vector_store.py:137), the best `limit` rows, ids only (vector_store.py:140).  No ANN index is ever
created (vector_store.py:55 is a scalar index), so the search is exact.

The distance arithmetic itself lives in the third-party `lancedb` package (unpinned, absent), so
this restatement is "parity unpinned" at that boundary.  Conventions fixed here:

* rows are float32; the query is cast to float32 first (the column type); products and sums are
  float64 (each row's sum is formed the same way whatever its position, so bit-identical rows get
  bit-identical scores);
* metric "cosine": similarity = x.q / (|x| |q|), 0.0 when the norm product is 0 — the formula of
  `MemorySystem._cosine_similarity` (memory_system.py:197-203); best = largest;
* metric "l2": squared Euclidean distance sum((x-q)^2) (LanceDB's default metric); best = smallest;
* ties are broken by ascending row index (= insertion order, which is also how the reference's
  stable sorts break ties: memory_system.py:470,827,865);
* the filter (user / tombstone mask) is applied BEFORE the top-k.
"""
from __future__ import annotations

import numpy as np

_CHUNK = 1 << 15


def row_norms(rows: np.ndarray) -> np.ndarray:
    """float64 Euclidean norm of each float32 row (memory_system.py:202: np.linalg.norm)."""
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    out = np.empty(rows.shape[0], dtype=np.float64)
    for s in range(0, rows.shape[0], _CHUNK):
        x = rows[s:s + _CHUNK].astype(np.float64)
        out[s:s + _CHUNK] = np.sqrt((x * x).sum(axis=1))
    return out


def flat_scores(rows: np.ndarray, query: np.ndarray, metric: str = "cosine") -> np.ndarray:
    """float64 score of every row against one query (cosine similarity or squared L2 distance)."""
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    q = np.asarray(query, dtype=np.float32).astype(np.float64)
    n = rows.shape[0]
    out = np.empty(n, dtype=np.float64)
    if metric == "cosine":
        qn = float(np.sqrt((q * q).sum()))
        for s in range(0, n, _CHUNK):
            x = rows[s:s + _CHUNK].astype(np.float64)
            dot = (x * q).sum(axis=1)
            den = np.sqrt((x * x).sum(axis=1)) * qn
            with np.errstate(divide="ignore", invalid="ignore"):
                sim = dot / den
            sim[~(den > 0)] = 0.0                      # memory_system.py:203
            out[s:s + _CHUNK] = sim
    elif metric == "l2":
        for s in range(0, n, _CHUNK):
            d = rows[s:s + _CHUNK].astype(np.float64) - q
            out[s:s + _CHUNK] = (d * d).sum(axis=1)
    else:
        raise ValueError(f"unknown metric {metric!r}")
    return out


def select_topk(scores: np.ndarray, k: int, metric: str = "cosine", mask: np.ndarray | None = None):
    """Indices and scores of the best k rows, ordered best first, ties by ascending index.

    `mask` (bool, True = row is eligible) is applied before selection.
    """
    idx = np.arange(scores.shape[0], dtype=np.int64)
    if mask is not None:
        idx = idx[mask]
    s = scores[idx]
    ok = ~np.isnan(s)
    idx, s = idx[ok], s[ok]
    if k <= 0 or idx.size == 0:
        return np.empty(0, np.int64), np.empty(0, np.float64)
    key = -s if metric == "cosine" else s              # ascending key = best first
    k = min(k, idx.size)
    if k < idx.size:
        kth = np.partition(key, k - 1)[k - 1]
        keep = key <= kth
        idx, s, key = idx[keep], s[keep], key[keep]
    order = np.lexsort((idx, key))[:k]
    return idx[order], s[order]


def flat_topk(rows, query, k, metric="cosine", mask=None):
    """Exhaustive top-k of one query: (row indices int64, scores float64)."""
    return select_topk(flat_scores(rows, query, metric), k, metric, mask)


def flat_topk_batch(rows, queries, k, metric="cosine", mask=None):
    """Exhaustive top-k of each query; returns (ids [B,k] int64 padded with -1, scores [B,k])."""
    queries = np.atleast_2d(np.asarray(queries, dtype=np.float32))
    ids = np.full((queries.shape[0], k), -1, dtype=np.int64)
    sc = np.full((queries.shape[0], k), np.nan, dtype=np.float64)
    for b, q in enumerate(queries):
        i, s = flat_topk(rows, q, k, metric, mask)
        ids[b, :i.size] = i
        sc[b, :s.size] = s
    return ids, sc


def merge_topk(id_lists, score_lists, k, metric="cosine"):
    """Merge per-partition top-k lists (global ids) into a global top-k — the associative merge the
    multi-GPU path relies on (order: best score first, ties by ascending id)."""
    ids = np.concatenate([np.asarray(i, dtype=np.int64) for i in id_lists])
    sc = np.concatenate([np.asarray(s, dtype=np.float64) for s in score_lists])
    ok = ids >= 0
    ids, sc = ids[ok], sc[ok]
    key = -sc if metric == "cosine" else sc
    order = np.lexsort((ids, key))[:k]
    return ids[order], sc[order]


def fast_topk_f32(rows: np.ndarray, queries: np.ndarray, k: int, metric: str = "cosine",
                  inv_norms: np.ndarray | None = None, chunk: int = 1 << 18):
    """Throughput-oriented CPU search used ONLY as the timed CPU baseline (bench.py): float32 BLAS
    GEMM (all host threads) + argpartition per chunk, the fastest numpy form of the same flat scan.
    Not used for parity (float32 accumulation)."""
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    q = np.ascontiguousarray(np.atleast_2d(queries), dtype=np.float32)
    B, n = q.shape[0], rows.shape[0]
    if metric == "cosine":
        qn = np.linalg.norm(q, axis=1, keepdims=True)
        q = q / np.where(qn > 0, qn, 1)
        if inv_norms is None:
            nr = np.linalg.norm(rows, axis=1)
            inv_norms = (1.0 / np.where(nr > 0, nr, np.inf)).astype(np.float32)
    best_s = np.full((B, 0), 0, np.float32)
    best_i = np.full((B, 0), 0, np.int64)
    for s0 in range(0, n, chunk):
        x = rows[s0:s0 + chunk]
        sc = q @ x.T                                                  # [B, chunk] sgemm
        if metric == "cosine":
            sc *= inv_norms[None, s0:s0 + chunk]
        else:
            sc = 2.0 * sc - (x * x).sum(axis=1)[None, :]              # -|x-q|^2 + |q|^2
        kk = min(k, sc.shape[1])
        part = np.argpartition(-sc, kk - 1, axis=1)[:, :kk]
        best_s = np.concatenate([best_s, np.take_along_axis(sc, part, 1)], axis=1)
        best_i = np.concatenate([best_i, part + s0], axis=1)
        if best_s.shape[1] > 4 * k:
            o = np.argpartition(-best_s, k - 1, axis=1)[:, :k]
            best_s, best_i = np.take_along_axis(best_s, o, 1), np.take_along_axis(best_i, o, 1)
    o = np.argsort(-best_s, axis=1, kind="stable")[:, :k]
    return np.take_along_axis(best_i, o, 1), np.take_along_axis(best_s, o, 1)


def two_level_topk(rows, shard_offsets, queries, k, nprobe, metric="cosine"):
    """Two-level search, restated exactly (north_star (1); level 1 = core/memory_system.py:464-470):
    centroid of a shard = float64 mean of its float32 member rows (memory_system.py:916-917), kept as float32;
    level 1 ranks the shards by cosine(query, centroid), best first, ties by lower shard index, and keeps `nprobe`;
    level 2 is the exhaustive top-k (`metric`, ties by row index) over the rows of exactly those shards.
    Returns (ids [B,k] int64 padded with -1, scores [B,k] float64, probed shards [B,nprobe] int64)."""
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    off = np.asarray(shard_offsets, dtype=np.int64)
    S = off.size - 1
    cen = np.stack([rows[off[s]:off[s + 1]].astype(np.float64).mean(axis=0) if off[s + 1] > off[s] else np.zeros(rows.shape[1])
                    for s in range(S)]).astype(np.float32)
    queries = np.atleast_2d(np.asarray(queries, dtype=np.float32))
    p = min(nprobe, S)
    ids = np.full((queries.shape[0], k), -1, dtype=np.int64)
    sc = np.full((queries.shape[0], k), np.nan, dtype=np.float64)
    probes = np.full((queries.shape[0], p), -1, dtype=np.int64)
    for b, q in enumerate(queries):
        ps, _ = flat_topk(cen, q, p, "cosine")
        probes[b, :ps.size] = ps
        mask = np.zeros(rows.shape[0], dtype=bool)
        for s in ps:
            mask[off[s]:off[s + 1]] = True
        i, v = flat_topk(rows, q, k, metric, mask)
        ids[b, :i.size] = i
        sc[b, :v.size] = v
    return ids, sc, probes
