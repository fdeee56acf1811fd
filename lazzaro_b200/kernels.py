"""Thin typed wrappers over the C-ABI (include/lazzaro_b200.h): torch tensors in, torch tensors out.

torch is used here for device memory and streams only; every computation below is a call into
liblazzaro_b200.so.  All calls are asynchronous on torch's current stream of the tensors' device.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


def _dev(t: torch.Tensor):
    if not t.is_cuda:
        raise _lib.LazzaroB200Error("lazzaro_b200 kernels take CUDA tensors only (no CPU fallback)")
    return t.device


def row_norms(rows: torch.Tensor, dim: Optional[int] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """float64 Euclidean norm of each float32 row (lz_row_norms)."""
    dev = _dev(rows)
    assert rows.dtype == torch.float32 and rows.dim() == 2 and rows.stride(1) == 1
    n, dim = rows.shape[0], dim or rows.shape[1]
    if out is None:
        out = torch.empty(n, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        check(_lib.load().lz_row_norms(ptr(rows), n, dim, rows.stride(0), ptr(out), stream_ptr(dev)), "lz_row_norms")
    return out


def default_chunks(n_rows: int, n_groups: int, device) -> int:
    with torch.cuda.device(device):
        return int(_lib.load().lz_search_exact_default_chunks(int(n_rows), int(n_groups)))


def search_exact(rows: torch.Tensor, dim: int, norms: Optional[torch.Tensor], tags: Optional[torch.Tensor],
                 queries: torch.Tensor, k: int, metric: int = _lib.COSINE,
                 row_range: Optional[Tuple[int, int]] = None,
                 group_lo: Optional[torch.Tensor] = None, group_hi: Optional[torch.Tensor] = None,
                 group_tag: Optional[torch.Tensor] = None, group_size: Optional[int] = None,
                 n_chunks: Optional[int] = None, scratch: Optional[torch.Tensor] = None, id_base: int = 0,
                 out: Optional[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]] = None):
    """Exhaustive float64 top-k (lz_search_exact). Returns (ids int64 [B,k], scores float64 [B,k],
    counts int32 [B]) on the device, unsynchronised."""
    dev = _dev(rows)
    assert rows.dtype == torch.float32 and queries.dtype == torch.float32 and queries.dim() == 2
    assert queries.stride(1) == 1 and rows.stride(1) == 1 and queries.device == dev
    B = queries.shape[0]
    lo, hi = row_range if row_range is not None else (0, rows.shape[0])
    if group_lo is not None:
        group_size = 1 if group_size is None else group_size
    if group_size is None:
        group_size = 4 if B >= 4 else (2 if B >= 2 else 1)
    n_groups = (B + group_size - 1) // group_size
    lib = _lib.load()
    with torch.cuda.device(dev):
        if n_chunks is None:
            span = (hi - lo) if group_lo is None else max(1, (hi - lo))
            n_chunks = int(lib.lz_search_exact_default_chunks(int(span), n_groups))
        need = int(lib.lz_search_exact_scratch_bytes(B, k, n_chunks))
        if scratch is None or scratch.numel() < need:
            scratch = torch.empty(need, dtype=torch.uint8, device=dev)
        if out is None:
            out = (torch.empty((B, k), dtype=torch.int64, device=dev),
                   torch.empty((B, k), dtype=torch.float64, device=dev),
                   torch.empty((B,), dtype=torch.int32, device=dev))
        ids, scores, counts = out
        check(lib.lz_search_exact(ptr(rows), rows.stride(0), dim, ptr(norms), ptr(tags), lo, hi,
                                  ptr(group_lo), ptr(group_hi), ptr(group_tag),
                                  ptr(queries), queries.stride(0), B, group_size, k, metric, n_chunks, int(id_base),
                                  ptr(scratch), scratch.numel(), ptr(ids), ptr(scores), ptr(counts),
                                  stream_ptr(dev)), "lz_search_exact")
    return ids, scores, counts


def merge_topk(in_ids: torch.Tensor, in_scores: torch.Tensor, metric: int = _lib.COSINE):
    """Merge [n_lists, B, k] per-partition lists into [B, k] (lz_merge_topk)."""
    dev = _dev(in_ids)
    assert in_ids.dtype == torch.int64 and in_scores.dtype == torch.float64
    assert in_ids.is_contiguous() and in_scores.is_contiguous() and in_ids.shape == in_scores.shape
    L, B, k = in_ids.shape
    ids = torch.empty((B, k), dtype=torch.int64, device=dev)
    scores = torch.empty((B, k), dtype=torch.float64, device=dev)
    counts = torch.empty((B,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.load().lz_merge_topk(ptr(in_ids), ptr(in_scores), L, B, k, metric, ptr(ids), ptr(scores),
                                        ptr(counts), stream_ptr(dev)), "lz_merge_topk")
    return ids, scores, counts


def merge_topk_packed(packed: torch.Tensor, metric: int = _lib.COSINE):
    """Merge from ONE gathered buffer: packed int64 [n_lists, 2, B, k], [:, 0] = ids, [:, 1] = the float64
    scores' bits (lz_merge_topk_strided with list_stride = 2*B*k)."""
    dev = _dev(packed)
    assert packed.dtype == torch.int64 and packed.is_contiguous() and packed.dim() == 4 and packed.shape[1] == 2
    L, _, B, k = packed.shape
    ids = torch.empty((B, k), dtype=torch.int64, device=dev)
    scores = torch.empty((B, k), dtype=torch.float64, device=dev)
    counts = torch.empty((B,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.load().lz_merge_topk_strided(packed.data_ptr(), packed.data_ptr() + B * k * 8, 2 * B * k, L, B, k, metric,
                                                ptr(ids), ptr(scores), ptr(counts), stream_ptr(dev)), "lz_merge_topk_strided")
    return ids, scores, counts


def score_pairs(rows: torch.Tensor, dim: int, norms: Optional[torch.Tensor], queries: torch.Tensor,
                pair_q: torch.Tensor, pair_row: torch.Tensor, metric: int = _lib.COSINE) -> torch.Tensor:
    """Exact float64 score of each (query, row) pair (lz_score_pairs)."""
    dev = _dev(rows)
    assert pair_q.dtype == torch.int32 and pair_row.dtype == torch.int64
    n = pair_q.numel()
    out = torch.empty(n, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        check(_lib.load().lz_score_pairs(ptr(rows), rows.stride(0), dim, ptr(norms), ptr(queries), queries.stride(0),
                                         ptr(pair_q), ptr(pair_row), n, metric, ptr(out), stream_ptr(dev)),
              "lz_score_pairs")
    return out


def edge_decay_prune(src, dst, weight, cooc, n: int, factor: float, threshold: float,
                     out=None, scratch: Optional[torch.Tensor] = None, kept: Optional[torch.Tensor] = None):
    """Fused decay + strict-threshold prune + stable compaction (lz_edge_decay_prune).
    `out` = (src, dst, weight, cooc) destination tensors; default is in place.
    Returns (kept int64[1] on device, out tensors)."""
    dev = _dev(weight)
    lib = _lib.load()
    if out is None:
        out = (src, dst, weight, cooc)
    with torch.cuda.device(dev):
        need = int(lib.lz_edge_decay_prune_scratch_bytes(n))
        if scratch is None or scratch.numel() < need:
            scratch = torch.empty(need, dtype=torch.uint8, device=dev)
        if kept is None:
            kept = torch.empty(1, dtype=torch.int64, device=dev)
        check(lib.lz_edge_decay_prune(ptr(src), ptr(dst), ptr(weight), ptr(cooc), n, float(factor), float(threshold),
                                      ptr(out[0]), ptr(out[1]), ptr(out[2]), ptr(out[3]), ptr(kept),
                                      ptr(scratch), scratch.numel(), stream_ptr(dev)), "lz_edge_decay_prune")
    return kept, out


def edge_decay(weight: torch.Tensor, n: int, factor: float):
    dev = _dev(weight)
    with torch.cuda.device(dev):
        check(_lib.load().lz_edge_decay(ptr(weight), n, float(factor), stream_ptr(dev)), "lz_edge_decay")


def salience_decay(salience: torch.Tensor, n: int, rate: float, floor: float = 0.2):
    dev = _dev(salience)
    with torch.cuda.device(dev):
        check(_lib.load().lz_salience_decay(ptr(salience), n, float(rate), float(floor), stream_ptr(dev)),
              "lz_salience_decay")


def shadow_dim(dim: int, metric: int = _lib.COSINE) -> int:
    return dim + 8 if metric == _lib.L2 else dim


def rows_to_shadow(rows: torch.Tensor, dim: int, norms: torch.Tensor, out: Optional[torch.Tensor] = None,
                 metric: int = _lib.COSINE, row_err: Optional[torch.Tensor] = None) -> torch.Tensor:
    """bf16 shadow of float32 rows (lz_rows_to_shadow): normalised rows for cosine; raw rows plus a three-way
    split of -|row|^2/2 (8 extra columns) for L2.  `row_err` (float32 [n], optional) receives each row's
    rounding-error norm — its maximum is the `det_row_err` of the deterministic certification margin."""
    dev = _dev(rows)
    n, sd = rows.shape[0], shadow_dim(dim, metric)
    if out is None:
        out = torch.empty((n, sd), dtype=torch.float16, device=dev)
    assert out.dtype == torch.float16 and out.is_contiguous() and out.shape[1] == sd
    assert row_err is None or (row_err.dtype == torch.float32 and row_err.is_contiguous() and row_err.numel() >= n)
    with torch.cuda.device(dev):
        check(_lib.load().lz_rows_to_shadow(ptr(rows), n, dim, rows.stride(0), ptr(norms), int(metric), ptr(out),
                                          ptr(row_err), stream_ptr(dev)), "lz_rows_to_shadow")
    return out


def search_batched(shadow: torch.Tensor, rows: torch.Tensor, dim: int, norms: torch.Tensor,
                   tags: Optional[torch.Tensor], want_tag: int, queries: torch.Tensor, k: int,
                   row_range: Optional[Tuple[int, int]] = None, margin_sigmas: float = 4.0, id_base: int = 0,
                   scratch: Optional[torch.Tensor] = None, n_rows_total: Optional[int] = None,
                   metric: int = _lib.COSINE, max_row_norm: Optional[float] = None, out=None,
                   det_row_err: Optional[float] = None):
    """Batched exact cosine top-k on the tensor cores (lz_search_batched).
    Returns (ids int64 [B,k], scores float64 [B,k], counts int32 [B], certified int32 [B]).
    det_row_err = None: statistical margin (margin_sigmas); a float >= 0: the deterministic (worst-case)
    margin with that bound of the rows' rounding-error norms (max of rows_to_shadow's row_err)."""
    dev = _dev(rows)
    assert shadow.dtype == torch.float16 and shadow.is_contiguous() and shadow.shape[1] == shadow_dim(dim, metric)
    if metric == _lib.L2 and max_row_norm is None:
        max_row_norm = float(norms[:n_rows_total if n_rows_total is not None else norms.shape[0]].max().item())
    assert queries.dtype == torch.float32 and queries.stride(1) == 1 and rows.stride(1) == 1
    B = queries.shape[0]
    n_total = n_rows_total if n_rows_total is not None else min(rows.shape[0], shadow.shape[0])
    lo, hi = row_range if row_range is not None else (0, n_total)
    lib = _lib.load()
    with torch.cuda.device(dev):
        need = int(lib.lz_search_batched_scratch_bytes(hi - lo, B, k, dim))
        if scratch is None or scratch.numel() < need:
            scratch = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
        if out is None:
            out = (torch.empty((B, k), dtype=torch.int64, device=dev), torch.empty((B, k), dtype=torch.float64, device=dev),
                   torch.empty((B,), dtype=torch.int32, device=dev))
        ids, scores, counts = out
        cert = torch.empty((B,), dtype=torch.int32, device=dev)
        check(lib.lz_search_batched(ptr(shadow), ptr(rows), rows.stride(0), dim, ptr(norms), ptr(tags), int(want_tag),
                                    int(n_total), int(lo), int(hi), ptr(queries), queries.stride(0), B, k,
                                    float(margin_sigmas), int(metric), float(max_row_norm or 0.0),
                                    -1.0 if det_row_err is None else float(det_row_err), int(id_base),
                                    ptr(scratch), scratch.numel(),
                                    ptr(ids), ptr(scores), ptr(counts), ptr(cert), stream_ptr(dev)), "lz_search_batched")
    return ids, scores, counts, cert


def search_exact_ranges(rows: torch.Tensor, dim: int, norms: Optional[torch.Tensor], queries: torch.Tensor,
                        range_lo: torch.Tensor, range_hi: torch.Tensor, k: int, metric: int = _lib.COSINE,
                        n_rows_total: Optional[int] = None, chunks_per_range: Optional[int] = None, id_base: int = 0):
    """Exact float64 top-k of query q over the union of the row ranges [range_lo[q, j], range_hi[q, j]) (int64 [B, R] on the
    device) — lz_search_exact_ranges.  Returns (ids, scores, counts)."""
    dev = _dev(rows)
    assert range_lo.dtype == torch.int64 and range_hi.dtype == torch.int64 and range_lo.is_contiguous() and range_hi.is_contiguous()
    B, R = range_lo.shape
    assert queries.shape[0] == B and queries.dtype == torch.float32 and queries.stride(1) == 1
    lib = _lib.load()
    n_total = rows.shape[0] if n_rows_total is None else n_rows_total
    with torch.cuda.device(dev):
        if chunks_per_range is None:
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            chunks_per_range = max(1, (2 * sms) // max(1, B * R))
        need = int(lib.lz_search_exact_scratch_bytes(B, k, R * chunks_per_range))
        scratch = torch.empty(max(need, 16), dtype=torch.uint8, device=dev)
        ids = torch.empty((B, k), dtype=torch.int64, device=dev)
        scores = torch.empty((B, k), dtype=torch.float64, device=dev)
        counts = torch.empty((B,), dtype=torch.int32, device=dev)
        check(lib.lz_search_exact_ranges(ptr(rows), rows.stride(0), dim, ptr(norms), None, ptr(range_lo), ptr(range_hi), R,
                                         int(n_total), ptr(queries), queries.stride(0), B, k, int(metric), int(chunks_per_range),
                                         int(id_base), ptr(scratch), scratch.numel(), ptr(ids), ptr(scores), ptr(counts),
                                         stream_ptr(dev)), "lz_search_exact_ranges")
    return ids, scores, counts


def search_two_level(shadow: torch.Tensor, rows: torch.Tensor, dim: int, norms: torch.Tensor, n_rows_total: int,
                     shard_offsets: torch.Tensor, cen_shadow: torch.Tensor, centroids: torch.Tensor, cen_norms: torch.Tensor,
                     queries: torch.Tensor, k: int, nprobe: int, metric: int = _lib.COSINE, margin_sigmas: float = 4.0,
                     max_row_norm: float = 0.0, det_row_err: Optional[float] = None, cen_det_row_err: Optional[float] = None,
                     id_base: int = 0, scratch: Optional[torch.Tensor] = None, want_probe: bool = False):
    """Two-level search on the tensor cores (lz_search_two_level): centroid routing, then the member rows of the nprobe
    best shards per query.  Returns (ids int64 [B,k], scores float64 [B,k], counts int32 [B], certified int32 [B],
    probe int64 [B,nprobe] or None)."""
    dev = _dev(rows)
    assert shadow.dtype == torch.float16 and cen_shadow.dtype == torch.float16 and shard_offsets.dtype == torch.int64
    assert centroids.dtype == torch.float32 and centroids.stride(1) == 1 and cen_norms.dtype == torch.float64
    assert queries.dtype == torch.float32 and queries.stride(1) == 1 and rows.stride(1) == 1
    B, S = queries.shape[0], shard_offsets.numel() - 1
    P = max(1, min(int(nprobe), S))
    lib = _lib.load()
    with torch.cuda.device(dev):
        need = int(lib.lz_search_two_level_scratch_bytes(S, B, k, P, dim))
        if scratch is None or scratch.numel() < need:
            scratch = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
        ids = torch.empty((B, k), dtype=torch.int64, device=dev)
        scores = torch.empty((B, k), dtype=torch.float64, device=dev)
        counts = torch.empty((B,), dtype=torch.int32, device=dev)
        cert = torch.empty((B,), dtype=torch.int32, device=dev)
        probe = torch.empty((B, P), dtype=torch.int64, device=dev) if want_probe else None
        check(lib.lz_search_two_level(ptr(shadow), ptr(rows), rows.stride(0), dim, ptr(norms), int(n_rows_total), ptr(shard_offsets), S,
                                      ptr(cen_shadow), ptr(centroids), centroids.stride(0), ptr(cen_norms),
                                      -1.0 if cen_det_row_err is None else float(cen_det_row_err),
                                      ptr(queries), queries.stride(0), B, k, P, float(margin_sigmas), int(metric), float(max_row_norm),
                                      -1.0 if det_row_err is None else float(det_row_err), int(id_base),
                                      ptr(scratch), scratch.numel(), ptr(ids), ptr(scores), ptr(counts), ptr(cert), ptr(probe),
                                      stream_ptr(dev)), "lz_search_two_level")
    return ids, scores, counts, cert, probe, scratch


def assign_nearest(table_shadow: torch.Tensor, table: torch.Tensor, table_norms: torch.Tensor, rows: torch.Tensor, dim: int,
                   table_det_row_err: Optional[float] = None, row_shadow: Optional[torch.Tensor] = None,
                   row_err: Optional[torch.Tensor] = None, margin_sigmas: float = 4.0, scratch: Optional[torch.Tensor] = None):
    """Nearest table row (centroid) of every row by cosine, ties to the lowest table index (lz_assign_nearest).
    Returns (idx int64 [n], cos float64 [n], certified int32 [n])."""
    dev = _dev(rows)
    assert table_shadow.dtype == torch.float16 and table.dtype == torch.float32 and rows.dtype == torch.float32
    assert rows.stride(1) == 1 and table.stride(1) == 1
    n, K = rows.shape[0], table.shape[0]
    if row_shadow is not None:
        assert row_shadow.dtype == torch.float16 and row_shadow.is_contiguous() and row_shadow.shape == (n, dim)
        assert row_err is not None and row_err.dtype == torch.float32 and row_err.numel() >= n
    lib = _lib.load()
    with torch.cuda.device(dev):
        need = int(lib.lz_assign_nearest_scratch_bytes(K, n, dim, int(row_shadow is not None)))
        if scratch is None or scratch.numel() < need:
            scratch = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
        idx = torch.empty(n, dtype=torch.int64, device=dev)
        cos = torch.empty(n, dtype=torch.float64, device=dev)
        cnt = torch.empty(n, dtype=torch.int32, device=dev)
        cert = torch.empty(n, dtype=torch.int32, device=dev)
        check(lib.lz_assign_nearest(ptr(table_shadow), ptr(table), table.stride(0), dim, ptr(table_norms), K,
                                    -1.0 if table_det_row_err is None else float(table_det_row_err),
                                    ptr(rows), rows.stride(0), n, ptr(row_shadow), ptr(row_err), float(margin_sigmas),
                                    ptr(scratch), scratch.numel(), ptr(idx), ptr(cos), ptr(cnt), ptr(cert), stream_ptr(dev)),
              "lz_assign_nearest")
    return idx, cos, cert


def group_by_label(labels: torch.Tensor, n_labels: int):
    """(order int64 [n], seg_offsets int64 [n_labels + 1]): row indices grouped by label, ascending inside a label
    (lz_group_by_label, a stable counting sort)."""
    dev = _dev(labels)
    assert labels.dtype == torch.int64 and labels.is_contiguous()
    n = labels.numel()
    lib = _lib.load()
    order = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
    seg = torch.empty(n_labels + 1, dtype=torch.int64, device=dev)
    nbytes = int(lib.lz_group_by_label_scratch_bytes(n, n_labels))
    scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.lz_group_by_label(ptr(labels), n, int(n_labels), ptr(scratch), nbytes, ptr(order), ptr(seg), stream_ptr(dev)),
              "lz_group_by_label")
    return order, seg


def segment_mean(rows: torch.Tensor, dim: int, seg_offsets: torch.Tensor, order: Optional[torch.Tensor] = None) -> torch.Tensor:
    """float64 [n_segments, dim] means of row groups (lz_segment_mean)."""
    dev = _dev(rows)
    assert seg_offsets.dtype == torch.int64 and (order is None or order.dtype == torch.int64)
    n_seg = seg_offsets.numel() - 1
    out = torch.empty((n_seg, dim), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        check(_lib.load().lz_segment_mean(ptr(rows), rows.stride(0), dim, ptr(order), ptr(seg_offsets), n_seg, ptr(out),
                                          stream_ptr(dev)), "lz_segment_mean")
    return out


def connected_components(src: torch.Tensor, dst: torch.Tensor, n_edges: int, n_nodes: int) -> torch.Tensor:
    """int32 [n_nodes] labels = smallest node index of each node's component (lz_connected_components)."""
    dev = _dev(src)
    labels = torch.empty(n_nodes, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.load().lz_connected_components(ptr(src), ptr(dst), int(n_edges), int(n_nodes), ptr(labels),
                                                  stream_ptr(dev)), "lz_connected_components")
    return labels


def edge_endpoints_to_global(src, dst, weight, n: int, slot_global: torch.Tensor, slot_resident: torch.Tensor,
                             out_src: torch.Tensor, out_dst: torch.Tensor, out_walk: torch.Tensor, min_weight: float = 0.0):
    """One shard's edge endpoints as global node indices + walk flags (lz_edge_endpoints_to_global); outputs are
    caller-provided slices of the concatenated arrays."""
    dev = _dev(slot_global)
    assert slot_global.dtype == torch.int32 and slot_resident.dtype == torch.uint8 and out_walk.dtype == torch.uint8
    assert out_src.dtype == torch.int32 and out_dst.dtype == torch.int32 and out_src.numel() >= n
    with torch.cuda.device(dev):
        check(_lib.load().lz_edge_endpoints_to_global(ptr(src), ptr(dst), ptr(weight), int(n), ptr(slot_global), ptr(slot_resident),
                                                      slot_global.numel(), float(min_weight), ptr(out_src), ptr(out_dst),
                                                      ptr(out_walk), stream_ptr(dev)), "lz_edge_endpoints_to_global")


def reach_components(src: torch.Tensor, dst: torch.Tensor, walk: Optional[torch.Tensor], n_edges: int, n_nodes: int) -> torch.Tensor:
    """int32 [n_nodes]: labels[v] = smallest index among the nodes that can reach v under the walk flags
    (lz_reach_components) — the reference traversal's components when indices follow its visiting order."""
    dev = _dev(src)
    lib = _lib.load()
    labels = torch.empty(n_nodes, dtype=torch.int32, device=dev)
    nbytes = int(lib.lz_reach_components_scratch_bytes(int(n_edges), int(n_nodes)))
    scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.lz_reach_components(ptr(src), ptr(dst), ptr(walk), int(n_edges), int(n_nodes), ptr(scratch), nbytes,
                                      ptr(labels), stream_ptr(dev)), "lz_reach_components")
    return labels


def evict_lowest(salience: torch.Tensor, access_count: torch.Tensor, last_accessed: torch.Tensor, now: float, k: int,
                 skip: Optional[torch.Tensor] = None, want_all: bool = False):
    """The k least important nodes (lz_evict_lowest): (idx int64 [count], importance float64 [count]) in
    (importance asc, index asc) order, plus every node's importance when want_all.  salience float32 or
    float64 [n]; access_count int32 [n]; last_accessed float64 [n]; skip uint8 [n] or None."""
    dev = _dev(salience)
    n = salience.numel()
    assert salience.dtype in (torch.float32, torch.float64) and access_count.dtype == torch.int32
    assert last_accessed.dtype == torch.float64 and (skip is None or skip.dtype == torch.uint8)
    lib = _lib.load()
    k = int(k)
    kk = max(k, 1)
    idx = torch.empty(kk, dtype=torch.int64, device=dev)
    imp = torch.empty(kk, dtype=torch.float64, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    all_imp = torch.empty(n, dtype=torch.float64, device=dev) if want_all else None
    nbytes = int(lib.lz_evict_lowest_scratch_bytes(n, k))
    scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.lz_evict_lowest(ptr(salience), int(salience.dtype == torch.float64), ptr(access_count), ptr(last_accessed),
                                  ptr(skip), n, float(now), k, ptr(scratch), nbytes, ptr(idx), ptr(imp), ptr(cnt),
                                  ptr(all_imp), stream_ptr(dev)), "lz_evict_lowest")
    c = int(cnt.item())
    return (idx[:c], imp[:c], all_imp) if want_all else (idx[:c], imp[:c])


def csr_build(src: torch.Tensor, dst: torch.Tensor, n_edges: int, n_nodes: int):
    """(offsets int64 [n_nodes+1], adj int32-as-uint32 [2*n_edges]) — lz_csr_build; node v's entries are
    adj[offsets[v]:offsets[v+1]] in ascending edge order, entry = 2*edge + (1 if v is the dst end)."""
    dev = _dev(src)
    lib = _lib.load()
    offsets = torch.empty(n_nodes + 1, dtype=torch.int64, device=dev)
    adj = torch.empty(max(2 * n_edges, 1), dtype=torch.int32, device=dev)       # uint32 payload, < 2^31
    nbytes = int(lib.lz_csr_build_scratch_bytes(int(n_edges), int(n_nodes)))
    scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.lz_csr_build(ptr(src), ptr(dst), int(n_edges), int(n_nodes), ptr(scratch), nbytes, ptr(offsets), ptr(adj),
                               stream_ptr(dev)), "lz_csr_build")
    return offsets, adj


def neighbors(offsets: torch.Tensor, adj: torch.Tensor, src: torch.Tensor, dst: torch.Tensor, weight: torch.Tensor,
              n_nodes: int, nodes: torch.Tensor, min_weight: float = 0.3, max_out: int = 64, want_edges: bool = False):
    """Per queried node the neighbours over edges with weight >= min_weight, in edge order (lz_neighbors).
    Returns (nbr int32 [q, width], count int32 [q][, edge int32 [q, width]]); width grows until nothing is truncated."""
    dev = _dev(offsets)
    assert nodes.dtype == torch.int32 and weight.dtype == torch.float32
    q = nodes.numel()
    lib = _lib.load()
    cnt = torch.zeros(q, dtype=torch.int32, device=dev)
    while True:
        nbr = torch.full((q, max(max_out, 1)), -1, dtype=torch.int32, device=dev)
        edge = torch.full((q, max(max_out, 1)), -1, dtype=torch.int32, device=dev) if want_edges else None
        with torch.cuda.device(dev):
            check(lib.lz_neighbors(ptr(offsets), ptr(adj), ptr(src), ptr(dst), ptr(weight), int(n_nodes), ptr(nodes), q,
                                   float(min_weight), int(max_out), ptr(nbr), ptr(edge), ptr(cnt), stream_ptr(dev)), "lz_neighbors")
        most = int(cnt.max().item()) if q else 0
        if most <= max_out:
            return (nbr, cnt, edge) if want_edges else (nbr, cnt)
        max_out = most


def edge_find(offsets: torch.Tensor, adj: torch.Tensor, dst: torch.Tensor, n_nodes: int, q_src: torch.Tensor,
              q_dst: torch.Tensor) -> torch.Tensor:
    """int32 [q]: index of the edge (q_src, q_dst) in the table indexed by csr_build, or -1 (lz_edge_find)."""
    dev = _dev(offsets)
    assert q_src.dtype == torch.int32 and q_dst.dtype == torch.int32
    out = torch.empty(q_src.numel(), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.load().lz_edge_find(ptr(offsets), ptr(adj), ptr(dst), int(n_nodes), ptr(q_src), ptr(q_dst), q_src.numel(),
                                       ptr(out), stream_ptr(dev)), "lz_edge_find")
    return out


def edge_reinforce(weight: torch.Tensor, cooc: torch.Tensor, n_edges: int, edge_idx: torch.Tensor,
                   times: Optional[torch.Tensor] = None) -> None:
    """weight[e] = min(1, weight[e] + 0.1) `times` times, cooc[e] += times, in place (lz_edge_reinforce)."""
    dev = _dev(weight)
    assert edge_idx.dtype == torch.int32 and (times is None or times.dtype == torch.int32)
    with torch.cuda.device(dev):
        check(_lib.load().lz_edge_reinforce(ptr(weight), ptr(cooc), int(n_edges), ptr(edge_idx), ptr(times), edge_idx.numel(),
                                            stream_ptr(dev)), "lz_edge_reinforce")
