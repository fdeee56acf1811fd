"""Episodic-buffer consolidation on the device: dedup, linking, shard (re)assignment, centroids.

Host-side mirror of the arithmetic inside the reference's consolidation path
(/root/reference/src/lazzaro/core/memory_system.py):

  dedup_against_store   lines 719-742  per new memory: top-1 of the store, merge if cosine > 0.95
  link_candidates       lines 838-891 / 816-836  per new memory: top-3 existing rows with cosine > 0.5,
                        edge weight 0.8 * cosine
  assign_to_centroids   no reference code: argmax_c cosine(x, centroid_c), ties to the
                        lowest centroid index — the k-means style shard reassignment north_star names
  centroid_means        lines 916-917  centroid = arithmetic mean of the member embeddings

The reference issues one search per new memory, sequentially, and new memories of one batch never see
each other (they are added to the store after the loop, memory_system.py:768-769), so a single batched
search per step is semantically identical.  All of it is the same tensor-core contraction as the
batched search (`lz_search_batched`, k = 1 or 3) followed by exact float64 rescoring, so the decisions
(which rows, which side of the 0.95 / 0.5 gates) are those of the exact float64 cosine; the host only
applies the reference's rules to the results.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib, kernels

DEDUP_GATE = 0.95       # memory_system.py:731
LINK_GATE = 0.5         # memory_system.py:829, 867
LINK_TOP = 3            # memory_system.py:828, 866
LINK_SCALE = 0.8        # memory_system.py:833, 881


def _queries_on_device(store, emb) -> torch.Tensor:
    q = np.ascontiguousarray(emb, dtype=np.float32)
    if q.shape[1] != store.dim:
        raise ValueError(f"embeddings have {q.shape[1]} dims, store has {store.dim}")
    out = torch.zeros((q.shape[0], store._dim_p), dtype=torch.float32, device=store.device)
    out[:, :store.dim] = torch.from_numpy(q).to(store.device, non_blocking=True)
    return out


MARGIN = "worst_case"   # certification margin of the tensor-core searches below: the deterministic one, because a
                        # wrong top-1 here silently merges (or fails to merge) a memory; "statistical" is narrower
                        # (slightly fewer rows rescored) but assumes uncorrelated rounding errors


def dedup_against_store(store, new_emb, user_id: str = "default", gate: float = DEDUP_GATE, margin: str = MARGIN):
    """For each new embedding: the store's best row for `user_id` (ranked by the STORE's metric, as the reference's
    search_nodes call is) and whether it is a semantic duplicate (cosine > gate, whatever the store metric).
    Returns (ids: list of str or None, cosine float64 [n], is_duplicate bool [n])."""
    n = len(new_emb)
    u = store._user(user_id)
    if n == 0 or store.n_rows == 0 or u < 0 or store._live[u] == 0:
        return [None] * n, np.zeros(n), np.zeros(n, bool)
    with store._lock:
        ids, sc, cnt = store.search_device(_queries_on_device(store, new_emb), u, 1, margin=margin)
        torch.cuda.current_stream(store.device).synchronize()
        ids, sc, cnt = ids.cpu().numpy()[:, 0], sc.cpu().numpy()[:, 0], cnt.cpu().numpy()
        if store.metric != "cosine":                    # the gate is a cosine whatever the search metric
            q = _queries_on_device(store, new_emb)
            pq = torch.arange(n, dtype=torch.int32, device=store.device)
            sc = kernels.score_pairs(store._rows, store._dim_p, store._norms, q, pq,
                                     torch.from_numpy(ids).to(store.device), _lib.COSINE).cpu().numpy()
        names = [store._id_of(int(r)) if c else None for r, c in zip(ids, cnt)]
        sims = np.where(cnt > 0, sc, 0.0)
        return names, sims, (cnt > 0) & (sims > gate)


def link_candidates(store, new_emb, user_id: str = "default", m: int = LINK_TOP, threshold: float = LINK_GATE,
                    scale: float = LINK_SCALE, row_range: Optional[Tuple[int, int]] = None, margin: str = MARGIN):
    """For each new embedding the up-to-m best existing rows with cosine > threshold, best first, as
    (id, cosine, edge_weight = scale * cosine).  `row_range` restricts the scan to one shard's rows
    (_link_within_shards); the default scans every row of the user (_link_to_existing_memories).
    Always ranked and gated by COSINE, as memory_system.py:858-867 does over the resident nodes — also on a
    store whose own metric is l2 (the store then keeps a cosine shadow beside its l2 one)."""
    n = len(new_emb)
    u = store._user(user_id)
    if n == 0 or store.n_rows == 0 or u < 0 or store._live[u] == 0:
        return [[] for _ in range(n)]
    with store._lock:
        q = _queries_on_device(store, new_emb)
        ids, sc, cnt = store.search_device(q, u, m, row_range=row_range, metric="cosine", margin=margin)
        torch.cuda.current_stream(store.device).synchronize()
        ids, sc, cnt = ids.cpu().numpy(), sc.cpu().numpy(), cnt.cpu().numpy()
        out = []
        for i in range(n):
            out.append([(store._id_of(int(r)), float(s), float(s) * scale)
                        for r, s in zip(ids[i, :cnt[i]], sc[i, :cnt[i]]) if s > threshold])
        return out


def _centroid_table(centroids: torch.Tensor, margin: str):
    if margin not in ("worst_case", "statistical"):
        raise ValueError(f"margin must be 'worst_case' or 'statistical', got {margin!r}")
    cen = centroids.contiguous()
    norms = kernels.row_norms(cen)
    err = torch.empty(cen.shape[0], dtype=torch.float32, device=cen.device)
    shadow = kernels.rows_to_shadow(cen, cen.shape[1], norms, row_err=err)
    det = float(err.max().item()) if margin == "worst_case" else None       # K floats: a small read-back
    return cen, norms, shadow, det


def assign_to_centroids(rows: torch.Tensor, centroids: torch.Tensor, margin_sigmas: float = 4.0, margin: str = MARGIN):
    """argmax_c cosine(row, centroid_c) for float32 CUDA tensors rows [n, dim], centroids [K, dim]
    (dim % 8 == 0).  Returns (index int64 [n], cosine float64 [n]) on the device; ties go to the lowest
    centroid index.  Rows are the query side of the contraction, centroids the scanned side."""
    assert rows.is_cuda and centroids.is_cuda and rows.dtype == torch.float32 and centroids.dtype == torch.float32
    dim = rows.shape[1]
    n = rows.shape[0]
    if n >= 8 and dim % 8 == 0:
        cen, norms, shadow, det = _centroid_table(centroids, margin)
        idx, cos, _ = kernels.assign_nearest(shadow, cen, norms, rows, dim, table_det_row_err=det, margin_sigmas=margin_sigmas)
        return idx, cos
    cen = centroids.contiguous()
    ids, sc, cnt = kernels.search_exact(cen, dim, kernels.row_norms(cen), None, rows, 1, _lib.COSINE)
    return ids[:, 0], sc[:, 0]


def assign_store_rows(store, centroids: torch.Tensor, row_range: Optional[Tuple[int, int]] = None, margin_sigmas: float = 4.0):
    """Shard reassignment of rows RESIDENT in a store: the same argmax, with the store's own fp16 cosine shadow and
    per-row rounding-error norms as the query side, so no conversion pass over the rows is needed (deterministic
    margin).  Returns (index int64 [n], cosine float64 [n]) for rows row_range (default: all rows)."""
    with store._lock:
        sh = store._shadow_for(_lib.COSINE)
        lo, hi = (0, store.n_rows) if row_range is None else (int(row_range[0]), int(row_range[1]))
        cen, norms, shadow, det = _centroid_table(centroids, "worst_case")
        idx, cos, _ = kernels.assign_nearest(shadow, cen, norms, store._rows[lo:hi], store._dim_p, table_det_row_err=det,
                                             row_shadow=sh.t[lo:hi], row_err=sh.err[lo:hi], margin_sigmas=margin_sigmas)
        return idx, cos


def centroid_means(rows: torch.Tensor, assign: Optional[torch.Tensor] = None, n_centroids: Optional[int] = None,
                   seg_offsets: Optional[torch.Tensor] = None) -> torch.Tensor:
    """float64 [K, dim] arithmetic means (memory_system.py:916-917).  Either `seg_offsets` (int64 [K+1],
    members of centroid c are the contiguous rows seg_offsets[c]:seg_offsets[c+1] — rows stored shard by
    shard) or `assign` (int64 [n] centroid of each row, n_centroids <= 16384).  With `assign`, the member lists come
    from a stable counting sort on the device (lz_group_by_label); the summation is lz_segment_mean."""
    dim = rows.shape[1]
    if seg_offsets is not None:
        return kernels.segment_mean(rows, dim, seg_offsets.to(torch.int64))
    assert assign is not None and n_centroids is not None
    order, seg = kernels.group_by_label(assign.contiguous(), n_centroids)
    return kernels.segment_mean(rows, dim, seg, order)
