"""Oracle: the Python-side arithmetic of the hot path, restated in numpy.
TEST INFRASTRUCTURE — see oracle/__init__.py.

Each function cites the reference lines it follows (paths under /root/reference/src/lazzaro/).
State that the reference keeps as Python float64 but persists as float32 (edge weight
vector_store.py:63, salience vector_store.py:42) is float32 here: one sweep reads float32, computes
in float64 exactly as the reference's Python does, and stores the float32 rounding.
"""
from __future__ import annotations

import numpy as np


def cosine(v1, v2) -> float:
    """core/memory_system.py:197-203 — float64 cosine; 0.0 for an empty vector or zero norm."""
    if v1 is None or v2 is None or len(v1) == 0 or len(v2) == 0:
        return 0.0
    a, b = np.array(v1, dtype=np.float64), np.array(v2, dtype=np.float64)
    norm = np.linalg.norm(a) * np.linalg.norm(b)
    return float(np.dot(a, b) / norm) if norm > 0 else 0.0


def cosine_matrix(a32: np.ndarray, b32: np.ndarray) -> np.ndarray:
    """All-pairs float64 cosine of float32 rows (memory_system.py:197-203 applied pairwise); each
    entry is formed as sum_i a_i*b_i over i in a position-independent way."""
    a = np.ascontiguousarray(a32, dtype=np.float32).astype(np.float64)
    b = np.ascontiguousarray(b32, dtype=np.float32).astype(np.float64)
    na = np.sqrt((a * a).sum(axis=1))
    nb = np.sqrt((b * b).sum(axis=1))
    out = np.empty((a.shape[0], b.shape[0]), dtype=np.float64)
    for i in range(a.shape[0]):
        dot = (b * a[i]).sum(axis=1)
        den = nb * na[i]
        with np.errstate(divide="ignore", invalid="ignore"):
            s = dot / den
        s[~(den > 0)] = 0.0
        out[i] = s
    return out


# --- decay / prune ------------------------------------------------------------------------------

def decay_factor(decay_rate: float = 0.01, sweeps: int = 1) -> float:
    """core/memory_shard.py:70: each sweep multiplies by (1 - decay_rate); `sweeps` folds Δ sweeps
    into one factor (the "exponential decay" of north_star)."""
    return float((1.0 - decay_rate) ** sweeps)


def decay_prune_edges(src, dst, weight32, cooc, factor: float, threshold: float):
    """core/memory_shard.py:69-70 then 79-84 fused: w *= factor (float64 product of the float32
    weight), drop edges with w < threshold (strict, compared in float64), keep survivors in their
    original order (dict order in the reference), store float32.
    Returns (src, dst, weight32, cooc, n_pruned)."""
    w64 = np.asarray(weight32, dtype=np.float32).astype(np.float64) * np.float64(factor)
    keep = ~(w64 < np.float64(threshold))
    return (np.asarray(src)[keep], np.asarray(dst)[keep], w64[keep].astype(np.float32),
            np.asarray(cooc)[keep], int((~keep).sum()))


def decay_salience(salience32, decay_rate: float = 0.01, floor: float = 0.2):
    """core/memory_shard.py:72-77: s = floor + (s - floor)*(1 - rate) if s > floor else floor."""
    s = np.asarray(salience32, dtype=np.float32).astype(np.float64)
    out = np.where(s > floor, floor + (s - floor) * (1.0 - decay_rate), floor)
    return out.astype(np.float32)


# --- consolidation ------------------------------------------------------------------------------

def top_m_above(sims: np.ndarray, m: int = 3, threshold: float = 0.5):
    """core/memory_system.py:827-836 / 865-867: stable sort descending, first m, keep sim > thr.
    Returns list of (index, sim)."""
    order = np.argsort(-sims, kind="stable")[:m]
    return [(int(i), float(sims[i])) for i in order if sims[i] > threshold]


def link_candidates(new32, existing32, m: int = 3, threshold: float = 0.5, edge_scale: float = 0.8):
    """core/memory_system.py:838-891 (and 816-836): for each new row, cosine vs every existing row,
    top-m above threshold; edge weight = edge_scale * sim (memory_system.py:833,881).
    Returns list per new row of (existing_index, sim, weight)."""
    sims = cosine_matrix(new32, existing32)
    return [[(i, s, s * edge_scale) for i, s in top_m_above(row, m, threshold)] for row in sims]


def dedup_matches(new32, graph32, mask=None, gate: float = 0.95, metric: str = "cosine"):
    """core/memory_system.py:719-742: per new memory, top-1 of the flat search, then cosine with
    that row; a match if cosine > gate.  New rows do not see each other (add_nodes happens after
    the loop, memory_system.py:768-769).  Returns (best_index int64 [-1 if none], cosine float64,
    is_duplicate bool)."""
    from . import flat_search
    n = np.asarray(new32).shape[0]
    best = np.full(n, -1, np.int64)
    sim = np.zeros(n, np.float64)
    for i in range(n):
        idx, _ = flat_search.flat_topk(graph32, new32[i], 1, metric, mask)
        if idx.size:
            best[i] = idx[0]
            sim[i] = cosine_matrix(new32[i:i + 1], np.asarray(graph32)[idx[0]:idx[0] + 1])[0, 0]
    return best, sim, sim > gate


def assign_to_centroids(rows32, centroids32):
    """No reference code: assign = argmax_c cosine(x, centroid_c), ties to the
    lowest centroid index.  Returns (index int64, cosine float64)."""
    sims = cosine_matrix(rows32, centroids32)
    idx = np.argmax(sims, axis=1)                      # first maximum = lowest index
    return idx.astype(np.int64), sims[np.arange(sims.shape[0]), idx]


def centroid_means(rows32, assign, n_centroids: int):
    """core/memory_system.py:916-917: centroid = np.mean(member embeddings, axis=0), un-normalised,
    float64 mean of the float32 rows; empty clusters give a zero vector. Returns float64 [K, D]."""
    rows = np.asarray(rows32, dtype=np.float32)
    out = np.zeros((n_centroids, rows.shape[1]), dtype=np.float64)
    for c in range(n_centroids):
        m = rows[np.asarray(assign) == c]
        if m.shape[0]:
            out[c] = np.mean(m.astype(np.float64), axis=0)
    return out


def importance(salience, access_count, days_old):
    """core/memory_system.py:545-550."""
    return (np.asarray(salience, np.float64) * 0.5
            + np.minimum(1.0, np.asarray(access_count, np.float64) / 10) * 0.3
            + (1.0 / (1.0 + np.asarray(days_old, np.float64))) * 0.2)


def evict_lowest(salience, access_count, last_accessed, now: float, k: int, skip=None):
    """core/memory_system.py:541-553 for one `now`: importance of every node in the caller's (= the
    reference's iteration) order, stable ascending sort, first k non-skipped. Returns (indices, importances)."""
    imp = importance(salience, access_count, (now - np.asarray(last_accessed, np.float64)) / 86400)
    cand = np.arange(len(imp)) if skip is None else np.nonzero(~np.asarray(skip, bool))[0]
    cand = cand[~np.isnan(imp[cand])]
    order = cand[np.argsort(imp[cand], kind="stable")][:k]
    return order.astype(np.int64), imp[order]


def neighbors(src, dst, weight, node: int, min_weight: float = 0.3):
    """core/memory_shard.py:54-62: neighbour ids of `node` in edge order (edge i = (src[i], dst[i])); an edge
    contributes once (its source end is tested first, so a self-loop yields the node itself once)."""
    out = []
    for s, t, w in zip(np.asarray(src).tolist(), np.asarray(dst).tolist(), np.asarray(weight, np.float64).tolist()):
        if s == node and w >= min_weight:
            out.append(t)
        elif t == node and w >= min_weight:
            out.append(s)
    return out


# --- connected components as the reference traverses them ------------------------------------------

def reference_components(super_ids, nodes, edges):
    """core/buffer_graph.py:99-120 restated: `nodes` = [(id, shard_key)] in insertion order, `edges` =
    [(source, target, weight)] in insertion order (an edge is stored in the shard of its SOURCE node, else in
    "default" if that shard exists — buffer_graph.py:51-61; a repeated key is reinforced, memory_shard.py:42-52).
    Visiting order = super-nodes, then shards in creation order with their nodes in insertion order
    (buffer_graph.py:28-34); from a node only the edges of ITS shard are followed, at weight >= 0.0
    (buffer_graph.py:72-77, 110).  Returns the list of components (sets) in visiting order."""
    shard_nodes, shard_edges = {}, {}
    for nid, key in nodes:
        shard_nodes.setdefault(key, []).append(nid)
        shard_edges.setdefault(key, {})
    home = {nid: key for nid, key in nodes}
    for s, t, w in edges:
        key = home.get(s, "default" if "default" in shard_nodes else None)
        if key is None:
            continue
        d = shard_edges[key]
        d[(s, t)] = min(1.0, d[(s, t)] + 0.1) if (s, t) in d else w
    def nbrs(nid):
        if nid not in home:
            return []
        out = []
        for (s, t), w in shard_edges[home[nid]].items():
            if s == nid and w >= 0.0:
                out.append(t)
            elif t == nid and w >= 0.0:
                out.append(s)
        return out
    order = list(super_ids) + [nid for key in shard_nodes for nid in shard_nodes[key]]
    seen, comps = set(), []
    for root in order:
        if root in seen:
            continue
        comp, stack = set(), [root]
        seen.add(root)
        while stack:                                   # iterative DFS: same visited sets as the recursive one
            cur = stack.pop()
            comp.add(cur)
            for nb in nbrs(cur):
                if nb not in seen:
                    seen.add(nb)
                    stack.append(nb)
        comps.append(comp)
    return comps


def min_ancestor_labels(n_nodes: int, src, dst, walk):
    """What lz_reach_components computes, by plain fixed-point iteration: labels[v] = the smallest index among the
    nodes that can reach v; walk bit 0 = src->dst allowed, bit 1 = dst->src allowed."""
    lab = np.arange(n_nodes, dtype=np.int64)
    changed = True
    while changed:
        changed = False
        for s, t, w in zip(src, dst, walk):
            if s < 0 or t < 0:
                continue
            if w & 1 and lab[s] < lab[t]:
                lab[t] = lab[s]; changed = True
            if w & 2 and lab[t] < lab[s]:
                lab[s] = lab[t]; changed = True
    return lab
