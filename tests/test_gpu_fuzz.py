"""GPU: seeded random sweep of shapes, k, metrics, filters and ranges — exact scan and tensor-core path
against the numpy oracle (ids bit-exact, scores to 1e-12)."""
import numpy as np
import pytest
import torch

from oracle import flat_search

pytestmark = pytest.mark.gpu


def _case(seed):
    rng = np.random.default_rng(seed)
    dim = int(rng.choice([8, 16, 24, 64, 72, 128, 200, 384, 520]))
    n = int(rng.choice([1, 7, 255, 256, 257, 1000, 4097, 20011]))
    B = int(rng.choice([1, 2, 5, 8, 9, 127, 129, 300]))
    k = int(rng.choice([1, 2, 3, 4, 10, 17, 64, 100]))
    return rng, dim, n, B, k


@pytest.mark.parametrize("seed", range(24))
def test_random_configurations(seed):
    from lazzaro_b200 import _lib, kernels as K, synth
    rng, dim, n, B, k = _case(seed)
    rows_np, _ = synth.clustered_rows(n, dim, max(1, min(8, n)), seed=seed, sigma=float(rng.choice([0.2, 0.6, 2.0])))
    rows_np *= rng.uniform(0.1, 5.0, size=(n, 1)).astype(np.float32)            # un-normalised rows
    if n > 3:
        rows_np[n // 2] = 0.0
        rows_np[n // 3] = rows_np[0]
    qs_np = synth.queries_near_rows(rows_np + 1e-3, B, seed=seed + 100, noise=float(rng.choice([0.05, 0.5, 3.0])))
    tags_np = rng.integers(-1, 3, size=n).astype(np.int32) if rng.random() < 0.5 else None
    want_tag = int(rng.integers(-1, 3)) if tags_np is not None else -1
    lo = int(rng.integers(0, max(1, n // 2))) if rng.random() < 0.5 else 0
    hi = int(rng.integers(lo, n + 1)) if rng.random() < 0.5 else n
    mask = np.zeros(n, bool); mask[lo:hi] = True
    if tags_np is not None:
        mask &= (tags_np >= 0) if want_tag < 0 else (tags_np == want_tag)
    rows, qs = torch.as_tensor(rows_np).cuda(), torch.as_tensor(qs_np).cuda()
    tags = None if tags_np is None else torch.as_tensor(tags_np).cuda()
    norms = K.row_norms(rows)
    kk = min(k, max(1, int(mask.sum())))
    for metric in ("cosine", "l2"):
        want_ids, want_sc = flat_search.flat_topk_batch(rows_np, qs_np, kk, metric, mask)
        gs = 1 if tags is not None else None
        gtag = None if tags is None else torch.full((B,), want_tag, dtype=torch.int32, device="cuda")
        ids, sc, cnt = K.search_exact(rows, dim, norms, tags, qs, kk, _lib.METRICS[metric], row_range=(lo, hi),
                                      group_tag=gtag, group_size=gs)
        np.testing.assert_array_equal(ids.cpu().numpy(), want_ids)
        ok = want_ids >= 0
        np.testing.assert_allclose(sc.cpu().numpy()[ok], want_sc[ok], rtol=1e-12, atol=1e-14)
        assert cnt.cpu().numpy().tolist() == ok.sum(1).tolist()
    if kk <= 128:
        for metric in ("cosine", "l2"):
            shadow = K.rows_to_shadow(rows, dim, norms, metric=_lib.METRICS[metric])
            want_ids, want_sc = flat_search.flat_topk_batch(rows_np, qs_np, kk, metric, mask)
            ids, sc, cnt, cert = K.search_batched(shadow, rows, dim, norms, tags, want_tag, qs, kk, row_range=(lo, hi),
                                                  metric=_lib.METRICS[metric])
            ids_np, ok = ids.cpu().numpy(), want_ids >= 0
            np.testing.assert_array_equal(ids_np, want_ids)
            np.testing.assert_allclose(sc.cpu().numpy()[ok], want_sc[ok], rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize("seed", range(12))
def test_random_graph_operations(seed):
    """Seeded sweep over the graph-side kernels against dict / numpy restatements of the reference rules:
    edge upsert (memory_shard.py:42-52), neighbours (54-62), decay + prune (64-84), components
    (buffer_graph.py:99-120), eviction order (memory_system.py:535-553)."""
    from lazzaro_b200 import kernels as K
    from lazzaro_b200.graph import Edge, MemoryShard, Node
    from oracle import graph_ops
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.choice([3, 17, 64, 300]))
    m = int(rng.choice([1, 40, 500, 3000]))
    sh = MemoryShard("s")
    for i in range(n):
        sh.add_node(Node(id=f"n{i}", content="", embedding=[]))
    table = {}                                                     # the reference's dict, restated on float32 storage
    src, dst = rng.integers(0, n, m), rng.integers(0, n, m)
    w = rng.uniform(0, 1, m).astype(np.float32)
    edges = [Edge(source=f"n{a}", target=f"n{b}", weight=float(x)) for a, b, x in zip(src, dst, w)]
    cut = int(rng.integers(0, m + 1))
    for e in edges[:cut]:                                          # one by one, then the rest as batches
        sh.add_edge(e)
    step = int(rng.choice([1, 7, 100, 5000]))
    for lo in range(cut, m, step):
        sh.add_edges(edges[lo:lo + step])
    for a, b, x in zip(src.tolist(), dst.tolist(), w.tolist()):
        if (a, b) in table:
            ww, cc = table[(a, b)]
            table[(a, b)] = (float(np.float32(min(1.0, ww + 0.1))), cc + 1)
        else:
            table[(a, b)] = (float(np.float32(x)), 1)
    sh._flush()
    es, ed, ew, ec = sh._table.to_numpy()
    assert list(zip(es.tolist(), ed.tolist())) == list(table)
    np.testing.assert_array_equal(ew, np.float32([v[0] for v in table.values()]))
    np.testing.assert_array_equal(ec, np.int32([v[1] for v in table.values()]))
    mw = float(rng.choice([0.0, 0.3, 0.55]))
    got = sh.get_neighbors_batch([f"n{i}" for i in range(n)], mw)
    for v in range(n):
        assert got[v] == [f"n{x}" for x in graph_ops.neighbors(es, ed, ew, v, mw)]
    # components before the sweep
    comps = sorted(sorted(int(x[1:]) for x in c) for c in sh.connected_components())
    parent = list(range(n))
    def find(x):
        while parent[x] != x:
            parent[x] = parent[parent[x]]; x = parent[x]
        return x
    for a, b in zip(es.tolist(), ed.tolist()):
        ra, rb = find(a), find(b)
        if ra != rb:
            parent[max(ra, rb)] = min(ra, rb)
    want = {}
    for v in range(n):
        want.setdefault(find(v), []).append(v)
    assert comps == sorted(want.values())
    # decay + prune, then the index must follow
    rate, thr = float(rng.choice([0.0, 0.01, 0.2])), float(rng.choice([0.1, 0.5, 0.9]))
    ks, kd, kw, kc, _ = graph_ops.decay_prune_edges(es, ed, ew, ec, 1.0 - rate, thr)
    pruned = sh.decay_and_prune(rate, thr)
    s2, d2, w2, c2 = sh._table.to_numpy()
    assert pruned == len(es) - len(ks)
    np.testing.assert_array_equal(s2, ks); np.testing.assert_array_equal(d2, kd)
    np.testing.assert_array_equal(w2, kw); np.testing.assert_array_equal(c2, kc)
    v = int(rng.integers(0, n))
    assert sh.get_neighbors(f"n{v}", 0.0) == [f"n{x}" for x in graph_ops.neighbors(s2, d2, w2, v, 0.0)]
    # eviction order on random node fields with many ties
    now = 1.7e9
    sal = rng.choice([0.2, 0.5, 0.55, 0.9], n)
    acc = rng.integers(0, 12, n).astype(np.int32)
    last = now - rng.choice([0.0, 86400.0, 86400.0 * 30], n)
    skip = rng.random(n) < 0.2
    k = int(rng.integers(1, n + 1))
    t = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a).astype(dt)).cuda()
    idx, imp = K.evict_lowest(t(sal, np.float64), t(acc, np.int32), t(last, np.float64), now, k, skip=t(skip, np.uint8))
    want_idx, want_imp = graph_ops.evict_lowest(sal, acc, last, now, k, skip=skip)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_idx)
    np.testing.assert_array_equal(imp.cpu().numpy(), want_imp)
