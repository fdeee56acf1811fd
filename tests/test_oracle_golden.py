"""The oracle (oracle/*.py) against outputs of the reference's own code (tests/golden/, written by
tests/golden/make_golden.py) and against the expectations of the reference's own tests.  CPU only."""
import json

import numpy as np
import pytest

from oracle import flat_search, graph_ops
from oracle.flat_store import FlatStore


def test_cosine_matches_reference(golden):
    # memory_system.py:197-203
    a, b, ref = golden["cos_a"], golden["cos_b"], golden["cos_out"]
    got = np.array([graph_ops.cosine(x.tolist(), y.tolist()) for x, y in zip(a, b)])
    np.testing.assert_array_equal(got, ref)
    assert ref[5] == 0.0 and float(golden["cos_empty"]) == 0.0
    assert graph_ops.cosine([], a[0].tolist()) == 0.0
    m = graph_ops.cosine_matrix(a, b)
    np.testing.assert_allclose(np.diag(m), ref, rtol=1e-13, atol=1e-15)
    # flat search scores are the same quantity
    for i in (0, 6, 17):
        s = flat_search.flat_scores(b, a[i], "cosine")
        np.testing.assert_allclose(s[i], ref[i], rtol=1e-13, atol=1e-15)


def test_decay_prune_matches_reference(golden):
    # memory_shard.py:64-84 run on Edge objects seeded with the same float32 weights
    g = golden
    src, dst, w, cooc, n = *graph_ops.decay_prune_edges(g["dp_src"], g["dp_dst"], g["dp_w"], g["dp_cooc"],
                                                       graph_ops.decay_factor(0.01), 0.5)[:4], None
    pruned = g["dp_w"].shape[0] - src.shape[0]
    assert pruned == int(g["dp_pruned"])
    np.testing.assert_array_equal(np.stack([src, dst], 1), g["dp_kept"])          # same set, same order
    np.testing.assert_array_equal(w, g["dp_kept_w"].astype(np.float32))           # fp32 rounding of ref fp64
    w64 = g["dp_w"].astype(np.float64) * 0.99
    np.testing.assert_array_equal(w64, g["dp_w_after"])                            # exact float64 product
    # the hand-placed boundary weights: 0.5 -> 0.495 pruned; 0.50505048*0.99 < 0.5 pruned; ...
    keep_head = ~(w64[:8] < 0.5)
    assert keep_head.tolist() == [False, False, False, True, False, True, False, False]
    assert w64[2] < 0.5 <= w64[3]                                                  # straddles the strict '<'


def test_salience_decay_matches_reference(golden):
    got = graph_ops.decay_salience(golden["dp_sal"], 0.01)
    np.testing.assert_array_equal(got, golden["dp_sal_after"].astype(np.float32))
    assert got[0] == np.float32(0.2) and got[1] == np.float32(0.2)


def _edges(js):
    return [(a, b, w, c) for a, b, w, c in json.loads(str(js))]


def test_cross_linking_matches_reference(golden):
    # memory_system.py:838-891 alone: top-3 > 0.5 against every existing row, weight 0.8*sim
    g = golden
    ref = [e for e in _edges(g["ln_cross"])]
    got = graph_ops.link_candidates(g["ln_new"], g["ln_rows"], 3, 0.5)
    flat = [(f"x{i}", f"e{j}", w) for i, lst in enumerate(got) for j, s, w in lst]
    assert [(a, b) for a, b, _, _ in ref] == [(a, b) for a, b, _ in flat]
    np.testing.assert_allclose([w for *_, w, _ in ref], [w for *_, w in flat], rtol=1e-12)


def test_within_shard_linking_matches_reference(golden):
    # memory_system.py:797-836: chain edges 0.5 between consecutive new nodes of a shard, then top-3 > 0.5
    # against the shard's other (non-new) nodes.
    g = golden
    ref = _edges(g["ln_within"])
    rows, shard, new, new_shard = g["ln_rows"], g["ln_shard"], g["ln_new"], g["ln_new_shard"]
    want = {}
    for s in np.unique(new_shard):
        ids = np.nonzero(new_shard == s)[0]
        for a, b in zip(ids[:-1], ids[1:]):
            want[(f"x{a}", f"x{b}")] = 0.5
        members = np.nonzero(shard == s)[0]
        links = graph_ops.link_candidates(new[ids], rows[members], 3, 0.5)
        for a, lst in zip(ids, links):
            for j, sim, w in lst:
                want[(f"x{a}", f"e{members[j]}")] = w
    assert {(a, b) for a, b, _, _ in ref} == set(want)
    for a, b, w, _ in ref:
        assert w == pytest.approx(want[(a, b)], rel=1e-12)


def test_dedup_matches_reference(golden):
    # memory_system.py:719-742
    g = golden
    best, sim, dup = graph_ops.dedup_matches(g["dd_new"], g["dd_rows"], gate=0.95)
    merged_into = np.zeros(g["dd_rows"].shape[0], np.int64)
    sal = np.full(g["dd_rows"].shape[0], 0.5)
    for i in np.nonzero(dup)[0]:
        merged_into[best[i]] += 1
        sal[best[i]] = max(sal[best[i]], g["dd_sal_new"][i])
    np.testing.assert_array_equal(merged_into, g["dd_access"])
    np.testing.assert_array_equal(sal, g["dd_sal_after"])
    assert g["dd_rows"].shape[0] + int((~dup).sum()) == int(g["dd_nodes_after"])
    # the planted near-duplicates are exactly the merged ones
    np.testing.assert_array_equal(best[dup], g["dd_dup_src"][dup])
    assert (g["dd_dup_src"] >= 0).sum() == dup.sum()


def test_centroid_matches_reference(golden):
    # memory_system.py:916-917
    got = graph_ops.centroid_means(golden["ce_rows"], np.zeros(golden["ce_rows"].shape[0], np.int64), 1)[0]
    np.testing.assert_allclose(got, golden["ce_centroid"], rtol=1e-14, atol=1e-17)


def test_search_memories_matches_reference(golden):
    # memory_system.py:1460-1472 over the same flat store: ids in order, non-resident rows dropped
    g = golden
    missing = set(g["sm_missing"].tolist())
    ref = json.loads(str(g["sm_ids"]))
    for q, want in zip(g["sm_queries"], ref):
        idx, _ = flat_search.flat_topk(g["sm_rows"], q, 10, "cosine")
        got = [f"node_{i}" for i in idx if i not in missing]
        assert got == want
    assert any(len(w) < 10 for w in ref)


def test_reference_test_pins(golden):
    g = golden
    # tests/test_consolidation_semantic.py:63-71
    assert int(g["pin_nodes"]) == 1 and int(g["pin_access"]) == 1
    assert float(g["pin_salience"]) == pytest.approx(0.893, abs=5e-4)
    s = graph_ops.decay_salience(np.float32([0.9]), 0.01)[0]
    assert float(s) == pytest.approx(0.893, abs=5e-4)
    # tests/test_profile_update.py:32-78: one component of 4, edges 0.8 survive prune 0.5
    assert json.loads(str(g["pin_components"])) == [["seed_0", "seed_1", "seed_2", "seed_3"]]
    assert int(g["pin_pruned"]) == 0
    # _merge_similar_nodes is a no-op as written
    assert int(g["pin_merge_noop"]) == 0


def _legacy_nodes():
    return [
        {"id": "node_1", "content": "I like apples", "embedding": [0.1] * 1536, "type": "semantic",
         "salience": 0.8, "shard_key": "food", "timestamp": 1234.5},
        {"id": "node_2", "content": "I hate oranges", "embedding": [-0.1] * 1536, "type": "semantic",
         "salience": 0.5, "shard_key": "food", "timestamp": 1235.0},
    ]


@pytest.mark.parametrize("metric", ["cosine", "l2"])
def test_flat_store_reference_vector_store_tests(metric):
    # tests/test_vector_store.py:22-68 retargeted at the legacy add/search/delete surface
    vs = FlatStore(metric=metric)
    vs.add(_legacy_nodes())
    assert vs.search([0.1] * 1536, limit=1) == ["node_1"]
    vs.delete(["node_1"])
    assert vs.search([0.1] * 1536, limit=1) == ["node_2"]
    vs.delete(["node_2"])
    assert vs.search([0.1] * 1536, limit=1) == []
    # empty id list = delete every row of the user (vector_store.py:143-145)
    vs.add_nodes(_legacy_nodes(), user_id="u1")
    vs.add_nodes(_legacy_nodes()[:1], user_id="u2")
    vs.delete_nodes([], user_id="u1")
    assert vs.search_nodes([0.1] * 1536, user_id="u1") == []
    assert vs.search_nodes([0.1] * 1536, user_id="u2") == ["node_1"]


def test_flat_search_ties_and_filter():
    rows = np.float32([[1, 0], [1, 0], [0, 1], [2, 0], [0, 0]])
    idx, sc = flat_search.flat_topk(rows, [1, 0], 4, "cosine")
    assert idx.tolist() == [0, 1, 3, 2] and sc[:3].tolist() == [1.0, 1.0, 1.0]     # ties by index; zero row -> 0.0
    idx, _ = flat_search.flat_topk(rows, [1, 0], 5, "cosine")
    assert idx.tolist() == [0, 1, 3, 2, 4]
    idx, _ = flat_search.flat_topk(rows, [1, 0], 2, "cosine", mask=np.array([0, 1, 1, 1, 1], bool))
    assert idx.tolist() == [1, 3]
    idx, d = flat_search.flat_topk(rows, [1, 0], 3, "l2")
    assert idx.tolist() == [0, 1, 3] and d.tolist() == [0.0, 0.0, 1.0]
    ids, sc = flat_search.merge_topk([[0, 3], [1, 2]], [[1.0, 0.5], [1.0, 0.7]], 3)
    assert ids.tolist() == [0, 1, 2]


def test_eviction_matches_reference(golden_next):
    # memory_system.py:535-578 run under a frozen clock: which nodes the reference deleted, in its order
    g = golden_next
    for c in range(3):
        sal, acc, last, sup = g[f"ev{c}_sal"], g[f"ev{c}_acc"], g[f"ev{c}_last"], g[f"ev{c}_super"]
        excess = len(sal) - int(g[f"ev{c}_limit"])
        idx, imp = graph_ops.evict_lowest(sal, acc, last, float(g["ev_now"]), excess, skip=sup)
        assert g[f"ev{c}_order"][idx].tolist() == g[f"ev{c}_deleted"].tolist()
        assert (np.diff(imp) >= 0).all() and not sup[idx].any()
        left = sorted(set(g[f"ev{c}_order"].tolist()) - set(g[f"ev{c}_deleted"].tolist()))
        assert left == g[f"ev{c}_left_nodes"].tolist()


def test_neighbors_match_reference(golden_next):
    # memory_shard.py:54-62 on 1500 edges (with reinforced repeats and weights exactly at the gate)
    import json
    g = golden_next
    lists = json.loads(str(g["nb_lists"]))
    for mw, want in lists.items():
        for node in range(0, int(g["nb_n"]), 3):
            assert graph_ops.neighbors(g["nb_src"], g["nb_dst"], g["nb_w"], node, float(mw)) == want[node]


def _components_case_arrays(case):
    """Global indices in the reference's visiting order + per-edge walk flags, as BufferGraph.get_connected_components
    of the device mirror derives them (residency of each endpoint in the edge's own shard)."""
    supers, nodes, edges = case["supers"], case["nodes"], case["edges"]
    shards = {}
    for nid, key in nodes:
        shards.setdefault(key, []).append(nid)
    home = {nid: key for nid, key in nodes}
    index = {}
    for nid in list(supers) + [nid for key in shards for nid in shards[key]]:
        index.setdefault(nid, len(index))
    n_res = len(index)
    seen, src, dst, walk = set(), [], [], []
    for s, t, w in edges:
        key = home.get(s, "default" if "default" in shards else None)
        if key is None or (key, s, t) in seen:
            continue
        seen.add((key, s, t))
        for x in (s, t):
            index.setdefault(x, len(index))
        src.append(index[s]); dst.append(index[t])
        walk.append((1 if home.get(s) == key else 0) | (2 if home.get(t) == key else 0))
    return index, n_res, np.array(src), np.array(dst), np.array(walk)


def test_components_oracle_and_min_ancestor_form_match_reference():
    """buffer_graph.py:99-120 recorded from the reference on graphs with cross-shard, dangling and super-node edges
    (tests/golden/make_golden_r2.py): the oracle's DFS restatement reproduces it, and so does the order-free
    formulation the CUDA kernel implements (label = smallest index that can reach the node)."""
    import json
    import os
    from oracle import graph_ops
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_r2.npz"))
    cases = json.loads(str(g["comps"]))
    order_dependent = 0
    for case in cases:
        want = [sorted(c) for c in case["components"]]
        got = [sorted(c) for c in graph_ops.reference_components(case["supers"], case["nodes"], case["edges"])]
        assert got == want
        index, n_res, src, dst, walk = _components_case_arrays(case)
        lab = graph_ops.min_ancestor_labels(len(index), src, dst, walk)
        names = list(index)
        groups = {}
        for i, l in enumerate(lab.tolist()):
            if i < n_res or l != i:
                groups.setdefault(l, []).append(names[i])
        assert [sorted(groups[k]) for k in sorted(groups)] == want
        undirected = graph_ops.min_ancestor_labels(len(index), src, dst, np.where(walk > 0, 3, 0))
        order_dependent += int(len(set(undirected[:n_res].tolist())) != len([c for c in want]))
    assert order_dependent >= 2, "the fixture should include graphs where one-way edges change the components"


def test_restated_reference_loops_match_reference(golden):
    """oracle/ref_loops.py (the per-object Python loops bench.py times as the reference-shaped CPU baseline) against
    what the reference's own loops produced: memory_shard.py:64-84 and memory_system.py:838-891."""
    from oracle import ref_loops
    g = golden
    edges = ref_loops.make_edge_dict(g["dp_src"], g["dp_dst"], g["dp_w"])
    nodes = {i: ref_loops.Node(i, [], float(s)) for i, s in enumerate(g["dp_sal"])}
    pruned = ref_loops.decay_and_prune(edges, nodes, 0.01, 0.5)
    assert pruned == int(g["dp_pruned"])
    np.testing.assert_array_equal(np.array([[a, b] for a, b in edges], np.int32), g["dp_kept"])
    np.testing.assert_array_equal(np.array([e.weight for e in edges.values()]), g["dp_kept_w"])
    np.testing.assert_array_equal(np.array([n.salience for n in nodes.values()]), g["dp_sal_after"])
    ref = _edges(g["ln_cross"])
    rows, new = g["ln_rows"], g["ln_new"]
    existing = [ref_loops.Node(f"e{i}", rows[i].astype(np.float64).tolist()) for i in range(rows.shape[0])]
    fresh = [ref_loops.Node(f"x{i}", new[i].astype(np.float64).tolist()) for i in range(new.shape[0])]
    made = {}
    assert ref_loops.link_to_existing(fresh, existing, made) == len(ref)
    assert [(a, b) for a, b, _, _ in ref] == list(made)
    np.testing.assert_array_equal([w for _, _, w, _ in ref], [e.weight for e in made.values()])
