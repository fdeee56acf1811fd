"""GPU: B200Store behind the reference's Store protocol — the reference's own store tests,
retargeted (tests/test_vector_store.py, tests/test_lancedb_integration.py), and call-for-call
equivalence with the oracle FlatStore."""
import numpy as np
import pytest

from oracle.flat_store import FlatStore

pytestmark = pytest.mark.gpu


def legacy_nodes():
    return [
        {"id": "node_1", "content": "I like apples", "embedding": [0.1] * 1536, "type": "semantic",
         "salience": 0.8, "shard_key": "food", "timestamp": 1234.5},
        {"id": "node_2", "content": "I hate oranges", "embedding": [-0.1] * 1536, "type": "semantic",
         "salience": 0.5, "shard_key": "food", "timestamp": 1235.0},
    ]


@pytest.mark.parametrize("metric", ["cosine", "l2"])
def test_reference_vector_store_tests(metric):
    # tests/test_vector_store.py:22-68
    from lazzaro_b200.store import LanceDBVectorStore
    vs = LanceDBVectorStore(db_dir="test_vector_db", metric=metric)
    vs.add(legacy_nodes())
    results = vs.search([0.1] * 1536, limit=1)
    assert len(results) == 1 and results[0] == "node_1"
    vs.delete(["node_1"])
    assert vs.search([0.1] * 1536, limit=1) == ["node_2"]
    vs.delete(["node_2"])
    assert vs.search([0.1] * 1536, limit=1) == []
    vs.close()


def test_store_matches_flat_store_call_for_call():
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    rows, shard = synth.clustered_rows(3000, 96, 6, seed=14)
    qs = synth.queries_near_rows(rows, 12, seed=15)
    ours, ref = B200Store(capacity=64), FlatStore()
    users = ["alice", "bob", "carol"]
    v0 = ours.get_latest_version()
    for st in (ours, ref):
        assert st.search_nodes(qs[0].tolist()) == [] and st.get_nodes("alice") == []
        for lo in range(0, 3000, 500):
            u = users[(lo // 500) % 3]
            st.add_nodes([{"id": f"n{i}", "content": f"c{i}", "embedding": rows[i].tolist(),
                           "shard_key": f"s{shard[i]}", "salience": 0.25 + (i % 3) / 4,
                           "child_ids": [f"n{i+1}"], "metadata": {"i": i}} for i in range(lo, lo + 500)], user_id=u)
    assert ours.get_latest_version() > v0
    for u in users + ["nobody"]:
        for q in qs:
            for limit in (1, 5, 40):
                assert ours.search_nodes(q.tolist(), user_id=u, limit=limit) == ref.search_nodes(q.tolist(), user_id=u, limit=limit)
    assert ours.search_batch(qs, "bob", 7) == [ref.search_nodes(q.tolist(), "bob", 7) for q in qs]
    for st in (ours, ref):
        st.delete_nodes([f"n{i}" for i in range(0, 3000, 7)], user_id="alice")
        st.delete_nodes([], user_id="carol")                      # empty list = all of the user
    for u in users:
        for q in qs[:4]:
            assert ours.search_nodes(q.tolist(), user_id=u, limit=10) == ref.search_nodes(q.tolist(), user_id=u, limit=10)
    a, b = ours.get_nodes("alice"), ref.get_nodes("alice")
    assert [x["id"] for x in a] == [x["id"] for x in b]
    for x, y in zip(a[:20], b[:20]):
        assert x["vector"] == y["vector"] and x["child_ids"] == y["child_ids"] and x["metadata"] == y["metadata"]
        assert x["salience"] == y["salience"] and x["shard_key"] == y["shard_key"]
    # edges / profiles round trip (vector_store.py:159-233)
    for st in (ours, ref):
        st.add_edges([{"source": "n1", "target": "n2", "weight": 0.9}, {"source_id": "n2", "target_id": "n3", "type": "x"}], "alice")
        st.save_profile({"a": 1}, "alice")
    assert ours.get_edges(user_id="alice") == ref.get_edges(user_id="alice")
    assert ours.get_edges("n2", "alice")[0]["type"] == "x" and ours.load_profile("alice") == {"a": 1}
    assert ours.load_profile("bob") is None
    ours.delete_edges("n1", "alice")
    assert [e["source_id"] for e in ours.get_edges(user_id="alice")] == ["n2"]


def test_search_memories_golden_through_b200_store(golden):
    """memory_system.py:1460-1472 recorded from the reference over the oracle store: our store
    returns the same id lists for the same rows and queries."""
    import json
    from lazzaro_b200.store import B200Store
    g = golden
    st = B200Store()
    st.add_nodes([{"id": f"node_{i}", "content": f"memory {i}", "embedding": r.tolist()} for i, r in enumerate(g["sm_rows"])])
    missing = {f"node_{i}" for i in g["sm_missing"].tolist()}
    want = json.loads(str(g["sm_ids"]))
    for q, w in zip(g["sm_queries"], want):
        got = [i for i in st.search_nodes(q.tolist(), limit=10) if i not in missing]
        assert got == w


def test_limit_bound_is_loud():
    from lazzaro_b200.store import B200Store
    st = B200Store()
    st.add_rows(np.random.default_rng(0).standard_normal((3000, 16)).astype(np.float32))
    assert len(st.search_nodes([0.1] * 16, limit=1024)) == 1024
    with pytest.raises(ValueError):
        st.search_nodes([0.1] * 16, limit=2000)
    with pytest.raises(ValueError):
        st.search_nodes([0.1] * 17, limit=1)


def test_per_user_batched_search_matches_oracle():
    """BASELINE configs[4] shape in miniature: many tenants, one query each, one launch."""
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    from oracle import flat_search
    n_users, per, dim = 60, 50, 64
    rows, _ = synth.clustered_rows(n_users * per, dim, 12, seed=33)
    st = B200Store()
    for u in range(n_users):
        st.add_rows(rows[u * per:(u + 1) * per], user_id=f"u{u}", id_prefix=f"u{u}_", first_id=0)
    st.delete_nodes([f"u3_{i}" for i in range(0, per, 2)], user_id="u3")
    qs = synth.queries_near_rows(rows, 40, seed=34)
    users = [f"u{(7 * i) % n_users}" for i in range(38)] + ["u3", "ghost"]
    got = st.search_users(qs, users, limit=8)
    for b, name in enumerate(users):
        if name == "ghost":
            assert got[b] == []
            continue
        u = int(name[1:])
        mask = np.zeros(n_users * per, bool); mask[u * per:(u + 1) * per] = True
        if u == 3:
            mask[u * per:(u + 1) * per:2] = False
        idx, _ = flat_search.flat_topk(rows, qs[b], 8, mask=mask)
        assert got[b] == [f"u{u}_{i - u * per}" for i in idx]
        assert got[b] == st.search_nodes(qs[b].tolist(), user_id=name, limit=8)


def test_two_level_search_parity_and_recall():
    """north_star (1): centroid routing then member scan. nprobe = all shards is the flat search."""
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    from oracle import flat_search
    n, dim, S = 40000, 128, 32
    rows, shard = synth.clustered_rows(n, dim, S, seed=44)
    st = B200Store()
    st.add_rows(rows, id_prefix="node_")
    off = np.searchsorted(shard, np.arange(S + 1))
    st.set_shards(off)
    np.testing.assert_allclose(st._centroids[3].cpu().numpy(), rows[off[3]:off[4]].astype(np.float64).mean(0), rtol=1e-6)
    qs = synth.queries_near_rows(rows, 24, seed=45)
    want, _ = flat_search.flat_topk_batch(rows, qs, 10)
    want = [[f"node_{i}" for i in w] for w in want.tolist()]
    assert st.search_two_level(qs, limit=10) == want                       # every shard visited: exact
    assert st.search_two_level(qs, limit=10, nprobe=S) == want
    approx = st.search_two_level(qs, limit=10, nprobe=2)
    recall = np.mean([len(set(a) & set(w)) / 10 for a, w in zip(approx, want)])
    assert recall > 0.95                                                     # neighbours live in the query's shard


def test_two_level_policy_and_recall_report():
    """the approximate mode as a store policy, reported with recall (never as parity)."""
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    n, dim, S = 30000, 64, 24
    rows, shard = synth.clustered_rows(n, dim, S, seed=51)
    st = B200Store()
    st.add_rows(rows, id_prefix="node_")
    qs = synth.queries_near_rows(rows, 16, seed=52)
    exact = [st.search_nodes(q.tolist(), limit=10) for q in qs]
    st.search_policy = "two_level"
    with pytest.raises(RuntimeError):
        st.search_nodes(qs[0].tolist(), limit=10)                  # no shard layout declared yet
    st.set_shards(np.searchsorted(shard, np.arange(S + 1)))
    assert [st.search_nodes(q.tolist(), limit=10) for q in qs] == exact          # nprobe None = every shard = exact
    st.nprobe = 2
    approx = [st.search_nodes(q.tolist(), limit=10) for q in qs]
    assert np.mean([len(set(a) & set(b)) / 10 for a, b in zip(approx, exact)]) > 0.9
    rep = st.recall_report(qs, limit=10, nprobes=(1, 4, S))
    assert [r["nprobe"] for r in rep] == [1, 4, S] and rep[-1]["recall_at_10"] == 1.0
    assert rep[0]["rows_scanned_fraction_upper"] < rep[1]["rows_scanned_fraction_upper"] <= 1.0
    st.search_policy = "exact"
    assert [st.search_nodes(q.tolist(), limit=10) for q in qs] == exact


def test_pipelined_host_search_equals_blocking():
    """ShardedSearch.submit_host keeps one search in flight (staging + H2D of the next batch overlap it): same
    ids, scores and counts as search_host and as the oracle, batch after batch, both staging slots."""
    from lazzaro_b200 import synth
    from lazzaro_b200.dist import ShardedSearch
    from lazzaro_b200.store import B200Store
    from oracle import flat_search
    rows, _ = synth.clustered_rows(20000, 72, 8, seed=91)         # dim 72: padded to 72? exercises the staging pad path if any
    st = B200Store()
    st.add_rows(rows, id_prefix="node_")
    ss = ShardedSearch(st, world=1)
    batches = [synth.queries_near_rows(rows, 16, seed=92 + i) for i in range(5)]
    pend, got = None, []
    for i, q in enumerate(batches):
        h = ss.submit_host(q, 10, slot=i & 1)
        if pend is not None:
            got.append(pend.result())
        pend = h
    got.append(pend.result())
    for q, (ids, sc, cnt) in zip(batches, got):
        b_ids, b_sc, b_cnt = ss.search_host(q, 10)
        np.testing.assert_array_equal(ids, b_ids); np.testing.assert_array_equal(sc, b_sc); np.testing.assert_array_equal(cnt, b_cnt)
        want, _ = flat_search.flat_topk_batch(rows, q, 10)
        np.testing.assert_array_equal(ids, want)


def test_graph_replay_survives_interleaved_shapes():
    """Cached CUDA graphs replay raw addresses: each graph owns its scratch, so a later search that needs a larger
    scratch (other batch size / limit) or any allocation in between must not disturb an earlier graph's replay."""
    import torch
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    rows, _ = synth.clustered_rows(6000, 64, 6, seed=31)
    qs = synth.queries_near_rows(rows, 8, seed=32)
    st, ref = B200Store(capacity=8192), FlatStore()
    for s in (st, ref):
        s.add_nodes([{"id": f"n{i}", "content": "c", "embedding": rows[i].tolist()} for i in range(3000)], user_id="a")
        s.add_nodes([{"id": f"m{i}", "content": "c", "embedding": rows[3000 + i].tolist()} for i in range(3000)], user_id="b")
    st.batched_search = "never"                     # every shape below takes the graph-replayed exact path
    want5 = [ref.search_nodes(q.tolist(), "a", 5) for q in qs]
    assert st.search_nodes(qs[0].tolist(), "a", 5) == want5[0]                    # captures graph (B=1, k=5)
    hog = []
    for round_ in range(3):
        assert st.search_batch(qs[:4], "b", 100) == [ref.search_nodes(q.tolist(), "b", 100) for q in qs[:4]]   # larger scratch
        assert st.search_nodes(qs[1].tolist(), "b", 50) == ref.search_nodes(qs[1].tolist(), "b", 50)
        # allocations that would land in a freed scratch block if a graph's scratch had been released
        hog.append(torch.full((1 << 18,), 7.0, device=st.device))
        guard = torch.full((1 << 16,), 3.0, device=st.device)
        for i, q in enumerate(qs):
            assert st.search_nodes(q.tolist(), "a", 5) == want5[i]                # replays the first graph
        assert float(guard.min()) == 3.0 and float(guard.max()) == 3.0
        assert all(float(h.min()) == 7.0 and float(h.max()) == 7.0 for h in hog)
    n_graphs = len(st._graphs)
    st.add_nodes([{"id": f"x{i}", "content": "c", "embedding": rows[i].tolist()} for i in range(4000)], user_id="c")   # grows the columns
    assert len(st._graphs) < n_graphs or st._rows.shape[0] == 8192
    assert st.search_nodes(qs[0].tolist(), "a", 5) == want5[0]


def test_tenant_search_routes_and_filters_by_user():
    from oracle import flat_search
    """configs[4]: equal-size tenants adopted in bulk; a routed batch returns, per query, the exact top-k of that
    user's rows only (the reference's `.where(user_id = ...)` pre-filter, vector_store.py:136-139)."""
    import torch
    from lazzaro_b200 import synth
    from lazzaro_b200.dist import TenantSearch
    from lazzaro_b200.store import B200Store
    n_users, per, dim = 40, 250, 64
    rows_np, _ = synth.clustered_rows(n_users * per, dim, 8, seed=61)
    world = 2
    stores = []
    for r in range(world):
        st = B200Store()
        st.adopt_tenants(torch.as_tensor(rows_np[r * 20 * per:(r + 1) * 20 * per]).cuda().contiguous(), per, first_user=r * 20)
        stores.append(TenantSearch(st, 20, r, world))
    users = np.random.default_rng(1).integers(0, n_users, 96)
    qs = synth.queries_near_rows(rows_np, 96, seed=62)
    got = {}
    for ts in stores:
        pos, ids, cnt = ts.search_host(qs, users, 7)
        for p, row, c in zip(pos.tolist(), ids, cnt):
            assert c == 7
            got[p] = row + ts.rank * 20 * per                       # local row -> global row
    assert sorted(got) == list(range(96))
    for i in range(96):
        u = users[i]
        mask = np.zeros(n_users * per, bool); mask[u * per:(u + 1) * per] = True
        want, _ = flat_search.flat_topk(rows_np, qs[i], 7, mask=mask)
        np.testing.assert_array_equal(got[i], want)
    assert stores[1].store.search_nodes(qs[0].tolist(), user_id="user_25", limit=3)[0].startswith("user_25/")
    # the pipelined host form: three batches in flight over three staging slots give the answers of search_host
    batches = [(synth.queries_near_rows(rows_np, 64, seed=70 + b), np.random.default_rng(b).integers(0, n_users, 64)) for b in range(5)]
    ts = stores[0]
    want_b = [ts.search_host(q, u, 5) for q, u in batches]
    pend = [ts.submit_host(q, u, 5, slot=b % 3) for b, (q, u) in enumerate(batches[:3])]
    got_b = [h.result() for h in pend]
    pend = [ts.submit_host(q, u, 5, slot=b % 3) for b, (q, u) in enumerate(batches[3:])]
    got_b += [h.result() for h in pend]
    for w, g in zip(want_b, got_b):
        for a, b_ in zip(w, g):
            np.testing.assert_array_equal(a, b_)
    none_here = ts.submit_host(batches[0][0], np.full(64, n_users - 1), 5).result()      # every query belongs to the other rank
    assert none_here[0].size == 0 and none_here[1].shape == (0, 5)


def test_single_query_fused_path_equals_the_two_launch_path():
    """lz_search_one (one launch, query and answer in pinned host memory, host polls a flag) against the regular exact
    path and the oracle: several users, limits, deletions in between, both metrics, limit above the live row count."""
    from oracle import flat_search
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    rows, _ = synth.clustered_rows(9000, 96, 6, seed=71)
    qs = synth.queries_near_rows(rows, 10, seed=72)
    for metric in ("cosine", "l2"):
        st, ref = B200Store(metric=metric), FlatStore(metric=metric)
        for s in (st, ref):
            s.add_nodes([{"id": f"a{i}", "content": "", "embedding": rows[i].tolist()} for i in range(6000)], user_id="a")
            s.add_nodes([{"id": f"b{i}", "content": "", "embedding": rows[6000 + i].tolist()} for i in range(3000)], user_id="b")
        assert st.use_fused_one
        for rounds in range(2):
            for u in ("a", "b", "nobody"):
                for q in qs:
                    for limit in (1, 10, 64):
                        got = st.search_nodes(q.tolist(), user_id=u, limit=limit)
                        assert got == ref.search_nodes(q.tolist(), user_id=u, limit=limit)
            st.use_fused_one = False
            slow = [st.search_nodes(q.tolist(), user_id="a", limit=10) for q in qs]
            st.use_fused_one = True
            assert slow == [st.search_nodes(q.tolist(), user_id="a", limit=10) for q in qs]
            for s in (st, ref):
                s.delete_nodes([f"a{i}" for i in range(0, 6000, 3)], user_id="a")
        few = B200Store(metric=metric)
        few.add_nodes([{"id": f"n{i}", "content": "", "embedding": rows[i].tolist()} for i in range(3)])
        assert few.search_nodes(qs[0].tolist(), limit=10) == [f"n{i}" for i in flat_search.flat_topk(rows[:3], qs[0], 3, metric)[0]]


def test_small_batches_keep_the_exact_scan_when_the_shadow_does_not_fit(monkeypatch):
    """A span large enough for single queries to take the fp16 shadow, and a shadow allocation that fails: the search
    still answers (exact scan on the GPU, same ids as with the shadow) and stops asking for the shadow."""
    import torch
    from oracle import flat_search
    from lazzaro_b200 import store as store_mod, synth
    from lazzaro_b200.store import B200Store
    n, dim = 520_000, 32
    rows_np, _ = synth.clustered_rows(n, dim, 16, seed=77)
    qs = synth.queries_near_rows(rows_np, 3, seed=78)
    a = B200Store(); a.adopt_rows(torch.as_tensor(rows_np).cuda(), id_prefix="n")
    with_shadow = [a.search_nodes(q.tolist(), limit=5) for q in qs]
    assert a._shadows and not a._shadow_oom                                  # the shadow path was taken
    b = B200Store(); b.adopt_rows(torch.as_tensor(rows_np).cuda(), id_prefix="n")

    def no_room(self, st):
        raise torch.cuda.OutOfMemoryError("no room for the shadow (simulated)")
    monkeypatch.setattr(store_mod._Shadow, "ensure", no_room)
    without = [b.search_nodes(q.tolist(), limit=5) for q in qs]
    assert b._shadow_oom and not b._shadows
    assert without == with_shadow
    want, _ = flat_search.flat_topk(rows_np, qs[0], 5)
    assert without[0] == [f"n{r}" for r in want.tolist()]
    with pytest.raises(torch.cuda.OutOfMemoryError):                          # a real batch needs the shadow: the error stands
        b.search_batch(np.repeat(qs, 4, axis=0), limit=5)


def test_single_queries_on_a_large_span_follow_adds_and_deletes():
    """Single queries over 520k rows take the fp16 shadow; rows added and deleted afterwards must show up / disappear exactly
    as with the exact scan (the shadow is extended for appended rows, tombstones are filtered in the exact pass)."""
    import torch
    from oracle import flat_search
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    n, dim = 520_000, 32
    rows_np, _ = synth.clustered_rows(n, dim, 16, seed=91)
    qs = synth.queries_near_rows(rows_np, 4, seed=92)
    st = B200Store(); st.adopt_rows(torch.as_tensor(rows_np).cuda(), id_prefix="n")
    ex = B200Store(); ex.adopt_rows(torch.as_tensor(rows_np).cuda(), id_prefix="n")
    ex.shadow_for_single_queries = False                                    # the same operations through the exact scan
    user = st._user_names[0]

    def same():
        for q in qs:
            a, b = st.search_nodes(q.tolist(), user, 8), ex.search_nodes(q.tolist(), user, 8)
            assert a == b and len(a) == 8
        return st.search_nodes(qs[0].tolist(), user, 8)

    first = same()
    assert st._shadows and not ex._shadows
    want, _ = flat_search.flat_topk(rows_np, qs[0], 8)
    assert first == [f"n{r}" for r in want.tolist()]
    near = [{"id": f"new{i}", "content": "", "embedding": (qs[0] * (1.0 + 1e-3 * i)).tolist()} for i in range(5)]   # collinear with query 0
    for s_ in (st, ex):
        s_.add_nodes(near, user_id=user)
    after_add = same()
    assert set(after_add[:5]) == {f"new{i}" for i in range(5)}
    for s_ in (st, ex):
        s_.delete_nodes(["new0", "new3", first[0]], user_id=user)
    after_del = same()
    assert "new0" not in after_del and "new3" not in after_del and first[0] not in after_del
    assert {"new1", "new2", "new4"} <= set(after_del)
