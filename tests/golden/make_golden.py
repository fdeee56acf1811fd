"""Regenerates tests/golden/*.npz by running the REFERENCE's own Python (imported from
/root/reference/src under a stub `lancedb`, see oracle/ref_harness.py) on seeded inputs.

Run in the build container only:   python tests/golden/make_golden.py
The fixtures travel with the repo; /root/reference does not exist on the GPU box.

What is recorded (inputs AND reference outputs, so tests need nothing but numpy):
  cosine        memory_system.py:197-203 on random / zero / empty pairs
  decay_prune   memory_shard.py:64-84 on Edge/Node objects seeded with float32 state
  linking       memory_system.py:797-891 (_link_within_shards + _link_to_existing_memories)
  dedup         memory_system.py:651-785 (_async_consolidate) merge decisions
  centroid      memory_system.py:893-933 (_create_super_nodes_for_shard)
  search_mem    memory_system.py:1460-1472 over an injected flat store, cfg1-like (2k x 64)
  pins          the expectations of the reference's own tests, reproduced by running its code
  msys          memory_system.py:580-785 end to end: three conversations (dedup, linking, super-nodes, decay, prune)
and, in ref_next_rows.npz:
  eviction      memory_system.py:535-578 (_enforce_buffer_limit) under a frozen clock
  neighbors     memory_shard.py:54-62 (get_neighbors) on a random edge set
  reload        memory_system.py:1304-1410 (_load_from_persistence) of the msys scenario's saved state
"""
from __future__ import annotations

import io
import json
import os
import sys
from contextlib import redirect_stdout

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)

from oracle import ref_harness  # noqa: E402
from oracle.flat_store import FlatStore  # noqa: E402
from lazzaro_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def quiet(fn, *a, **k):
    with redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def gold_cosine(ms):
    rng = np.random.default_rng(11)
    a = rng.standard_normal((64, 48)).astype(np.float32)
    b = rng.standard_normal((64, 48)).astype(np.float32)
    b[5] = 0.0                                          # zero-norm guard, memory_system.py:203
    b[6] = a[6] * 3.0                                   # collinear
    sims = np.array([ms._cosine_similarity(x.tolist(), y.tolist()) for x, y in zip(a, b)])
    empty = ms._cosine_similarity([], a[0].tolist())    # memory_system.py:199
    return dict(cos_a=a, cos_b=b, cos_out=sims, cos_empty=np.float64(empty))


def gold_decay_prune(lz):
    from lazzaro.core.memory_shard import MemoryShard
    from lazzaro.models.graph import Edge, Node
    n_nodes, n_edges = 300, 5000
    src, dst, w, cooc = synth.edge_table(n_edges, n_nodes, seed=synth.SEED_EDGES)
    # unique (src,dst) keys: the shard's dict is keyed by the pair (memory_shard.py:47)
    _, first = np.unique(src.astype(np.int64) * n_nodes + dst, return_index=True)
    keep = np.sort(first)
    src, dst, w, cooc = src[keep], dst[keep], w[keep], cooc[keep]
    w[:8] = np.float32([0.5, 0.50505048, 0.5050505, 0.5050506, 0.0, 1.0, 0.49999997, 0.50000006])
    rng = np.random.default_rng(5)
    sal = rng.random(n_nodes, dtype=np.float32)
    sal[:4] = np.float32([0.2, 0.19, 0.20000002, 1.0])
    shard = MemoryShard("s")
    for i in range(n_nodes):
        shard.add_node(Node(id=f"n{i}", content="c", salience=float(sal[i])))
    for s, d, ww, c in zip(src, dst, w, cooc):
        shard.edges[(f"n{s}", f"n{d}")] = Edge(source=f"n{s}", target=f"n{d}", weight=float(ww),
                                              co_occurrence=int(c))
    shard.apply_temporal_decay(0.01)                                   # memory_shard.py:64-77
    w_after = np.array([e.weight for e in shard.edges.values()], dtype=np.float64)
    sal_after = np.array([n.salience for n in shard.nodes.values()], dtype=np.float64)
    pruned = shard.prune_weak_edges(0.5)                               # memory_shard.py:79-84
    kept = np.array([[int(k[0][1:]), int(k[1][1:])] for k in shard.edges.keys()], dtype=np.int32)
    kept_w = np.array([e.weight for e in shard.edges.values()], dtype=np.float64)
    return dict(dp_src=src, dp_dst=dst, dp_w=w, dp_cooc=cooc, dp_sal=sal,
                dp_w_after=w_after, dp_sal_after=sal_after, dp_pruned=np.int64(pruned),
                dp_kept=kept, dp_kept_w=kept_w)


def gold_linking(lz):
    """40 new rows against 400 existing rows in 4 shards, through the reference's own linkers."""
    from lazzaro.models.graph import Node
    dim, n_exist, n_new, n_shards = 32, 400, 40, 4
    rows, shard = synth.clustered_rows(n_exist, dim, n_shards, seed=21, sigma=0.8)
    new, _ = synth.clustered_rows(n_new, dim, n_shards, seed=21, sigma=0.8)   # same centres
    new = synth._unit(new + 0.3 * np.random.default_rng(22).standard_normal(new.shape) / np.sqrt(dim)).astype(np.float32)
    new_shard = (np.arange(n_new) * n_shards // n_new).astype(np.int32)
    def build():
        ms = quiet(ref_harness.make_memory_system, FlatStore(), enable_hierarchy=False, max_buffer_size=10**9)
        for i in range(n_exist):
            ms.buffer.add_node(Node(id=f"e{i}", content=f"e{i}", embedding=rows[i].astype(np.float64).tolist(),
                                    shard_key=f"s{shard[i]}"))
        new_nodes = []
        for i in range(n_new):
            ms.buffer.add_node(Node(id=f"x{i}", content=f"x{i}", embedding=new[i].astype(np.float64).tolist(),
                                    shard_key=f"s{new_shard[i]}"))
            new_nodes.append((f"x{i}", f"s{new_shard[i]}"))
        return ms, new_nodes

    dump = lambda ms: [(k[0], k[1], e.weight, e.co_occurrence) for s in ms.shards.values() for k, e in s.edges.items()]
    ms, new_nodes = build()
    quiet(ms._link_within_shards, new_nodes)                            # memory_system.py:797-836
    within = dump(ms)
    quiet(ms._link_to_existing_memories, new_nodes)                     # memory_system.py:838-891
    both = dump(ms)
    ms, new_nodes = build()
    quiet(ms._link_to_existing_memories, new_nodes)                     # cross linker alone
    cross = dump(ms)
    enc = lambda L: json.dumps(L)
    return dict(ln_rows=rows, ln_shard=shard, ln_new=new, ln_new_shard=new_shard,
                ln_within=np.array(enc(within)), ln_both=np.array(enc(both)),
                ln_cross=np.array(enc(cross)))


def gold_dedup(lz):
    """_async_consolidate on 30 candidate memories, 12 of them near-duplicates (cos≈0.97) of graph
    rows; records which merged, the resulting salience/access_count, and the new node count."""
    dim, n_graph, n_new = 48, 600, 30
    rows, shard = synth.clustered_rows(n_graph, dim, 6, seed=31)
    dups, srcs = synth.near_duplicates(rows, 12, seed=32, noise=0.25)
    fresh = synth._unit(np.random.default_rng(33).standard_normal((n_new - 12, dim))).astype(np.float32)
    new = np.concatenate([dups, fresh])
    perm = np.random.default_rng(34).permutation(n_new)
    new, is_dup_src = new[perm], np.concatenate([srcs, -np.ones(n_new - 12, np.int64)])[perm]
    store = FlatStore(metric="cosine")
    emb = ref_harness.TableEmbedder({f"new fact number {i}": new[i].astype(np.float64).tolist() for i in range(n_new)}, dim)
    llm = ref_harness.ScriptedLLM()
    ms = quiet(ref_harness.make_memory_system, store, llm, emb, enable_hierarchy=False,
               max_buffer_size=10**9, auto_consolidate=False, auto_prune=False)
    from lazzaro.models.graph import Node
    nodes = []
    for i in range(n_graph):
        nd = Node(id=f"g{i}", content=f"graph memory {i}", embedding=rows[i].astype(np.float64).tolist(),
                  shard_key=f"s{shard[i]}", salience=0.5)
        ms.buffer.add_node(nd)
        nodes.append({"id": nd.id, "content": nd.content, "embedding": nd.embedding, "shard_key": nd.shard_key})
    store.add_nodes(nodes, user_id=ms.user_id)
    sal_new = np.round(np.random.default_rng(35).random(n_new), 3)
    llm.push_memories([{"content": f"new fact number {i}", "type": "semantic", "salience": float(sal_new[i]),
                        "topic": "other"} for i in range(n_new)])
    ms.consolidation_queue.append({"memories": [{"content": "hello"}], "timestamp": 0.0})
    quiet(ms._async_consolidate)                                        # memory_system.py:651-785
    merged = np.array([ms.buffer.get_node(f"g{i}").access_count for i in range(n_graph)], dtype=np.int64)
    sal_after = np.array([ms.buffer.get_node(f"g{i}").salience for i in range(n_graph)], dtype=np.float64)
    n_nodes_after = ms.buffer.size()[0]
    return dict(dd_rows=rows, dd_new=new, dd_sal_new=sal_new, dd_dup_src=is_dup_src,
                dd_access=merged, dd_sal_after=sal_after, dd_nodes_after=np.int64(n_nodes_after))


def gold_centroid(lz):
    from lazzaro.models.graph import Node
    dim, n = 40, 50
    rows, _ = synth.clustered_rows(n, dim, 1, seed=41)
    ms = quiet(ref_harness.make_memory_system, FlatStore(), super_node_threshold=20)
    for i in range(n):
        ms.buffer.add_node(Node(id=f"c{i}", content=f"c{i}", embedding=rows[i].astype(np.float64).tolist(), shard_key="s0"))
    quiet(ms._create_super_nodes_for_shard, "s0")                       # memory_system.py:893-933
    sn = next(iter(ms.super_nodes.values()))
    return dict(ce_rows=rows, ce_centroid=np.array(sn.embedding, dtype=np.float64),
                ce_children=np.array(json.dumps(sn.child_ids)))


def gold_search_memories(lz):
    """search_memories (memory_system.py:1460-1472) and _optimized_retrieval (453-514) over the
    injected flat store: 2000 x 64 rows, 16 queries, limit 10; 25 rows are not resident in the
    in-memory graph, so the returned lists may be shorter than `limit` (memory_system.py:1468-1471)."""
    from lazzaro.models.graph import Node
    dim, n, nq = 64, 2000, 16
    rows, shard = synth.clustered_rows(n, dim, 8, seed=51)
    qs = synth.queries_near_rows(rows, nq, seed=52)
    store = FlatStore(metric="cosine")
    emb = ref_harness.TableEmbedder({f"q{i}": qs[i].astype(np.float64).tolist() for i in range(nq)}, dim)
    ms = quiet(ref_harness.make_memory_system, store, None, emb, enable_hierarchy=False, max_buffer_size=10**9)
    missing = set(np.random.default_rng(53).choice(n, 25, replace=False).tolist())
    batch = []
    for i in range(n):
        nd = Node(id=f"node_{i}", content=f"memory {i}", embedding=rows[i].astype(np.float64).tolist(), shard_key=f"s{shard[i]}")
        if i not in missing:
            ms.buffer.add_node(nd)
        batch.append({"id": nd.id, "content": nd.content, "embedding": nd.embedding, "shard_key": nd.shard_key})
    store.add_nodes(batch, user_id=ms.user_id)
    sm = [[nd.id for nd in ms.search_memories(f"q{i}", limit=10)] for i in range(nq)]
    orr = [ms._optimized_retrieval(qs[i].astype(np.float64).tolist(), f"q{i}") for i in range(nq)]
    return dict(sm_rows=rows, sm_shard=shard, sm_queries=qs, sm_missing=np.array(sorted(missing), np.int64),
                sm_ids=np.array(json.dumps(sm)), sm_retrieval=np.array(json.dumps(orr)))


def gold_pins(lz):
    """Re-run the reference tests' scenarios through its own code with the injected store."""
    # tests/test_consolidation_semantic.py:27-71
    store = FlatStore(metric="cosine")
    llm = ref_harness.ScriptedLLM()
    emb = ref_harness.TableEmbedder(dim=1536)
    ms = quiet(ref_harness.make_memory_system, store, llm, emb)
    f1, f2 = "User prefers Python for data science", "The user has a preference for the Python programming language in data science tasks"
    emb.table[f1], emb.table[f2] = [0.1] * 1536, [0.1001] * 1536
    llm.push_memories([{"content": f1, "type": "semantic", "salience": 0.8, "topic": "work"}])
    quiet(ms.start_conversation); ms.add_to_short_term("I like Python for DS"); quiet(ms.end_conversation)
    nid = list(ms.buffer.nodes.keys())[0]
    llm.push_memories([{"content": f2, "type": "semantic", "salience": 0.9, "topic": "work"}])
    quiet(ms.start_conversation); ms.add_to_short_term("Python is my go-to for data work"); quiet(ms.end_conversation)
    node = ms.buffer.get_node(nid)
    # tests/test_profile_update.py:32-78 — chain of 4 nodes with edges 0.8: one component, avg > 0.3
    from lazzaro.models.graph import Edge, Node
    ms2 = quiet(ref_harness.make_memory_system, FlatStore(), auto_consolidate=False)
    for i in range(4):
        ms2.buffer.add_node(Node(id=f"seed_{i}", content=f"m{i}", embedding=[0.1] * 10, shard_key="default"))
    for i in range(3):
        ms2.buffer.add_edge(Edge(source=f"seed_{i}", target=f"seed_{i+1}", weight=0.8))
    comps = ms2.buffer.get_connected_components()
    pruned = ms2.buffer.prune_weak_edges(0.5)
    # _merge_similar_nodes is a no-op as written (memory_system.py:1073-1077)
    ms3 = quiet(ref_harness.make_memory_system, FlatStore())
    for i in range(4):
        ms3.buffer.add_node(Node(id=f"d{i}", content="same", embedding=[0.3] * 16, shard_key="default"))
    merged = quiet(ms3._merge_similar_nodes, 0.95)
    return dict(pin_nodes=np.int64(ms.buffer.size()[0]), pin_salience=np.float64(node.salience),
                pin_access=np.int64(node.access_count),
                pin_components=np.array(json.dumps(sorted(sorted(c) for c in comps))),
                pin_pruned=np.int64(pruned), pin_merge_noop=np.int64(merged))


def _run_msys_scenario():
    """Three conversations through the reference's own MemorySystem with scripted facts and table embeddings."""
    dim, per = 32, 9
    topics = ["work", "health", "learning"]
    rows, shard = synth.clustered_rows(3 * 3 * per, dim, 3, seed=61, sigma=0.6)
    order = np.random.default_rng(62).permutation(rows.shape[0])
    facts = [(f"fact number {i} about {topics[shard[i]]}", topics[shard[i]], rows[i]) for i in order]
    dups, src = synth.near_duplicates(rows[order[:per * 3]], 5, seed=63, noise=0.15)       # re-told facts, conv 3
    table = {c: v.astype(np.float64).tolist() for c, _, v in facts}
    for j, v in enumerate(dups):
        table[f"retold fact {j}"] = v.astype(np.float64).tolist()
    queries = synth.queries_near_rows(rows, 6, seed=64)
    for j, v in enumerate(queries):
        table[f"query {j}"] = v.astype(np.float64).tolist()
    store, llm, emb = FlatStore(metric="cosine"), ref_harness.ScriptedLLM(), ref_harness.TableEmbedder(table, dim)
    kw = dict(max_buffer_size=10**6, super_node_threshold=12, auto_consolidate=True, consolidate_every=3, prune_threshold=0.45)
    ms = quiet(ref_harness.make_memory_system, store, llm, emb, **kw)
    convs = [facts[:per * 3], facts[per * 3:per * 6], facts[per * 6:]]
    script, states = [], []
    for ci, conv in enumerate(convs):
        mems = [{"content": c, "type": "semantic", "salience": round(0.3 + 0.05 * (k % 9), 2), "topic": t}
                for k, (c, t, _) in enumerate(conv)]
        if ci == 2:
            mems += [{"content": f"retold fact {j}", "type": "semantic", "salience": 0.95, "topic": facts[int(src[j])][1]}
                     for j in range(len(dups))]
        script.append(mems)
        llm.push_memories(mems)
        quiet(ms.start_conversation); ms.add_to_short_term(f"conversation {ci}"); quiet(ms.end_conversation)
        states.append({
            "nodes": {nid: [n.content, n.salience, n.access_count, n.shard_key, n.parent_id]
                      for sh in ms.shards.values() for nid, n in sh.nodes.items()},
            "edges": {f"{a}|{b}": [e.weight, e.co_occurrence] for sh in ms.shards.values() for (a, b), e in sh.edges.items()},
            "super": {sid: [n.shard_key, n.child_ids, n.embedding] for sid, n in ms.super_nodes.items()},
        })
    return dict(ms=ms, store=store, llm=llm, emb=emb, kw=kw, table=table, script=script, states=states,
                queries=queries, dim=dim)


def gold_memory_system(lz):
    """Three conversations through the reference's own MemorySystem (end_conversation -> consolidation,
    dedup, linking, super-nodes, decay, prune) with scripted facts and table embeddings; the state after
    each conversation and a few retrievals are recorded for the device mirror to reproduce."""
    r = _run_msys_scenario()
    ms, table, queries = r["ms"], r["table"], r["queries"]
    sm = [[n.id for n in ms.search_memories(f"query {j}", limit=5)] for j in range(len(queries))]
    orr = [ms._optimized_retrieval(table[f"query {j}"], f"query {j}") for j in range(len(queries))]
    comps = sorted(sorted(c) for c in ms.buffer.get_connected_components() if len(c) >= 3)
    return dict(msys=np.array(json.dumps({"table": table, "script": r["script"], "states": r["states"], "search": sm,
                                          "retrieval": orr, "components": comps, "dim": r["dim"]})))


def gold_reload(lz):
    """_load_from_persistence (memory_system.py:1304-1410) of the state the scenario above saved: a second reference
    MemorySystem on the same store with load_from_disk=True; what it restored is recorded."""
    r = _run_msys_scenario()
    ms2 = quiet(ref_harness.make_memory_system, r["store"], r["llm"], r["emb"], load_from_disk=True, **r["kw"])
    state = {
        "shards": {k: list(sh.nodes) for k, sh in ms2.shards.items()},
        "nodes": {nid: [n.content, n.salience, n.access_count, n.shard_key, n.parent_id, n.is_super_node, n.child_ids]
                  for nid, n in ms2.buffer.nodes.items()},
        "super": list(ms2.super_nodes),
        "edges": {k: [[a, b, e.weight, e.co_occurrence] for (a, b), e in sh.edges.items()] for k, sh in ms2.shards.items()},
        "node_counter": ms2.node_counter,
        "search": [[n.id for n in ms2.search_memories(f"query {j}", limit=5)] for j in range(len(r["queries"]))],
    }
    return dict(reload=np.array(json.dumps(state)))


def gold_eviction(lz):
    """_enforce_buffer_limit (memory_system.py:535-578) with the clock frozen: three shards, super-nodes mixed
    in, ties in importance (equal fields) to pin the stable order, edges touching evicted nodes."""
    import types
    from lazzaro.core import memory_system as ref_ms
    from lazzaro.models.graph import Edge, Node
    NOW = 1_750_000_000.25
    rng = np.random.default_rng(71)
    out = {}
    for case, (n, limit) in enumerate([(64, 40), (300, 120), (12, 11)]):
        store = FlatStore()
        ms = quiet(ref_harness.make_memory_system, store, max_buffer_size=limit)
        sal = np.round(rng.uniform(0.05, 1.0, n), 2)
        acc = rng.integers(0, 15, n)
        last = NOW - rng.choice([0.0, 3600.0, 86400.0, 5 * 86400.0, 40 * 86400.0, 400 * 86400.0], n) - rng.uniform(0, 50, n).round(1)
        sup = rng.random(n) < 0.08
        sal[5] = sal[3]; acc[5] = acc[3]; last[5] = last[3]; sup[3] = sup[5] = False      # exact ties
        if n > 40:
            sal[33] = sal[3]; acc[33] = acc[3]; last[33] = last[3]; sup[33] = False
        shard_of = rng.integers(0, 3, n)
        for i in range(n):
            ms.buffer.add_node(Node(id=f"n{i}", content=f"c{i}", embedding=[0.0] * 4, salience=float(sal[i]),
                                    access_count=int(acc[i]), last_accessed=float(last[i]), is_super_node=bool(sup[i]),
                                    shard_key=f"s{shard_of[i]}"))
        for _ in range(2 * n):
            a, b = rng.integers(0, n, 2)
            if a != b:
                ms.buffer.add_edge(Edge(source=f"n{a}", target=f"n{b}", weight=0.7))
        order = [nid for sh in ms.shards.values() for nid in sh.nodes]          # the reference's iteration order
        store.add_nodes([{"id": nid, "content": nid, "embedding": [0.0] * 4} for nid in order])
        deleted = []
        store.delete_nodes = lambda ids, user_id="default", _d=deleted: _d.extend(ids)
        real_time = ref_ms.time
        ref_ms.time = types.SimpleNamespace(time=lambda: NOW)
        try:
            quiet(ms._enforce_buffer_limit)
        finally:
            ref_ms.time = real_time
        idx = [int(nid[1:]) for nid in order]
        out.update({f"ev{case}_order": np.array(idx, np.int64), f"ev{case}_sal": sal[idx], f"ev{case}_acc": acc[idx].astype(np.int32),
                    f"ev{case}_last": last[idx], f"ev{case}_super": sup[idx], f"ev{case}_limit": np.int64(limit),
                    f"ev{case}_deleted": np.array([int(x[1:]) for x in deleted], np.int64),
                    f"ev{case}_left_nodes": np.array(sorted(int(nid[1:]) for sh in ms.shards.values() for nid in sh.nodes), np.int64),
                    f"ev{case}_left_edges": np.array(json.dumps(sorted([a, b] for sh in ms.shards.values() for (a, b) in sh.edges)))})
    out["ev_now"] = np.float64(NOW)
    return out


def gold_neighbors(lz):
    """MemoryShard.get_neighbors (memory_shard.py:54-62): neighbour ids in edge-dict order, weight >= min_weight."""
    from lazzaro.core.memory_shard import MemoryShard
    from lazzaro.models.graph import Edge, Node
    rng = np.random.default_rng(72)
    n, ne = 200, 1500
    sh = MemoryShard("s")
    for i in range(n):
        sh.add_node(Node(id=f"n{i}", content="", embedding=[]))
    src, dst = rng.integers(0, n, ne), rng.integers(0, n, ne)
    w = rng.uniform(0, 1, ne).astype(np.float32)
    w[::7] = np.float32(0.3)                                                     # exactly at the default gate
    for a, b, x in zip(src, dst, w):
        sh.add_edge(Edge(source=f"n{a}", target=f"n{b}", weight=float(x)))     # repeats reinforce (memory_shard.py:42-52)
    keys = list(sh.edges.keys())
    e_src = np.array([int(a[1:]) for a, _ in keys], np.int32)
    e_dst = np.array([int(b[1:]) for _, b in keys], np.int32)
    e_w = np.array([sh.edges[k].weight for k in keys], np.float64)
    res = {}
    for mw in (0.3, 0.5, 0.0):
        res[str(mw)] = [[int(x[1:]) for x in sh.get_neighbors(f"n{i}", mw)] for i in range(n)]
    return dict(nb_src=e_src, nb_dst=e_dst, nb_w=e_w, nb_n=np.int64(n), nb_lists=np.array(json.dumps(res)))


def main():
    lz = ref_harness.import_reference()
    ms = quiet(ref_harness.make_memory_system, FlatStore())
    out = {}
    for fn, arg in [(gold_cosine, ms), (gold_decay_prune, lz), (gold_linking, lz), (gold_dedup, lz),
                    (gold_centroid, lz), (gold_search_memories, lz), (gold_pins, lz), (gold_memory_system, lz)]:
        out.update(fn(arg))
        print("ok", fn.__name__)
    path = os.path.join(OUT, "ref_python_path.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")
    nxt = {}
    for fn in (gold_eviction, gold_neighbors, gold_reload):
        nxt.update(fn(lz))
        print("ok", fn.__name__)
    path = os.path.join(OUT, "ref_next_rows.npz")
    np.savez_compressed(path, **nxt)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
