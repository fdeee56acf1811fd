"""Round-2 fixtures (tests/golden/ref_r2.npz), recorded by running the REFERENCE's own Python (imported from
/root/reference/src under a stub `lancedb`, oracle/ref_harness.py) in the build container:

    python tests/golden/make_golden_r2.py

  comps      BufferGraph.get_connected_components (buffer_graph.py:99-120) on graphs with cross-shard edges in
             both directions, edges to ids that are not resident, super-nodes and several visiting orders — the
             cases in which the reference's components are NOT plain undirected connectivity
  xsuper     _async_consolidate -> _link_to_existing_memories (memory_system.py:838-891) when the store already
             holds a super-node's centroid row that is among the new memories' nearest rows
  state_json the reference's save_state file for the three-conversation scenario and what its load_state restores
  msys_l2    the three-conversation scenario of make_golden.py over an L2-ranked store with un-normalised
             embeddings (LanceDB's default metric; cosine is still what the consolidation code gates on)
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from oracle import ref_harness  # noqa: E402
from oracle.flat_store import FlatStore  # noqa: E402
from lazzaro_b200 import synth  # noqa: E402
from make_golden import quiet  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def gold_components(lz):
    from lazzaro.models.graph import Edge, Node
    cases = []
    rng = np.random.default_rng(81)
    for case in range(6):
        n, n_sh = [12, 40, 40, 90, 90, 25][case], [2, 3, 4, 5, 3, 1][case]
        shard_of = rng.integers(0, n_sh, n)
        order = rng.permutation(n)                                 # insertion order = the reference's visiting order
        ms = quiet(ref_harness.make_memory_system, FlatStore(), max_buffer_size=10 ** 9)
        supers = [f"super_{j}" for j in range(2 if case in (1, 3) else 0)]
        for sid in supers:
            ms.super_nodes[sid] = Node(id=sid, content=sid, embedding=[], is_super_node=True, shard_key="s0")
        nodes = []
        for i in order.tolist():
            ms.buffer.add_node(Node(id=f"n{i}", content="", embedding=[], shard_key=f"s{shard_of[i]}"))
            nodes.append([f"n{i}", f"s{shard_of[i]}"])
        edges = []
        n_edges = int(n * [0.5, 0.7, 0.9, 0.6, 1.2, 0.8][case])
        for _ in range(n_edges):
            a, b = rng.integers(0, n, 2)
            if a == b:
                continue
            src, dst = f"n{a}", f"n{b}"
            r = rng.random()
            if r < 0.08:
                dst = f"ghost{rng.integers(0, 4)}"                  # an id that is not resident anywhere
            elif r < 0.12 and supers:
                dst = supers[int(rng.integers(0, len(supers)))]     # an edge into a super-node
            w = float(np.float32(rng.uniform(0.0, 1.0)))
            ms.buffer.add_edge(Edge(source=src, target=dst, weight=w))     # stored in the SOURCE's shard (buffer_graph.py:51-61)
            edges.append([src, dst, w])
        comps = [sorted(c) for c in ms.buffer.get_connected_components()]
        cases.append({"supers": supers, "nodes": nodes, "edges": edges, "components": comps})
    return dict(comps=np.array(json.dumps(cases)))


def gold_cross_super(lz):
    """30 'work' memories -> a super-node (threshold 20) saved to the store; then a conversation with 6 new facts
    that sit close to that shard's centroid.  The centroid row is among their nearest store rows; the reference
    links each new fact to its 3 nearest NON-super resident nodes regardless."""
    dim = 24
    rows, _ = synth.clustered_rows(30, dim, 1, seed=91, sigma=0.9)
    centre = rows.astype(np.float64).mean(axis=0)
    rng = np.random.default_rng(92)
    new = synth._unit(centre / np.linalg.norm(centre) + 0.25 * rng.standard_normal((6, dim)) / np.sqrt(dim)).astype(np.float32)
    table = {f"work fact {i}": rows[i].astype(np.float64).tolist() for i in range(30)}
    table.update({f"late work fact {j}": new[j].astype(np.float64).tolist() for j in range(6)})
    store, llm, emb = FlatStore(metric="cosine"), ref_harness.ScriptedLLM(), ref_harness.TableEmbedder(table, dim)
    kw = dict(max_buffer_size=10 ** 6, super_node_threshold=20, auto_consolidate=False, prune_threshold=0.3)
    ms = quiet(ref_harness.make_memory_system, store, llm, emb, **kw)
    script = [[{"content": f"work fact {i}", "type": "semantic", "salience": 0.5, "topic": "work"} for i in range(30)],
              [{"content": f"late work fact {j}", "type": "semantic", "salience": 0.6, "topic": "other"} for j in range(6)]]
    states = []
    for ci, mems in enumerate(script):
        llm.push_memories(mems)
        quiet(ms.start_conversation); ms.add_to_short_term(f"conversation {ci}"); quiet(ms.end_conversation)
        states.append({"edges": {f"{a}|{b}": [e.weight, e.co_occurrence] for sh in ms.shards.values() for (a, b), e in sh.edges.items()},
                       "super": sorted(n.shard_key for n in ms.super_nodes.values()),
                       "nodes": [nid for sh in ms.shards.values() for nid in sh.nodes]})
    assert states[0]["super"] == ["work"]
    # the point of the fixture: the centroid row really is among the new facts' 3 nearest store rows
    cen = np.asarray(next(iter(ms.super_nodes.values())).embedding)
    closer = [(rows.astype(np.float64) @ v / np.linalg.norm(rows, axis=1) > cen @ v / np.linalg.norm(cen)).sum() for v in new.astype(np.float64)]
    assert min(closer) < 3, closer
    return dict(xsuper=np.array(json.dumps({"table": table, "script": script, "states": states, "dim": dim, "kw": kw})))


def gold_msys_l2(lz):
    """make_golden._run_msys_scenario with un-normalised embeddings and an L2-ranked store."""
    dim, per = 32, 9
    topics = ["work", "health", "learning"]
    rows, shard = synth.clustered_rows(3 * 3 * per, dim, 3, seed=61, sigma=0.6)
    scale = np.random.default_rng(65).uniform(0.5, 2.0, rows.shape[0]).astype(np.float32)
    rows = rows * scale[:, None]                                     # L2 and cosine rankings now differ
    order = np.random.default_rng(62).permutation(rows.shape[0])
    facts = [(f"fact number {i} about {topics[shard[i]]}", topics[shard[i]], rows[i]) for i in order]
    dups, src = synth.near_duplicates(rows[order[:per * 3]], 5, seed=63, noise=0.15)
    dups = dups * scale[order[:per * 3]][src][:, None]
    table = {c: v.astype(np.float64).tolist() for c, _, v in facts}
    for j, v in enumerate(dups):
        table[f"retold fact {j}"] = v.astype(np.float64).tolist()
    queries = synth.queries_near_rows(rows, 6, seed=64) * np.float32(1.3)
    for j, v in enumerate(queries):
        table[f"query {j}"] = v.astype(np.float64).tolist()
    store, llm, emb = FlatStore(metric="l2"), ref_harness.ScriptedLLM(), ref_harness.TableEmbedder(table, dim)
    kw = dict(max_buffer_size=10 ** 6, super_node_threshold=12, auto_consolidate=True, consolidate_every=3, prune_threshold=0.45)
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
    sm = [[n.id for n in ms.search_memories(f"query {j}", limit=5)] for j in range(len(queries))]
    orr = [ms._optimized_retrieval(table[f"query {j}"], f"query {j}") for j in range(len(queries))]
    comps = [sorted(c) for c in ms.buffer.get_connected_components()]
    # the metric must matter for the fixture to mean anything
    cos_store = FlatStore(metric="cosine")
    cos_store.add_nodes(store.get_nodes(ms.user_id) and [dict(n, embedding=n["vector"]) for n in store.get_nodes(ms.user_id)], ms.user_id)
    differ = sum(store.search_nodes(table[f"query {j}"], ms.user_id, 5) != cos_store.search_nodes(table[f"query {j}"], ms.user_id, 5)
                 for j in range(len(queries)))
    assert differ > 0
    return dict(msys_l2=np.array(json.dumps({"table": table, "script": script, "states": states, "search": sm, "retrieval": orr,
                                             "components": comps, "dim": dim, "kw": kw})))


def gold_state_json(lz):
    """The reference's own save_state file (memory_system.py:1216-1239) after the three-conversation scenario, and
    what its own load_state (1241-1273) makes of it in a fresh MemorySystem."""
    import tempfile
    from make_golden import _run_msys_scenario
    r = _run_msys_scenario()
    ms = r["ms"]
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "memory_state.json")
        quiet(ms.save_state, path)
        text = open(path).read()
        ms2 = quiet(ref_harness.make_memory_system, FlatStore(), r["llm"], r["emb"])
        quiet(ms2.load_state, path)
    loaded = {"shards": {k: {"nodes": list(sh.nodes), "edges": [[a, b, e.weight, e.co_occurrence] for (a, b), e in sh.edges.items()]}
                         for k, sh in ms2.shards.items()},
              "super": list(ms2.super_nodes), "node_counter": ms2.node_counter, "conversation_count": ms2.conversation_count,
              "max_buffer_size": ms2.max_buffer_size, "prune_threshold": ms2.prune_threshold,
              # quirk: load_state rebinds self.shards / self.super_nodes to NEW dicts (memory_system.py:1246, 1257) but
              # self.buffer keeps the old ones, so the reference's own buffer is empty after a load; recorded as is,
              # next to the components of the graph that was saved (what a loader that re-points the buffer sees)
              "components_after_reference_load": [sorted(c) for c in ms2.buffer.get_connected_components()],
              "components_of_saved_graph": [sorted(c) for c in ms.buffer.get_connected_components()]}
    return dict(state_json=np.array(text), state_loaded=np.array(json.dumps(loaded)))


def main():
    lz = ref_harness.import_reference()
    out = {}
    for fn in (gold_components, gold_cross_super, gold_msys_l2, gold_state_json):
        out.update(fn(lz))
        print("ok", fn.__name__)
    path = os.path.join(OUT, "ref_r2.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
