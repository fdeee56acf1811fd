"""Import / export in the reference's table layout (core/vector_store.py:33-85)."""
import os
import re

import numpy as np
import pytest

REF_VS = "/root/reference/src/lazzaro/core/vector_store.py"

WANT = {
    "nodes": ["id:string", "user_id:string", "content:string", "vector:fixed_size_list<float>", "type:string",
              "timestamp:double", "access_count:int32", "last_accessed:double", "salience:float",
              "is_super_node:bool", "child_ids:string", "parent_id:string", "shard_key:string", "metadata:string"],
    "edges": ["id:string", "user_id:string", "source_id:string", "target_id:string", "weight:float",
              "edge_type:string", "co_occurrence:int32", "last_updated:double", "metadata:string"],
    "profiles": ["user_id:string", "data:string", "updated_at:double"],
}


def _sig(schema):
    out = []
    for f in schema:
        t = str(f.type)
        out.append(f"{f.name}:{'fixed_size_list<float>' if t.startswith('fixed_size_list') else t}")
    return out


def test_schemas_are_the_references():
    from lazzaro_b200 import persist
    assert _sig(persist.node_schema(1536)) == WANT["nodes"]
    assert persist.node_schema(1536).field("vector").type.list_size == 1536          # vector_store.py:37
    assert _sig(persist.EDGE_SCHEMA) == WANT["edges"]
    assert _sig(persist.PROFILE_SCHEMA) == WANT["profiles"]
    assert (persist.NODES, persist.EDGES, persist.PROFILES) == ("nodes", "edges", "profiles")      # vector_store.py:20-22


@pytest.mark.skipif(not os.path.exists(REF_VS), reason="/root/reference is only present in the build container")
def test_schemas_match_reference_source():
    """Field names and Arrow types read out of the reference's own schema literals, in order."""
    src = open(REF_VS).read()
    fields = re.findall(r'pa\.field\("(\w+)",\s*pa\.(\w+)\(', src)
    names = {"string": "string", "float64": "double", "float32": "float", "int32": "int32", "bool_": "bool",
             "list_": "fixed_size_list<float>"}
    got = [f"{n}:{names[t]}" for n, t in fields]
    assert got == WANT["nodes"] + WANT["edges"] + WANT["profiles"]
    names_in_ref = re.findall(r'self\.(\w+)_table_name = "(\w+)"', src)
    assert dict(names_in_ref) == {"nodes": "nodes", "edges": "edges", "profile": "profiles"}


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", ["arrow", "parquet"])
def test_round_trip_keeps_search_results(tmp_path, fmt):
    from lazzaro_b200 import persist, synth
    from lazzaro_b200.store import B200Store
    rows, shard = synth.clustered_rows(1200, 48, 5, seed=81)
    st = B200Store(metric="cosine")
    for u, (lo, hi) in {"alice": (0, 700), "bob": (700, 1200)}.items():
        st.add_nodes([{"id": f"n{i}", "content": f"c{i}", "embedding": rows[i].tolist(), "salience": 0.25 + (i % 4) / 8,
                       "shard_key": f"s{shard[i]}", "child_ids": [f"n{i + 1}"], "metadata": {"i": i},
                       "access_count": i % 7, "last_accessed": 1000.5 + i, "timestamp": 10.0 + i,
                       "is_super_node": i % 50 == 0, "parent_id": "p" if i % 3 == 0 else None} for i in range(lo, hi)], user_id=u)
    st.delete_nodes([f"n{i}" for i in range(0, 700, 9)], user_id="alice")
    st.add_edges([{"source": "n1", "target": "n2", "weight": 0.9, "metadata": {"k": 1}},
                  {"source_id": "n2", "target_id": "n4", "type": "x", "co_occurrence": 3}], "alice")
    st.save_profile({"likes": ["tea"]}, "bob")
    paths = persist.save(st, str(tmp_path), fmt=fmt)
    assert sorted(paths) == ["edges", "nodes", "profiles"]
    tables = persist.export_tables(st)
    assert tables["nodes"].schema == persist.node_schema(48) and tables["nodes"].num_rows == 1200 - len(range(0, 700, 9))
    assert tables["edges"].schema == persist.EDGE_SCHEMA and tables["profiles"].schema == persist.PROFILE_SCHEMA
    back = persist.load(str(tmp_path), fmt=fmt, metric="cosine")
    qs = synth.queries_near_rows(rows, 8, seed=82)
    for u in ("alice", "bob", "nobody"):
        for q in qs:
            assert back.search_nodes(q.tolist(), user_id=u, limit=7) == st.search_nodes(q.tolist(), user_id=u, limit=7)
        a, b = back.get_nodes(u), st.get_nodes(u)
        assert a == b
    assert back.get_edges(user_id="alice") == st.get_edges(user_id="alice")
    assert back.load_profile("bob") == {"likes": ["tea"]} and back.load_profile("alice") is None
    # a dump written under the old table name still loads
    import os
    os.rename(paths["profiles"], os.path.join(str(tmp_path), f"user_profile.{fmt}"))
    assert persist.load(str(tmp_path), fmt=fmt, metric="cosine").load_profile("bob") == {"likes": ["tea"]}


@pytest.mark.gpu
def test_empty_store_round_trip(tmp_path):
    from lazzaro_b200 import persist
    from lazzaro_b200.store import B200Store
    st = B200Store()
    persist.save(st, str(tmp_path))
    back = persist.load(str(tmp_path))
    assert back.get_nodes("default") == [] and back.search_nodes([0.0] * 4) == []


@pytest.mark.gpu
def test_json_state_round_trip_and_reference_layout(tmp_path):
    """save_state / load_state (memory_system.py:1216-1273): the file has the reference's layout and a second
    MemorySystem loaded from it has the same graph (edge weights come back from HBM as float32 values)."""
    import json
    from lazzaro_b200.graph import Edge, Node
    from lazzaro_b200.memory_system import MemorySystem

    class _LLM:
        def completion(self, messages, response_format=None):
            return "{}"

    class _Emb:
        def embed(self, t):
            return [0.0] * 8

        def batch_embed(self, ts):
            return [[0.0] * 8 for _ in ts]

    rng = np.random.default_rng(3)
    ms = MemorySystem(llm_provider=_LLM(), embedding_provider=_Emb(), max_buffer_size=77, prune_threshold=0.4)
    for i in range(30):
        ms.buffer.add_node(Node(id=f"node_{i}", content=f"c{i}", embedding=rng.standard_normal(8).tolist(), salience=0.3 + i / 100,
                                access_count=i % 5, shard_key=f"s{i % 3}"))
    for _ in range(60):
        a, b = rng.integers(0, 30, 2)
        ms.buffer.add_edge(Edge(source=f"node_{a}", target=f"node_{b}", weight=float(np.float32(rng.uniform(0.2, 0.9)))))
    ms.super_nodes["super_s0_1"] = Node(id="super_s0_1", content="topic", embedding=[0.5] * 8, is_super_node=True,
                                        child_ids=["node_0", "node_3"], shard_key="s0")
    ms.node_counter, ms.conversation_count = 30, 4
    path = str(tmp_path / "state.json")
    assert ms.save_state(path) == f"✓ State saved to {path}"
    state = json.load(open(path))
    assert list(state) == ["shards", "super_nodes", "profile", "node_counter", "conversation_count", "settings"]
    assert list(state["shards"]["s0"]) == ["nodes", "edges"]
    assert list(state["shards"]["s0"]["nodes"][0]) == ["id", "content", "embedding", "type", "timestamp", "access_count",
                                                       "last_accessed", "salience", "is_super_node", "child_ids", "parent_id", "shard_key"]
    assert list(state["shards"]["s0"]["edges"][0]) == ["source", "target", "weight", "edge_type", "co_occurrence", "last_updated"]
    assert list(state["settings"]) == ["auto_consolidate", "consolidate_every", "auto_prune", "prune_threshold", "max_buffer_size"]
    ms2 = MemorySystem(llm_provider=_LLM(), embedding_provider=_Emb())
    assert ms2.load_state(str(tmp_path / "missing.json")).startswith("⚠")
    assert ms2.load_state(path) == f"✓ State loaded from {path}"
    assert (ms2.node_counter, ms2.conversation_count, ms2.max_buffer_size, ms2.prune_threshold) == (30, 4, 77, 0.4)
    assert list(ms2.shards) == list(ms.shards) and list(ms2.super_nodes) == ["super_s0_1"]
    for k in ms.shards:
        assert list(ms2.shards[k].nodes) == list(ms.shards[k].nodes)
        e1, e2 = ms.shards[k].edges, ms2.shards[k].edges
        assert list(e1) == list(e2)
        for key in e1:
            assert (e1[key].weight, e1[key].co_occurrence, e1[key].edge_type) == (e2[key].weight, e2[key].co_occurrence, e2[key].edge_type)
    assert ms2.buffer.get_connected_components() == ms.buffer.get_connected_components()
    assert ms2.buffer.decay_and_prune(0.01, 0.5) == ms.buffer.decay_and_prune(0.01, 0.5)


@pytest.mark.gpu
def test_load_state_reads_the_references_own_file(tmp_path):
    """A state file written by the reference's save_state (tests/golden/make_golden_r2.py gold_state_json) loads into
    the mirror as the reference's own load_state loads it: same shards, node order, edges (weights to float32),
    counters, settings and components; saving it again reproduces the file's structure."""
    import json
    from lazzaro_b200.memory_system import MemorySystem
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_r2.npz"))
    want = json.loads(str(g["state_loaded"]))
    path = str(tmp_path / "memory_state.json")
    open(path, "w").write(str(g["state_json"]))

    class _P:
        def completion(self, *a, **k):
            return "{}"

        def embed(self, t):
            return [0.0] * 32

        def batch_embed(self, ts):
            return [[0.0] * 32 for _ in ts]

    ms = MemorySystem(llm_provider=_P(), embedding_provider=_P())
    assert ms.load_state(path).startswith("✓")
    assert {k: list(sh.nodes) for k, sh in ms.shards.items()} == {k: v["nodes"] for k, v in want["shards"].items()}
    for k, v in want["shards"].items():
        got = [[a, b, e.weight, e.co_occurrence] for (a, b), e in ms.shards[k].edges.items()]
        assert [x[:2] + x[3:] for x in got] == [x[:2] + x[3:] for x in v["edges"]]
        np.testing.assert_allclose([x[2] for x in got], [x[2] for x in v["edges"]], rtol=0, atol=3e-8)
    assert list(ms.super_nodes) == want["super"]
    assert (ms.node_counter, ms.conversation_count, ms.max_buffer_size, ms.prune_threshold) == \
        (want["node_counter"], want["conversation_count"], want["max_buffer_size"], want["prune_threshold"])
    # The reference's load_state leaves self.buffer bound to the dicts it had before the load (memory_system.py:1246,
    # 1257 rebind self.shards / self.super_nodes only), so ITS buffer is empty afterwards; the mirror re-points the
    # buffer at what was loaded and therefore sees the saved graph's components.
    assert want["components_after_reference_load"] == []
    norm = lambda c: sorted("_".join(x.split("_")[:2]) if x.startswith("super_") else x for x in c)
    got_c, want_c = [norm(c) for c in ms.buffer.get_connected_components()], [norm(c) for c in want["components_of_saved_graph"]]
    assert [c for c in got_c if not c[0].startswith("super_")] == [c for c in want_c if not c[0].startswith("super_")]
    assert sorted(c for c in got_c if c[0].startswith("super_")) == sorted(c for c in want_c if c[0].startswith("super_"))
    again = str(tmp_path / "again.json")
    ms.save_state(again)
    a, b = json.load(open(path)), json.load(open(again))
    assert list(a) == list(b) and list(a["shards"]) == list(b["shards"]) and a["settings"] == b["settings"]
    assert a["super_nodes"] == b["super_nodes"] and a["profile"]["data"] == b["profile"]["data"]
    for k in a["shards"]:
        assert a["shards"][k]["nodes"] == b["shards"][k]["nodes"]
        assert [(e["source"], e["target"], e["co_occurrence"], e["edge_type"]) for e in a["shards"][k]["edges"]] == \
               [(e["source"], e["target"], e["co_occurrence"], e["edge_type"]) for e in b["shards"][k]["edges"]]
