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
    "user_profile": ["user_id:string", "data:string", "updated_at:double"],
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
    assert _sig(persist.PROFILE_SCHEMA) == WANT["user_profile"]


@pytest.mark.skipif(not os.path.exists(REF_VS), reason="/root/reference is only present in the build container")
def test_schemas_match_reference_source():
    """Field names and Arrow types read out of the reference's own schema literals, in order."""
    src = open(REF_VS).read()
    fields = re.findall(r'pa\.field\("(\w+)",\s*pa\.(\w+)\(', src)
    names = {"string": "string", "float64": "double", "float32": "float", "int32": "int32", "bool_": "bool",
             "list_": "fixed_size_list<float>"}
    got = [f"{n}:{names[t]}" for n, t in fields]
    assert got == WANT["nodes"] + WANT["edges"] + WANT["user_profile"]


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
    assert sorted(paths) == ["edges", "nodes", "user_profile"]
    tables = persist.export_tables(st)
    assert tables["nodes"].schema == persist.node_schema(48) and tables["nodes"].num_rows == 1200 - len(range(0, 700, 9))
    assert tables["edges"].schema == persist.EDGE_SCHEMA and tables["user_profile"].schema == persist.PROFILE_SCHEMA
    back = persist.load(str(tmp_path), fmt=fmt, metric="cosine")
    qs = synth.queries_near_rows(rows, 8, seed=82)
    for u in ("alice", "bob", "nobody"):
        for q in qs:
            assert back.search_nodes(q.tolist(), user_id=u, limit=7) == st.search_nodes(q.tolist(), user_id=u, limit=7)
        a, b = back.get_nodes(u), st.get_nodes(u)
        assert a == b
    assert back.get_edges(user_id="alice") == st.get_edges(user_id="alice")
    assert back.load_profile("bob") == {"likes": ["tea"]} and back.load_profile("alice") is None


@pytest.mark.gpu
def test_empty_store_round_trip(tmp_path):
    from lazzaro_b200 import persist
    from lazzaro_b200.store import B200Store
    st = B200Store()
    persist.save(st, str(tmp_path))
    back = persist.load(str(tmp_path))
    assert back.get_nodes("default") == [] and back.search_nodes([0.0] * 4) == []
