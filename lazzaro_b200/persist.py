"""Import / export of the store's contents in the reference's table layout.

The reference persists three LanceDB tables whose Arrow schemas are fixed in core/vector_store.py:33-85
(`nodes`, `edges`, `profiles`).  LanceDB itself is not installed here, but its tables are Arrow tables:
what this module writes (Arrow IPC files, or Parquet) has exactly those schemas — same field names, order and
types, `vector` as FixedSizeList<float32>[dim], `child_ids` / `metadata` as JSON strings — so a site that has
lancedb can hand them straight to `db.create_table(name, data=table)`, and `table.to_arrow()` of an existing
dump loads back through `import_tables`.

Vectors leave and enter HBM in one copy each way; no Python list of floats is built.
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, Optional

import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pyarrow.parquet as pq
import torch

from .store import B200Store

NODES, EDGES, PROFILES = "nodes", "edges", "profiles"              # vector_store.py:20-22 table names
_OLD_PROFILES = "user_profile"                                      # name used by dumps this module wrote before; still read


def node_schema(dim: int = 1536) -> pa.Schema:
    """vector_store.py:33-48 (the reference fixes dim = 1536)."""
    return pa.schema([
        pa.field("id", pa.string()), pa.field("user_id", pa.string()), pa.field("content", pa.string()),
        pa.field("vector", pa.list_(pa.float32(), dim)), pa.field("type", pa.string()),
        pa.field("timestamp", pa.float64()), pa.field("access_count", pa.int32()),
        pa.field("last_accessed", pa.float64()), pa.field("salience", pa.float32()),
        pa.field("is_super_node", pa.bool_()), pa.field("child_ids", pa.string()),
        pa.field("parent_id", pa.string()), pa.field("shard_key", pa.string()), pa.field("metadata", pa.string()),
    ])


EDGE_SCHEMA = pa.schema([                                           # vector_store.py:59-69
    pa.field("id", pa.string()), pa.field("user_id", pa.string()), pa.field("source_id", pa.string()),
    pa.field("target_id", pa.string()), pa.field("weight", pa.float32()), pa.field("edge_type", pa.string()),
    pa.field("co_occurrence", pa.int32()), pa.field("last_updated", pa.float64()), pa.field("metadata", pa.string()),
])

PROFILE_SCHEMA = pa.schema([                                        # vector_store.py:76-80
    pa.field("user_id", pa.string()), pa.field("data", pa.string()), pa.field("updated_at", pa.float64()),
])


def export_tables(store: B200Store) -> Dict[str, pa.Table]:
    """The live rows, edges and profiles of every user as Arrow tables with the reference's schemas."""
    with store._lock:
        dim = store.dim or 0
        live = np.nonzero(store._row_user[:store._n] >= 0)[0] if store._n else np.empty(0, np.int64)
        if live.size:
            idx = torch.from_numpy(live).to(store.device)
            vecs = store._rows.index_select(0, idx)[:, :dim].contiguous().cpu().numpy()
        else:
            vecs = np.empty((0, dim), np.float32)
        cols: Dict[str, list] = {f.name: [] for f in node_schema(max(dim, 1)) if f.name != "vector"}
        for r in live.tolist():
            m = store._row_meta.get(r, {})
            cols["id"].append(store._id_of(r))
            cols["user_id"].append(store._user_names[int(store._row_user[r])])
            cols["content"].append(m.get("content", ""))
            cols["type"].append(m.get("type", "semantic"))
            cols["timestamp"].append(float(m.get("timestamp", 0.0)))
            cols["access_count"].append(int(m.get("access_count", 0)))
            cols["last_accessed"].append(float(m.get("last_accessed", 0.0)))
            cols["salience"].append(float(m.get("salience", 0.5)))
            cols["is_super_node"].append(bool(m.get("is_super_node", False)))
            cols["child_ids"].append(m.get("child_ids", "[]"))
            cols["parent_id"].append(m.get("parent_id", "") or "")
            cols["shard_key"].append(m.get("shard_key", "default"))
            cols["metadata"].append(m.get("metadata", "{}"))
        schema = node_schema(max(dim, 1))
        arrays = []
        for f in schema:
            if f.name == "vector":
                flat = pa.array(vecs.reshape(-1), type=pa.float32())
                arrays.append(pa.FixedSizeListArray.from_arrays(flat, max(dim, 1)) if dim else pa.array([], type=f.type))
            else:
                arrays.append(pa.array(cols[f.name], type=f.type))
        nodes = pa.Table.from_arrays(arrays, schema=schema)
        edges = pa.Table.from_pylist([{k: e[k] for k in EDGE_SCHEMA.names} for e in store._edges], schema=EDGE_SCHEMA)
        profiles = pa.Table.from_pylist([{"user_id": u, "data": p["data"], "updated_at": float(p["updated_at"])}
                                         for u, p in store._profiles.items()], schema=PROFILE_SCHEMA)
        return {NODES: nodes, EDGES: edges, PROFILES: profiles}


def import_tables(tables: Dict[str, pa.Table], store: Optional[B200Store] = None, chunk: int = 65536,
                  **store_kwargs: Any) -> B200Store:
    """Loads tables with the reference's schemas (e.g. `lancedb_table.to_arrow()`) into a store, user by user
    in table order (so row order — the tie-break of every search — is the dump's)."""
    nodes = tables.get(NODES)
    st = store or B200Store(**store_kwargs)
    if nodes is not None and nodes.num_rows:
        vec = nodes.column("vector").combine_chunks()
        dim = vec.type.list_size
        flat = vec.flatten().to_numpy(zero_copy_only=False).astype(np.float32, copy=False).reshape(-1, dim)
        other = {name: nodes.column(name).to_pylist() for name in nodes.schema.names if name != "vector"}
        users = other["user_id"]
        lo = 0
        while lo < nodes.num_rows:                                   # runs of one user, at most `chunk` rows each
            hi = lo + 1
            while hi < nodes.num_rows and hi - lo < chunk and users[hi] == users[lo]:
                hi += 1
            batch = []
            for i in range(lo, hi):
                batch.append({"id": other["id"][i], "content": other["content"][i], "embedding": flat[i],
                              "type": other["type"][i], "timestamp": other["timestamp"][i],
                              "access_count": other["access_count"][i], "last_accessed": other["last_accessed"][i],
                              "salience": other["salience"][i], "is_super_node": other["is_super_node"][i],
                              "child_ids": json.loads(other["child_ids"][i] or "[]"), "parent_id": other["parent_id"][i] or None,
                              "shard_key": other["shard_key"][i], "metadata": json.loads(other["metadata"][i] or "{}")})
            st.add_nodes(batch, user_id=users[lo])
            lo = hi
    edges = tables.get(EDGES)
    if edges is not None and edges.num_rows:
        by_user: Dict[str, list] = {}
        for e in edges.to_pylist():
            e = dict(e)
            e["metadata"] = json.loads(e.get("metadata") or "{}")
            by_user.setdefault(e.pop("user_id"), []).append(e)
        for u, es in by_user.items():
            st.add_edges(es, user_id=u)
    profiles = tables.get(PROFILES, tables.get(_OLD_PROFILES))
    if profiles is not None:
        for p in profiles.to_pylist():
            st.save_profile(json.loads(p["data"]), user_id=p["user_id"])
    return st


def save(store: B200Store, directory: str, fmt: str = "arrow") -> Dict[str, str]:
    """Writes `<directory>/{nodes,edges,profiles}.{arrow|parquet}`; returns the paths."""
    if fmt not in ("arrow", "parquet"):
        raise ValueError("fmt must be 'arrow' or 'parquet'")
    os.makedirs(directory, exist_ok=True)
    out = {}
    for name, table in export_tables(store).items():
        path = os.path.join(directory, f"{name}.{fmt}")
        if fmt == "arrow":
            with ipc.new_file(path, table.schema) as w:
                w.write_table(table)
        else:
            pq.write_table(table, path)
        out[name] = path
    return out


def load(directory: str, fmt: str = "arrow", **store_kwargs: Any) -> B200Store:
    """Reads what `save` wrote (or files of the same names and schemas) into a new store."""
    tables = {}
    for name in (NODES, EDGES, PROFILES, _OLD_PROFILES):
        path = os.path.join(directory, f"{name}.{fmt}")
        if os.path.exists(path):
            tables[name] = ipc.open_file(path).read_all() if fmt == "arrow" else pq.read_table(path)
    return import_tables(tables, **store_kwargs)


# ---- the JSON graph state (memory_system.py:1216-1273) ------------------------------------------------------

def state_dict(ms) -> Dict[str, Any]:
    """What the reference's `save_state` serialises (memory_system.py:1217-1237): per shard the nodes and edges as
    `asdict` dictionaries (models/graph.py:58-60, 102-104), the super-nodes, the profile, the counters and five
    settings.  `ms` is a lazzaro_b200.memory_system.MemorySystem; edge weights are read back from HBM."""
    node = lambda n: {"id": n.id, "content": n.content, "embedding": list(n.embedding), "type": n.type,
                      "timestamp": n.timestamp, "access_count": n.access_count, "last_accessed": n.last_accessed,
                      "salience": n.salience, "is_super_node": n.is_super_node, "child_ids": list(n.child_ids),
                      "parent_id": n.parent_id, "shard_key": n.shard_key}
    edge = lambda e: {"source": e.source, "target": e.target, "weight": e.weight, "edge_type": e.edge_type,
                      "co_occurrence": e.co_occurrence, "last_updated": e.last_updated}
    profile = ms.profile if isinstance(ms.profile, dict) and "data" in ms.profile else {"data": ms.profile or {}, "last_updated": 0.0}
    return {
        "shards": {k: {"nodes": [node(n) for n in v.nodes.values()], "edges": [edge(e) for e in v.edges.values()]}
                   for k, v in ms.shards.items()},
        "super_nodes": [node(n) for n in ms.super_nodes.values()],
        "profile": profile,
        "node_counter": ms.node_counter,
        "conversation_count": ms.conversation_count,
        "settings": {"auto_consolidate": ms.auto_consolidate, "consolidate_every": ms.consolidate_every,
                     "auto_prune": ms.auto_prune, "prune_threshold": ms.prune_threshold,
                     "max_buffer_size": ms.max_buffer_size},
    }


def save_state(ms, filename: str = "memory_state.json") -> str:
    """memory_system.py:1216-1239: same file layout, so either implementation reads the other's file."""
    with open(filename, "w") as f:
        json.dump(state_dict(ms), f, indent=2)
    return f"✓ State saved to {filename}"


def load_state(ms, filename: str = "memory_state.json") -> str:
    """memory_system.py:1241-1273: rebuilds shards (nodes, then edges through add_edge, so a repeated key is
    reinforced exactly as the reference's loader does), super-nodes, profile, counters and settings.  The edge
    tables are uploaded to HBM shard by shard in one batch each."""
    from .graph import Edge, MemoryShard, Node
    from .memory_system import BufferGraph
    try:
        with open(filename, "r") as f:
            state = json.load(f)
    except FileNotFoundError:
        return f"⚠ File {filename} not found"
    nf, ef = Node.__dataclass_fields__, Edge.__dataclass_fields__
    ms.shards = {}
    for key, data in state.get("shards", {}).items():
        shard = MemoryShard(key, device=str(ms.store.device))
        for nd in data.get("nodes", []):
            shard.add_node(Node(**{k: v for k, v in nd.items() if k in nf}))
        shard.add_edges([Edge(**{k: v for k, v in ed.items() if k in ef}) for ed in data.get("edges", [])])
        ms.shards[key] = shard
    ms.super_nodes = {n["id"]: Node(**{k: v for k, v in n.items() if k in nf}) for n in state.get("super_nodes", [])}
    ms.buffer = BufferGraph(ms.shards, ms.super_nodes)
    ms.buffer._device = str(ms.store.device)
    prof = state.get("profile", {})
    if prof:
        ms.profile = {"data": prof.get("data", {}), "last_updated": prof.get("last_updated", 0.0)}
    ms.node_counter = state.get("node_counter", 0)
    ms.conversation_count = state.get("conversation_count", 0)
    for key, val in state.get("settings", {}).items():
        if hasattr(ms, key):
            setattr(ms, key, val)
    return f"✓ State loaded from {filename}"
