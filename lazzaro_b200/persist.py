"""Import / export of the store's contents in the reference's table layout.

The reference persists three LanceDB tables whose Arrow schemas are fixed in core/vector_store.py:33-85
(`nodes`, `edges`, `user_profile`).  LanceDB itself is not installed here, but its tables are Arrow tables:
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

NODES, EDGES, PROFILES = "nodes", "edges", "user_profile"          # vector_store.py:21-23 table names


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
    profiles = tables.get(PROFILES)
    if profiles is not None:
        for p in profiles.to_pylist():
            st.save_profile(json.loads(p["data"]), user_id=p["user_id"])
    return st


def save(store: B200Store, directory: str, fmt: str = "arrow") -> Dict[str, str]:
    """Writes `<directory>/{nodes,edges,user_profile}.{arrow|parquet}`; returns the paths."""
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
    for name in (NODES, EDGES, PROFILES):
        path = os.path.join(directory, f"{name}.{fmt}")
        if os.path.exists(path):
            tables[name] = ipc.open_file(path).read_all() if fmt == "arrow" else pq.read_table(path)
    return import_tables(tables, **store_kwargs)
