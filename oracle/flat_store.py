"""Oracle: an in-memory `Store` (interfaces.py:55-102) with the semantics of `LanceDBStore`.
TEST INFRASTRUCTURE — see oracle/__init__.py.

Follows /root/reference/src/lazzaro/core/vector_store.py method by method so that the reference's
own `MemorySystem` can run on top of it (injected through `MemorySystem(store=...)`,
memory_system.py:83,104-107) and so that the CUDA-backed store can be compared with it call for
call.  Search is `oracle.flat_search` (exhaustive, float32 rows, float64 accumulation).
"""
from __future__ import annotations

import json
import time
from typing import Any, Dict, List, Optional

import numpy as np

from . import flat_search


class FlatStore:
    def __init__(self, db_dir: str = "db", dim: Optional[int] = None, metric: str = "cosine"):
        self.db_dir = db_dir
        self.dim = dim
        self.metric = metric
        self._rows: List[Dict[str, Any]] = []          # node rows in insertion order
        self._vecs: List[np.ndarray] = []
        self._edges: List[Dict[str, Any]] = []
        self._profiles: Dict[str, Dict[str, Any]] = {}
        self._version = 1
        self._nodes_table = self                        # memory_system.py:1434-1439 pokes at this

    # --- nodes (vector_store.py:89-148) ---
    def add_nodes(self, nodes: List[Dict[str, Any]], user_id: str = "default"):
        if not nodes:                                   # vector_store.py:90-91
            return
        for node in nodes:
            v = np.asarray(node["embedding"], dtype=np.float32)
            if self.dim is None:
                self.dim = int(v.shape[0])
            if v.shape[0] != self.dim:
                raise ValueError(f"embedding has {v.shape[0]} dims, store has {self.dim}")
            self._rows.append({                         # vector_store.py:95-110
                "id": node["id"], "user_id": user_id, "content": node["content"],
                "type": node.get("type", "semantic"),
                "timestamp": node.get("timestamp", 0.0),
                "access_count": node.get("access_count", 0),
                "last_accessed": node.get("last_accessed", 0.0),
                "salience": float(np.float32(node.get("salience", 0.5))),
                "is_super_node": node.get("is_super_node", False),
                "child_ids": json.dumps(node.get("child_ids", [])),
                "parent_id": node.get("parent_id") or "",
                "shard_key": node.get("shard_key", "default"),
                "metadata": json.dumps(node.get("metadata", {})),
            })
            self._vecs.append(v)
        self._version += 1

    def get_nodes(self, user_id: str = "default") -> List[Dict[str, Any]]:
        out = []
        for r, v in zip(self._rows, self._vecs):       # vector_store.py:118-129
            if r["user_id"] != user_id:
                continue
            d = dict(r)
            d["vector"] = v.tolist()
            d["child_ids"] = json.loads(d["child_ids"])
            d["metadata"] = json.loads(d["metadata"])
            out.append(d)
        return out

    def search_nodes(self, query_emb: List[float], user_id: str = "default", limit: int = 5) -> List[str]:
        if not self._rows:                              # vector_store.py:133-134
            return []
        rows = np.stack(self._vecs)
        mask = np.fromiter((r["user_id"] == user_id for r in self._rows), dtype=bool, count=len(self._rows))
        idx, _ = flat_search.flat_topk(rows, query_emb, limit, self.metric, mask)
        return [self._rows[i]["id"] for i in idx]       # vector_store.py:140 (ids only)

    def delete_nodes(self, node_ids: List[str], user_id: str = "default"):
        if node_ids is None or len(node_ids) == 0:      # vector_store.py:143-145: all of the user
            dead = [i for i, r in enumerate(self._rows) if r["user_id"] == user_id]
        else:
            ids = set(node_ids)
            dead = [i for i, r in enumerate(self._rows) if r["user_id"] == user_id and r["id"] in ids]
        for i in reversed(dead):
            del self._rows[i]
            del self._vecs[i]
        self._version += 1

    def get_latest_version(self) -> int:                # vector_store.py:150-156
        return self._version

    # --- edges (vector_store.py:159-212) ---
    def add_edges(self, edges: List[Dict[str, Any]], user_id: str = "default"):
        if not edges:
            return
        for edge in edges:
            src = edge.get("source") or edge.get("source_id")
            tgt = edge.get("target") or edge.get("target_id")
            self._edges.append({
                "id": edge.get("id", f"{src}_{tgt}"), "user_id": user_id,
                "source_id": src, "target_id": tgt,
                "weight": float(np.float32(edge.get("weight", 1.0))),     # float32 column, vector_store.py:63
                "edge_type": edge.get("edge_type") or edge.get("type", "relates_to"),
                "co_occurrence": edge.get("co_occurrence", 1),
                "last_updated": edge.get("last_updated", 0.0),
                "metadata": json.dumps(edge.get("metadata", {})),
            })

    def delete_edges(self, source_id: Optional[str] = None, user_id: str = "default"):
        self._edges = [e for e in self._edges
                       if not (e["user_id"] == user_id and (not source_id or e["source_id"] == source_id))]

    def get_edges(self, source_id: Optional[str] = None, user_id: str = "default") -> List[Dict[str, Any]]:
        out = []
        for e in self._edges:
            if e["user_id"] != user_id or (source_id and e["source_id"] != source_id):
                continue
            d = dict(e)
            d["type"] = d["edge_type"]                  # vector_store.py:204-205
            d["metadata"] = json.loads(d["metadata"])
            out.append(d)
        return out

    # --- profiles (vector_store.py:216-233) ---
    def save_profile(self, profile_data: Dict[str, Any], user_id: str = "default"):
        self._profiles[user_id] = {"data": json.dumps(profile_data), "updated_at": time.time()}

    def load_profile(self, user_id: str = "default") -> Optional[Dict[str, Any]]:
        p = self._profiles.get(user_id)
        return json.loads(p["data"]) if p else None

    def close(self):
        pass

    # legacy vector-store surface (tests/test_vector_store.py:15,43,46,66)
    def add(self, nodes):
        return self.add_nodes(nodes)

    def search(self, query_emb, limit: int = 5):
        return self.search_nodes(query_emb, limit=limit)

    def delete(self, ids):
        return self.delete_nodes(ids)

    # memory_system.py:1438 does store._nodes_table.search().select(["user_id"]).to_list()
    def search(self, query_emb=None, limit: int = 5):          # noqa: F811  (table-style when no query)
        if query_emb is None:
            return _UserScan(self)
        return self.search_nodes(query_emb, limit=limit)


class _UserScan:
    def __init__(self, store):
        self._s = store

    def select(self, cols):
        return self

    def to_list(self):
        return [{"user_id": r["user_id"]} for r in self._s._rows]
