"""Oracle: the reference's own per-object Python loops, restated.  TEST INFRASTRUCTURE — see oracle/__init__.py.

`/root/reference` does not travel to the GPU box, so the CPU baselines that BASELINE.md §4 asks for (C1, C3, C4: what the
reference itself does, object by object) are restated here line for line in behaviour — plain Python over small objects,
`np.array(list)` conversions and all — and timed by bench.py beside the vectorised numpy ports.  Nothing in the product
imports this module.
"""
from __future__ import annotations

import numpy as np


class Edge:                                  # models/graph.py:90-95 (the fields the loops touch)
    __slots__ = ("source", "target", "weight", "co_occurrence")

    def __init__(self, source, target, weight, co_occurrence=1):
        self.source, self.target, self.weight, self.co_occurrence = source, target, weight, co_occurrence


class Node:                                  # models/graph.py:39-50 (the fields the loops touch)
    __slots__ = ("id", "embedding", "salience", "is_super_node")

    def __init__(self, id, embedding, salience=0.5, is_super_node=False):
        self.id, self.embedding, self.salience, self.is_super_node = id, embedding, salience, is_super_node


def make_edge_dict(src, dst, weight):
    """{(source, target): Edge} as MemoryShard keeps it (memory_shard.py:42-52)."""
    return {(int(s), int(t)): Edge(int(s), int(t), float(w)) for s, t, w in zip(src, dst, weight)}


def decay_and_prune(edges: dict, nodes: dict, decay_rate: float = 0.01, threshold: float = 0.5) -> int:
    """memory_shard.py:64-77 then 79-84, as written: a loop over every Edge object multiplying its weight, a loop over
    every Node adjusting its salience, then a list comprehension of the weak keys and one `del` per key."""
    for edge in edges.values():
        edge.weight *= (1 - decay_rate)
    for node in nodes.values():
        if node.salience > 0.2:
            node.salience = 0.2 + (node.salience - 0.2) * (1 - decay_rate)
        else:
            node.salience = 0.2
    to_remove = [key for key, edge in edges.items() if edge.weight < threshold]
    for key in to_remove:
        del edges[key]
    return len(to_remove)


def cosine_similarity(v1, v2) -> float:
    """memory_system.py:197-203, including the two list -> array conversions it performs per call."""
    if not v1 or not v2:
        return 0.0
    a, b = np.array(v1), np.array(v2)
    norm = np.linalg.norm(a) * np.linalg.norm(b)
    return float(np.dot(a, b) / norm) if norm > 0 else 0.0


def link_to_existing(new_nodes, existing_nodes, edges: dict):
    """memory_system.py:838-891 for one shard table: per new node the cosine against EVERY existing node, a sort, the
    first three above 0.5, an existence check in both directions, an Edge of weight 0.8 * similarity."""
    created = 0
    for new in new_nodes:
        sims = [(ex.id, cosine_similarity(new.embedding, ex.embedding)) for ex in existing_nodes if not ex.is_super_node]
        sims.sort(key=lambda x: x[1], reverse=True)
        for ex_id, sim in sims[:3]:
            if sim > 0.5 and (new.id, ex_id) not in edges and (ex_id, new.id) not in edges:
                edges[(new.id, ex_id)] = Edge(new.id, ex_id, sim * 0.8)
                created += 1
    return created


def search_memories(store, resident: dict, query_emb, user_id: str = "default", limit: int = 5):
    """memory_system.py:1460-1472 after the embedding call: the store's search, then id -> resident node."""
    ids = store.search_nodes(query_emb, user_id=user_id, limit=limit)
    return [resident[i] for i in ids if i in resident]
