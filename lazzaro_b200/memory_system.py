"""MemorySystem — host-side mirror of the reference orchestrator's hot-path entry points.

Same names, arguments and results as /root/reference/src/lazzaro/core/memory_system.py for the part
of `MemorySystem` that SURVEY §8 puts on the hot path, written from scratch around the device store
and graph of this package:

    search_memories            memory_system.py:1460-1472
    _optimized_retrieval       memory_system.py:453-514   (super-node gate 0.4, first 10 children, vector fallback)
    end_conversation           memory_system.py:580-649   (consolidate, decay 0.01, prune, buffer limit)
    _async_consolidate         memory_system.py:651-785   (extract facts, dedup > 0.95, insert, link, super-nodes)
    _link_within_shards        memory_system.py:797-836
    _link_to_existing_memories memory_system.py:838-891
    _create_super_nodes_for_shard  memory_system.py:893-933
    _enforce_buffer_limit      memory_system.py:535-578
    _merge_similar_nodes       memory_system.py:1065-1120 (a no-op as written in the reference — SURVEY Q1 — and here)
    run_consolidation          memory_system.py:935-1010  (merge, components, prune; profile extraction is out of scope)
    _load_from_persistence / check_for_updates / switch_user   memory_system.py:1304-1428, 1474-1486

What differs is where the arithmetic runs: every cosine, top-k, mean, decay and prune is a kernel call
(batched where the reference loops), and the reference's one-search-per-memory loops become one batched
search per step.  Chat / prompt building / profile extraction / persistence formats are out of scope
 and absent.  LLM and embedding providers follow the reference protocols
(core/interfaces.py:3-53) and are always supplied by the caller — there is no network here.
"""
from __future__ import annotations

import json
import time
from collections import defaultdict
from typing import Dict, List, Optional, Set, Tuple

import numpy as np
import torch

from . import _lib, consolidate, kernels
from .graph import BufferGraph as _DeviceBufferGraph
from .graph import Edge, MemoryShard, Node
from .store import B200Store

_TOPIC_KEYWORDS = {          # memory_system.py:157-162 (behavioural data: which words route to which shard)
    "work": ["work", "project", "meeting", "deadline", "client", "colleague"],
    "personal": ["family", "friend", "hobby", "home", "personal"],
    "learning": ["learn", "study", "course", "book", "tutorial", "practice"],
    "health": ["health", "exercise", "diet", "sleep", "medical", "fitness"],
}


class BufferGraph(_DeviceBufferGraph):
    """The union view over shards and super-nodes (core/buffer_graph.py:28-120) on top of the device
    decay / prune fan-out of lazzaro_b200.graph.BufferGraph."""

    @property
    def nodes(self) -> Dict[str, Node]:
        out = dict(self.super_nodes)
        for shard in self.shards.values():
            out.update(shard.nodes)
        return out

    @property
    def edges(self) -> Dict[Tuple[str, str], Edge]:
        out: Dict[Tuple[str, str], Edge] = {}
        for shard in self.shards.values():
            out.update(shard.edges)
        return out

    def add_node(self, node: Node):
        key = node.shard_key or "default"
        if key not in self.shards:
            self.shards[key] = MemoryShard(key, device=self._device)
        self.shards[key].add_node(node)

    def add_edge(self, edge: Edge):                       # buffer_graph.py:51-61
        for shard in self.shards.values():
            if edge.source in shard.nodes:
                shard.add_edge(edge)
                return
        if "default" in self.shards:
            self.shards["default"].add_edge(edge)

    def get_node(self, node_id: str) -> Optional[Node]:   # buffer_graph.py:63-70
        if node_id in self.super_nodes:
            return self.super_nodes[node_id]
        for shard in self.shards.values():
            if node_id in shard.nodes:
                return shard.nodes[node_id]
        return None

    def get_neighbors(self, node_id: str, min_weight: float = 0.3) -> List[str]:
        for shard in self.shards.values():
            if node_id in shard.nodes:
                return shard.get_neighbors(node_id, min_weight)
        return []

    def update_access(self, node_id: str):                # buffer_graph.py:79-85
        node = self.get_node(node_id)
        if node:
            node.access_count += 1
            node.last_accessed = time.time()
            node.salience = min(1.0, node.salience + 0.05)

    def get_connected_components(self) -> List[Set[str]]:
        """buffer_graph.py:99-120 on the device (lz_reach_components): one launch sequence over the concatenated
        shard edge tables, no host traversal.

        Node indices follow the reference's visiting order (`self.nodes`: super-nodes, then the shards' nodes in
        dict order; ids that appear only as edge endpoints come last).  The reference follows an edge from a node
        only if the edge sits in the shard the node lives in (buffer_graph.py:72-77), so every edge gets walk
        flags from the residency of its endpoints IN ITS OWN SHARD: both resident = two-way; a cross-shard link or
        an endpoint that is not resident = one-way, away from the resident end; the kernel then assigns each node
        the smallest index that can reach it, which is the root the reference's DFS gives it.  Components come
        back in the reference's order (by root)."""
        index: Dict[str, int] = {}
        for nid in self.nodes:                                   # visiting order of buffer_graph.py:114
            index.setdefault(nid, len(index))
        n_resident = len(index)
        plan, total = [], 0
        for shard in self.shards.values():
            shard._flush()
            if shard._table.n == 0:
                continue
            for name in shard._names:                            # edge-only ids: indexed after every resident node
                index.setdefault(name, len(index))
            plan.append((shard, total))
            total += shard._table.n
        names = list(index)
        if not names:
            return []
        if total == 0:
            return [{nid} for nid in names[:n_resident]]
        dev = plan[0][0]._table.device
        gsrc = torch.empty(total, dtype=torch.int32, device=dev)
        gdst = torch.empty(total, dtype=torch.int32, device=dev)
        walk = torch.empty(total, dtype=torch.uint8, device=dev)
        for shard, off in plan:
            t = shard._table
            slot_global = torch.tensor([index[nm] for nm in shard._names], dtype=torch.int32).to(dev, non_blocking=True)
            resident = torch.tensor([1 if nm in shard.nodes else 0 for nm in shard._names], dtype=torch.uint8).to(dev, non_blocking=True)
            kernels.edge_endpoints_to_global(t.src, t.dst, t.weight, t.n, slot_global, resident,
                                             gsrc[off:off + t.n], gdst[off:off + t.n], walk[off:off + t.n], min_weight=0.0)
        labels = kernels.reach_components(gsrc, gdst, walk, total, len(names)).cpu().numpy()
        groups: Dict[int, Set[str]] = {}
        for i, lab in enumerate(labels.tolist()):
            if i < n_resident or lab != i:                        # an edge-only id nobody reaches is not a node
                groups.setdefault(lab, set()).add(names[i])
        return [groups[k] for k in sorted(groups)]


class MemorySystem:
    def __init__(self, llm_provider=None, embedding_provider=None, store: Optional[B200Store] = None,
                 user_id: str = "default", enable_sharding: bool = True, enable_hierarchy: bool = True,
                 super_node_threshold: int = 20, auto_consolidate: bool = True, consolidate_every: int = 3,
                 auto_prune: bool = True, prune_threshold: float = 0.5, max_buffer_size: int = 10,
                 metric: str = "cosine", device: Optional[str] = None, load_from_disk: bool = False, **_ignored):
        if llm_provider is None or embedding_provider is None:
            raise ValueError("lazzaro_b200.MemorySystem needs llm_provider and embedding_provider (no network here)")
        self.llm, self.embedder, self.user_id = llm_provider, embedding_provider, user_id
        self.store = store if store is not None else B200Store(metric=metric, device=device)
        self.vector_store = self.store                    # memory_system.py:110
        self.shards: Dict[str, MemoryShard] = {}
        self.super_nodes: Dict[str, Node] = {}
        self.buffer = BufferGraph(self.shards, self.super_nodes)
        self.buffer._device = str(self.store.device)
        self.enable_sharding, self.enable_hierarchy = enable_sharding, enable_hierarchy
        self.super_node_threshold = super_node_threshold
        self.auto_consolidate, self.consolidate_every = auto_consolidate, consolidate_every
        self.auto_prune, self.prune_threshold, self.max_buffer_size = auto_prune, prune_threshold, max_buffer_size
        self.conversation_active = False
        self.short_term_memory: List[Dict] = []
        self.consolidation_queue: List[Dict] = []
        self.node_counter = 0
        self.conversation_count = 0
        self.metrics = {"embedding_calls": 0, "llm_calls": 0, "retrieval_times": [], "consolidation_times": []}
        self.profile: Optional[dict] = None
        self._last_nodes_version = 0
        if load_from_disk:                                # memory_system.py:131-132
            self._load_from_persistence()

    # ------------------------------------------------------------------ small helpers
    def _generate_node_id(self) -> str:
        self.node_counter += 1
        return f"node_{self.node_counter}"

    def _infer_shard_key(self, content: str) -> str:      # memory_system.py:152-169
        if not self.enable_sharding:
            return "default"
        low = content.lower()
        for key, words in _TOPIC_KEYWORDS.items():
            if any(w in low for w in words):
                return key
        return time.strftime("%Y-%m")

    def _get_or_create_shard(self, key: str) -> MemoryShard:
        if key not in self.shards:
            self.shards[key] = MemoryShard(key, device=str(self.store.device))
        return self.shards[key]

    def _get_embedding(self, text: str) -> List[float]:
        self.metrics["embedding_calls"] += 1
        return self.embedder.embed(text)

    def _batch_embed(self, texts: List[str]) -> List[List[float]]:
        if not texts:
            return []
        self.metrics["embedding_calls"] += 1
        return self.embedder.batch_embed(texts)

    def _cosine_similarity(self, v1, v2) -> float:
        """memory_system.py:197-203, one pair, on the device (lz_score_pairs)."""
        if v1 is None or v2 is None or len(v1) == 0 or len(v2) == 0:
            return 0.0
        return float(self._cosine_pairs(np.asarray([v1], np.float32), np.asarray([v2], np.float32))[0])

    def _cosine_pairs(self, a: np.ndarray, b: np.ndarray) -> np.ndarray:
        """cos(a[i], b[i]) in float64 for float32 arrays [n, d]."""
        n, d = a.shape
        dp = (d + 7) // 8 * 8
        dev = self.store.device
        rows = torch.zeros((n, dp), dtype=torch.float32, device=dev)
        qs = torch.zeros((n, dp), dtype=torch.float32, device=dev)
        rows[:, :d] = torch.from_numpy(np.ascontiguousarray(b)).to(dev)
        qs[:, :d] = torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        norms = kernels.row_norms(rows)
        idx = torch.arange(n, device=dev)
        return kernels.score_pairs(rows, dp, norms, qs, idx.to(torch.int32), idx.to(torch.int64), _lib.COSINE).cpu().numpy()

    # ------------------------------------------------------------------ conversation life cycle
    def start_conversation(self) -> str:
        self.conversation_active = True
        self.short_term_memory = []
        return "✓ Conversation started"

    def add_to_short_term(self, content: str, memory_type: str = "semantic", salience: float = 0.5):
        if not self.conversation_active:
            raise RuntimeError("No active conversation")
        self.short_term_memory.append({"content": content, "type": memory_type, "salience": salience,
                                       "timestamp": time.time()})

    def end_conversation(self) -> str:
        """memory_system.py:580-649: consolidate, decay, prune, enforce the buffer limit."""
        if not self.conversation_active:
            return "⚠ No active conversation to end."
        self.conversation_active = False
        if not self.short_term_memory:
            return "✓ Conversation ended. No memories to consolidate."
        results = [self._consolidate_to_buffer()]
        if self.auto_prune:                               # decay + prune as one fused sweep per shard
            pruned = self.buffer.decay_and_prune(0.01, self.prune_threshold)
            results.append("✓ Applied temporal decay")
            if pruned > 0:
                results.append(f"✓ Auto-pruned {pruned} weak edges")
        else:
            self.buffer.apply_temporal_decay(decay_rate=0.01)
            results.append("✓ Applied temporal decay")
        self._enforce_buffer_limit()
        self.conversation_count += 1
        if self.auto_consolidate and self.conversation_count % self.consolidate_every == 0:
            results.append(self.run_consolidation())
        self.short_term_memory = []
        self._save_to_persistence()
        return "\n".join(results)

    def _consolidate_to_buffer(self) -> str:
        self.consolidation_queue.append({"memories": list(self.short_term_memory), "timestamp": time.time()})
        self._async_consolidate()
        n, e = self.buffer.size()
        return f"✓ Consolidation complete. Memory: {n} nodes, {e} edges"

    # ------------------------------------------------------------------ consolidation (memory_system.py:651-785)
    def _extract_memories(self, exchanges: List[Dict]) -> List[Dict]:
        self.metrics["llm_calls"] += 1
        messages = [{"role": "system", "content": "Extract distinct, atomic facts from this conversation. Return JSON: "
                                                  "{\"memories\": [{\"content\", \"type\", \"salience\", \"topic\"}]}"},
                    {"role": "user", "content": json.dumps(exchanges)}]
        response = self.llm.completion(messages, response_format={"type": "json_object"})
        try:
            if "```json" in response:
                response = response.split("```json")[1].split("```")[0].strip()
            data = json.loads(response)
        except (json.JSONDecodeError, TypeError):
            return []
        mems = data.get("memories", []) if isinstance(data, dict) else data if isinstance(data, list) else []
        return [m for m in mems if isinstance(m, dict)]

    def _async_consolidate(self):
        if not self.consolidation_queue:
            return
        t0 = time.time()
        exchanges = [m for batch in self.consolidation_queue for m in batch["memories"]]
        self.consolidation_queue.clear()
        memories = self._extract_memories(exchanges)
        contents = [m.get("content", "") for m in memories if m.get("content")]
        embeddings = self._batch_embed(contents)
        # the reference indexes embeddings by position in `memories`; kept as is
        cand = []
        for i, mem in enumerate(memories):
            content = mem.get("content", "")
            if not content or len(content) < 5:
                continue
            cand.append((mem, content, embeddings[i] if i < len(embeddings) else []))
        # dedup: one batched top-1 + cosine gate instead of one search per memory (memory_system.py:719-742);
        # memories of one batch never see each other (rows are added after the loop, 768-769)
        with_emb = [j for j, c in enumerate(cand) if len(c[2])]
        match: Dict[int, str] = {}
        if with_emb and self.store.n_rows:
            names, sims, dup = consolidate.dedup_against_store(self.store, np.asarray([cand[j][2] for j in with_emb], np.float32),
                                                               self.user_id)
            for j, name, d in zip(with_emb, names, dup):
                if d and name is not None and self.buffer.get_node(name) is not None:
                    match[j] = name
        new_nodes: List[Tuple[str, str]] = []
        rows_to_add = []
        for j, (mem, content, emb) in enumerate(cand):
            shard_key = mem.get("topic", self._infer_shard_key(content))
            shard = self._get_or_create_shard(shard_key)
            if j in match:                                # memory_system.py:735-742
                node = self.buffer.get_node(match[j])
                node.salience = max(node.salience, mem.get("salience", 0.5))
                node.last_accessed = time.time()
                node.access_count += 1
                continue
            node = Node(id=self._generate_node_id(), content=content, embedding=list(emb), type=mem.get("type", "semantic"),
                        salience=mem.get("salience", 0.5), shard_key=shard_key)
            shard.add_node(node)
            new_nodes.append((node.id, shard_key))
            rows_to_add.append({"id": node.id, "content": content, "embedding": list(emb), "type": node.type,
                                "salience": node.salience, "shard_key": shard_key, "timestamp": node.timestamp})
        # links to existing memories are searched BEFORE the new rows enter the store: the reference adds
        # them first but then excludes them from the scan (memory_system.py:842-847), which is the same set
        cross = self._cross_link_candidates(new_nodes)
        if rows_to_add:
            self.store.add_nodes(rows_to_add, user_id=self.user_id)
        self._link_within_shards(new_nodes)
        self._apply_cross_links(new_nodes, cross)
        self._enforce_buffer_limit()
        if self.enable_hierarchy:
            for key in {k for _, k in new_nodes}:
                if key in self.shards and len(self.shards[key].nodes) > self.super_node_threshold:
                    self._create_super_nodes_for_shard(key)
        self.metrics["consolidation_times"].append(time.time() - t0)
        self._save_to_persistence()

    def _save_to_persistence(self):
        """memory_system.py:1275-1302: the user's rows and edges in the store are replaced by the whole
        in-memory graph (super-nodes first, then the shards in insertion order), so after every conversation
        the store holds exactly the resident nodes — including the super-nodes and their centroids."""
        nodes = [{"id": n.id, "content": n.content, "embedding": list(n.embedding), "type": n.type,
                  "timestamp": n.timestamp, "access_count": n.access_count, "last_accessed": n.last_accessed,
                  "salience": n.salience, "is_super_node": n.is_super_node, "child_ids": list(n.child_ids),
                  "parent_id": n.parent_id, "shard_key": n.shard_key}
                 for n in self.buffer.nodes.values()]
        self.store.delete_nodes([], user_id=self.user_id)
        self.store.delete_edges(user_id=self.user_id)
        dims = {len(n["embedding"]) for n in nodes}
        if nodes and len(dims) == 1 and 0 not in dims:
            self.store.add_nodes(nodes, user_id=self.user_id)
        edges = [{"source": e.source, "target": e.target, "weight": e.weight, "edge_type": e.edge_type,
                  "co_occurrence": e.co_occurrence, "last_updated": e.last_updated}
                 for shard in self.shards.values() for e in shard.edges.values()]
        if edges:
            self.store.add_edges(edges, user_id=self.user_id)

    def _load_from_persistence(self):
        """memory_system.py:1304-1410: rebuild shards, super-nodes, edges and the id counter from the store's rows
        of this user.  Nodes keep the store's row order; an edge goes to the shard its source node (or super-node)
        names; a shard's edges are upserted in one device batch (same outcome as the reference's add_edge loop)."""
        node_dicts = self.store.get_nodes(user_id=self.user_id)
        if not node_dicts:
            return
        self.shards, self.super_nodes = {}, {}
        fields = Node.__dataclass_fields__
        for nd in node_dicts:
            nd = dict(nd)
            if "vector" in nd:
                nd["embedding"] = nd.pop("vector")
            node = Node(**{k: v for k, v in nd.items() if k in fields})
            if node.is_super_node:
                self.super_nodes[node.id] = node
            else:
                if node.shard_key not in self.shards:
                    self.shards[node.shard_key] = MemoryShard(node.shard_key, device=str(self.store.device))
                self.shards[node.shard_key].add_node(node)
        per_shard: Dict[str, List[Edge]] = defaultdict(list)
        efields = Edge.__dataclass_fields__
        edge_dicts = self.store.get_edges(user_id=self.user_id)
        for ed in edge_dicts:
            ed = dict(ed)
            if "source_id" in ed:
                ed["source"] = ed.pop("source_id")
            if "target_id" in ed:
                ed["target"] = ed.pop("target_id")
            if "type" in ed and "edge_type" not in ed:
                ed["edge_type"] = ed.pop("type")
            edge = Edge(**{k: v for k, v in ed.items() if k in efields})
            src = self.super_nodes.get(edge.source)
            if src is None:
                for shard in self.shards.values():
                    if edge.source in shard.nodes:
                        src = shard.nodes[edge.source]
                        break
            if src is not None and src.shard_key in self.shards:
                per_shard[src.shard_key].append(edge)
        for key, batch in per_shard.items():
            self.shards[key].add_edges(batch)
        self.profile = self.store.load_profile(user_id=self.user_id)
        self._last_nodes_version = self.store.get_latest_version()
        self.buffer = BufferGraph(self.shards, self.super_nodes)
        self.buffer._device = str(self.store.device)
        max_id = 0
        for nid in self.buffer.nodes:
            if nid.startswith("node_"):
                try:
                    max_id = max(max_id, int(nid.split("_")[1]))
                except ValueError:
                    pass
        if self.buffer.nodes:
            self.node_counter = max_id

    def check_for_updates(self) -> bool:
        """memory_system.py:1412-1428."""
        current = self.store.get_latest_version()
        if current > self._last_nodes_version:
            self._load_from_persistence()
            return True
        return False

    def get_all_users(self) -> List[str]:
        """memory_system.py:1430-1458 (the reference reaches into the store's node table; here the store answers)."""
        table = getattr(self.store, "_nodes_table", None)
        if table is None:
            return [self.user_id]
        return sorted({r["user_id"] for r in table.search().select(["user_id"]).to_list()})

    def switch_user(self, new_user_id: str):
        """memory_system.py:1474-1486."""
        if self.conversation_active:
            self.end_conversation()
        else:
            self._save_to_persistence()
        self.user_id = new_user_id
        self._load_from_persistence()      # as in the reference, a user with no saved rows leaves the resident graph as it is

    def _link_within_shards(self, new_nodes: List[Tuple[str, str]]):
        """memory_system.py:797-836: chain edges (0.5) between consecutive new nodes of a shard, then for
        each new node the 3 most similar older nodes of the shard above 0.5 (weight 0.8 * cosine).  The
        cosines of all (new, older) pairs of a shard are one lz_score_pairs launch."""
        groups: Dict[str, List[str]] = defaultdict(list)
        for nid, key in new_nodes:
            groups[key].append(nid)
        for key, ids in groups.items():
            if len(ids) < 2:
                continue
            shard = self.shards[key]
            batch = [Edge(source=a, target=b, weight=0.5, edge_type="relates_to") for a, b in zip(ids[:-1], ids[1:])]
            new_set = set(ids)
            older = [(oid, n) for oid, n in shard.nodes.items() if oid not in new_set]
            dims = {len(n.embedding) for _, n in older} | {len(shard.nodes[i].embedding) for i in ids}
            if not older or len(dims) != 1 or 0 in dims:
                shard.add_edges(batch)
                continue
            a = np.repeat(np.asarray([shard.nodes[i].embedding for i in ids], np.float32), len(older), axis=0)
            b = np.tile(np.asarray([n.embedding for _, n in older], np.float32), (len(ids), 1))
            sims = self._cosine_pairs(a, b).reshape(len(ids), len(older))
            for i, nid in enumerate(ids):
                order = np.argsort(-sims[i], kind="stable")[:3]         # stable: ties by insertion order
                for j in order:
                    if sims[i, j] > 0.5:
                        batch.append(Edge(source=nid, target=older[j][0], weight=float(sims[i, j]) * 0.8,
                                          edge_type="relates_to"))
            shard.add_edges(batch)                      # one device lookup + upsert for the shard, same outcome as in order

    def _cross_link_candidates(self, new_nodes):
        """memory_system.py:838-867: per new node the 3 best existing rows above 0.5 (one batched search)."""
        if not new_nodes or self.store.n_rows == 0:
            return {}
        embs, ids = [], []
        for nid, _ in new_nodes:
            node = self.buffer.get_node(nid)
            if node is not None and len(node.embedding):
                embs.append(node.embedding)
                ids.append(nid)
        if not embs:
            return {}
        # The reference ranks only resident non-super nodes (memory_system.py:842-847) and takes the best 3 of THOSE;
        # the store also holds the super-nodes' centroid rows (written by _save_to_persistence), which are often among
        # a new memory's nearest rows.  So ask for 3 + (number of super-nodes) candidates, drop super-nodes and rows
        # that are not resident, and keep the first 3 — the same three the reference finds.
        m = min(_lib.MAX_K, consolidate.LINK_TOP + len(self.super_nodes))
        links = consolidate.link_candidates(self.store, np.asarray(embs, np.float32), self.user_id, m=m)
        out = {}
        for nid, lst in zip(ids, links):
            keep = []
            for cand in lst:
                node = self.buffer.get_node(cand[0])
                if node is not None and not node.is_super_node:
                    keep.append(cand)
                    if len(keep) == consolidate.LINK_TOP:
                        break
            out[nid] = keep
        return out

    def _apply_cross_links(self, new_nodes, cross):
        """memory_system.py:866-889: skip a link whose edge exists in either direction in any shard; the
        edge goes to the new node's shard."""
        cand = []
        for nid, key in new_nodes:
            for target, sim, weight in cross.get(nid, []):
                node = self.buffer.get_node(target)
                if node is not None and not node.is_super_node:
                    cand.append((nid, key, target, weight))
        if not cand:
            return
        # the existence checks of all candidates in one device lookup per shard; edges added inside this loop
        # are tracked in `added`, so the outcome is that of the reference's one-by-one loop
        pairs = [(nid, target) for nid, _, target, _ in cand] + [(target, nid) for nid, _, target, _ in cand]
        exists = [False] * len(cand)
        for s in self.shards.values():
            got = s.has_edges(pairs)
            for i in range(len(cand)):
                exists[i] = exists[i] or got[i] or got[len(cand) + i]
        added = set()
        per_shard: Dict[str, List[Edge]] = defaultdict(list)
        for i, (nid, key, target, weight) in enumerate(cand):
            if exists[i] or (nid, target) in added or (target, nid) in added:
                continue
            if key in self.shards:
                per_shard[key].append(Edge(source=nid, target=target, weight=weight, edge_type="relates_to"))
                added.add((nid, target))
        for key, batch in per_shard.items():
            self.shards[key].add_edges(batch)

    def _link_to_existing_memories(self, new_nodes):
        """Reference-shaped entry point (memory_system.py:838-891) for nodes already in the store: the scan
        excludes the new nodes themselves by asking for 3 + len(new) candidates."""
        if not new_nodes or self.store.n_rows == 0:
            return
        new_set = {nid for nid, _ in new_nodes}
        embs = [self.buffer.get_node(nid).embedding for nid, _ in new_nodes]
        k = min(_lib.MAX_K, 3 + len(new_set) + len(self.super_nodes))
        links = consolidate.link_candidates(self.store, np.asarray(embs, np.float32), self.user_id, m=k)
        keep = lambda c: c[0] not in new_set and (lambda n: n is not None and not n.is_super_node)(self.buffer.get_node(c[0]))
        cross = {nid: [l for l in lst if keep(l)][:3] for (nid, _), lst in zip(new_nodes, links)}
        self._apply_cross_links(new_nodes, cross)

    def _create_super_nodes_for_shard(self, shard_key: str):
        """memory_system.py:893-933: one super-node per shard, embedding = mean of the members (lz_segment_mean)."""
        shard = self.shards[shard_key]
        if len(shard.nodes) < self.super_node_threshold:
            return
        if any(n.shard_key == shard_key for n in self.super_nodes.values()):
            return
        nodes = list(shard.nodes.values())
        embs = [n.embedding for n in nodes if len(n.embedding)]
        centroid: List[float] = []
        if embs:
            d = len(embs[0])
            dp = (d + 7) // 8 * 8
            rows = torch.zeros((len(embs), dp), dtype=torch.float32, device=self.store.device)
            rows[:, :d] = torch.from_numpy(np.asarray(embs, np.float32)).to(self.store.device)
            seg = torch.tensor([0, len(embs)], dtype=torch.int64, device=self.store.device)
            centroid = kernels.segment_mean(rows, dp, seg)[0, :d].cpu().tolist()
        sid = f"super_{shard_key}_{int(time.time())}"
        self.super_nodes[sid] = Node(id=sid, content=f"Topic: {shard_key}. Contains memories about: "
                                     + "; ".join(n.content for n in nodes[:3]), embedding=centroid, type="semantic",
                                     is_super_node=True, child_ids=[n.id for n in nodes], shard_key=shard_key)
        for n in nodes:
            n.parent_id = sid

    def _enforce_buffer_limit(self):
        """memory_system.py:535-578: evict the least important non-super nodes beyond max_buffer_size.
        Importance scoring and the lowest-`excess` selection run on the device (lz_evict_lowest, in rounds of
        LZ_MAX_K); node fields are gathered in the reference's iteration order (shards, then nodes, in dict
        order) so index order = its stable-sort order.  One `now` for the sweep (the reference reads the clock
        per node)."""
        total, _ = self.buffer.size()
        if total <= self.max_buffer_size:
            return
        excess = total - self.max_buffer_size
        now = time.time()
        order = [n for shard in self.shards.values() for n in shard.nodes.values()]
        if not order:
            return
        dev = self.store.device
        sal = torch.tensor([n.salience for n in order], dtype=torch.float64, device=dev)
        acc = torch.tensor([n.access_count for n in order], dtype=torch.int32, device=dev)
        last = torch.tensor([n.last_accessed for n in order], dtype=torch.float64, device=dev)
        skip = torch.tensor([1 if n.is_super_node else 0 for n in order], dtype=torch.uint8, device=dev)
        chosen: List[int] = []
        while len(chosen) < excess:
            idx, _ = kernels.evict_lowest(sal, acc, last, now, min(excess - len(chosen), _lib.MAX_K), skip=skip)
            if idx.numel() == 0:
                break
            skip[idx] = 1
            chosen.extend(idx.tolist())
        removed = []
        for i in chosen:
            n = order[i]
            shard = self.shards.get(n.shard_key)
            if shard is not None and n.id in shard.nodes:
                shard.remove_node(n.id)
                removed.append(n.id)
        if removed:
            self.store.delete_nodes([order[i].id for i in chosen], user_id=self.user_id)

    def _merge_similar_nodes(self, similarity_threshold: float = 0.95) -> int:
        """memory_system.py:1065-1120.  As written in the reference the pair loop sits outside the outer
        loop and never runs, so the function returns 0 and changes nothing; parity means the
        same here.  The intended all-pairs dedup is available as consolidate.dedup_against_store."""
        return 0

    def run_consolidation(self, weight_threshold: float = 0.6, merge_similar: bool = True) -> str:
        """memory_system.py:935-1010 without the LLM profile extraction (out of scope)."""
        results = []
        if merge_similar:
            merged = self._merge_similar_nodes(0.95)
            if merged > 0:
                results.append(f"✓ Merged {merged} similar nodes")
        self.last_components = [c for c in self.buffer.get_connected_components() if len(c) >= 3]
        pruned = self.buffer.prune_weak_edges(threshold=self.prune_threshold)
        if pruned > 0:
            results.append(f"✓ Pruned {pruned} weak edges")
        return "\n".join(results) if results else "✓ No consolidation actions needed"

    # ------------------------------------------------------------------ retrieval
    def search_memories(self, query: str, limit: int = 5) -> List[Node]:
        """memory_system.py:1460-1472: embed, search the store, map ids to resident nodes (missing ones are dropped)."""
        ids = self.store.search_nodes(self._get_embedding(query), user_id=self.user_id, limit=limit)
        return [n for n in (self.buffer.get_node(i) for i in ids) if n is not None]

    def _optimized_retrieval(self, query_emb: List[float], query_text: str) -> List[str]:
        """memory_system.py:453-514."""
        retrieved: List[str] = []
        if self.enable_hierarchy and self.super_nodes:
            supers = [(sid, n) for sid, n in self.super_nodes.items() if len(n.embedding) == len(query_emb)]
            if supers:
                q = np.tile(np.asarray([query_emb], np.float32), (len(supers), 1))
                sims = self._cosine_pairs(q, np.asarray([n.embedding for _, n in supers], np.float32))
                best = int(np.argsort(-sims, kind="stable")[0])
                if sims[best] > 0.4:
                    for cid in supers[best][1].child_ids[:10]:
                        child = self.buffer.get_node(cid)
                        if child and not child.is_super_node:
                            retrieved.append(cid)
                    if len(retrieved) >= 5:
                        return retrieved[:5]
        vector_ids = self.store.search_nodes(query_emb, user_id=self.user_id, limit=10 if not retrieved else 5)
        seen, seen_content, out = set(retrieved), set(), []
        for rid in retrieved:
            node = self.buffer.get_node(rid)
            if node:
                seen_content.add(node.content)
                out.append(rid)
        for rid in vector_ids:
            if rid in seen:
                continue
            node = self.buffer.get_node(rid)
            if node and node.content not in seen_content:
                seen_content.add(node.content)
                out.append(rid)
                seen.add(rid)
        return out[:5]

    def save_state(self, filename: str = "memory_state.json") -> str:
        """memory_system.py:1216-1239 (same JSON layout; lazzaro_b200/persist.py)."""
        from . import persist
        return persist.save_state(self, filename)

    def load_state(self, filename: str = "memory_state.json") -> str:
        """memory_system.py:1241-1273."""
        from . import persist
        return persist.load_state(self, filename)

    def close(self):
        self.store.close()
