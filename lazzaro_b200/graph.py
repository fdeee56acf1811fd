"""Device-resident graph containers: the decay / prune side of the hot path.

Host-side mirror of /root/reference/src/lazzaro/core/memory_shard.py (`MemoryShard`) and
core/buffer_graph.py (`BufferGraph.apply_temporal_decay / prune_weak_edges`, lines 87-97), with the
same method names and return values, over a struct-of-arrays edge table in HBM:

    src int32 | dst int32 | weight float32 | co_occurrence int32        (16 B per edge)

Strings (node ids, contents) stay on the host; edges refer to nodes by a per-shard integer slot.
`apply_temporal_decay` and `prune_weak_edges` are single kernel sweeps (lz_edge_decay,
lz_salience_decay, lz_edge_decay_prune); the fused `decay_and_prune` does both in one pass, which is
what `MemorySystem.end_conversation` does back to back (core/memory_system.py:624-630).
State is float32 (the reference persists weight and salience as float32, vector_store.py:42,63);
each sweep computes in float64 exactly as the reference's Python floats do and stores the rounding.
"""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np
import torch

from . import _lib, kernels


@dataclass
class Node:
    """Same fields as the reference's models/graph.py:39-50."""
    id: str
    content: str
    embedding: List[float] = field(default_factory=list)
    type: str = "semantic"
    timestamp: float = field(default_factory=time.time)
    access_count: int = 0
    last_accessed: float = field(default_factory=time.time)
    salience: float = 0.5
    is_super_node: bool = False
    child_ids: List[str] = field(default_factory=list)
    parent_id: Optional[str] = None
    shard_key: str = "default"


@dataclass
class Edge:
    """Same fields as the reference's models/graph.py:90-95."""
    source: str
    target: str
    weight: float = 1.0
    edge_type: str = "relates_to"
    co_occurrence: int = 1
    last_updated: float = field(default_factory=time.time)


class EdgeTable:
    """Edge struct-of-arrays in HBM with append, decay, fused decay+prune."""

    def __init__(self, device=None, capacity: int = 1024, in_place: bool = False):
        _lib.require_cuda()
        _lib.load()
        self._in_place = in_place
        self._alt = None
        self.device = torch.device(device or f"cuda:{torch.cuda.current_device()}")
        self.n = 0
        self._alloc(max(16, capacity))
        self._scratch: Optional[torch.Tensor] = None
        self._kept = torch.zeros(1, dtype=torch.int64, device=self.device)
        self._kept_host = torch.zeros(1, dtype=torch.int64, pin_memory=True)

    def _alloc(self, cap: int):
        new = [torch.empty(cap, dtype=dt, device=self.device)
               for dt in (torch.int32, torch.int32, torch.float32, torch.int32)]
        if self.n:
            for a, b in zip(new, (self.src, self.dst, self.weight, self.cooc)):
                a[:self.n] = b[:self.n]
        self.src, self.dst, self.weight, self.cooc = new
        self._alt = None

    @classmethod
    def from_arrays(cls, src, dst, weight, cooc, device=None) -> "EdgeTable":
        t = cls(device, capacity=len(src))
        t.append(src, dst, weight, cooc)
        return t

    def append(self, src, dst, weight, cooc):
        k = len(src)
        if self.n + k > self.src.shape[0]:
            self._alloc(max(self.n + k, 2 * self.src.shape[0]))
        for col, vals, dt in ((self.src, src, torch.int32), (self.dst, dst, torch.int32),
                              (self.weight, weight, torch.float32), (self.cooc, cooc, torch.int32)):
            col[self.n:self.n + k] = torch.as_tensor(vals, dtype=dt).to(self.device, non_blocking=True)
        self.n += k

    def decay(self, factor: float):
        """weight *= factor (memory_shard.py:69-70), one elementwise sweep."""
        kernels.edge_decay(self.weight, self.n, factor)

    def decay_prune(self, factor: float, threshold: float, sync: bool = True) -> int:
        """Fused sweep: decay by `factor`, drop weight < threshold (strict), compact in place keeping
        order.  Returns the number pruned (memory_shard.py:79-84) — this reads one int64 back."""
        n0 = self.n
        if n0 == 0:
            return 0
        lib = _lib.load()
        with torch.cuda.device(self.device):
            need = int(lib.lz_edge_decay_prune_scratch_bytes(n0))
            if self._scratch is None or self._scratch.numel() < need:
                self._scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
            # out of place into a second set of columns (then swap): the streamed kernel then counts survivors ahead of
            # its data path (0.85 of the HBM copy peak); `in_place=True` at construction keeps one copy of the table: a tile
            # then also waits until the (at most two) tiles its survivors land on are on chip (0.73 of the copy peak)
            if self._in_place:
                out = None
            else:
                if self._alt is None or self._alt[0].shape[0] < self.src.shape[0]:
                    self._alt = [torch.empty_like(c) for c in (self.src, self.dst, self.weight, self.cooc)]
                out = tuple(self._alt)
            kernels.edge_decay_prune(self.src, self.dst, self.weight, self.cooc, n0, factor, threshold,
                                     out=out, scratch=self._scratch, kept=self._kept)
            if out is not None:
                self._alt = [self.src, self.dst, self.weight, self.cooc]
                self.src, self.dst, self.weight, self.cooc = out
            if not sync:
                return -1
            self._kept_host.copy_(self._kept, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        self.n = int(self._kept_host[0])
        return n0 - self.n

    def to_numpy(self):
        n = self.n
        return (self.src[:n].cpu().numpy(), self.dst[:n].cpu().numpy(),
                self.weight[:n].cpu().numpy(), self.cooc[:n].cpu().numpy())


class MemoryShard:
    """Mirror of the reference's MemoryShard (core/memory_shard.py:7-88) over a device edge table."""

    def __init__(self, shard_key: str, device=None):
        self.shard_key = shard_key
        self.nodes: Dict[str, Node] = {}
        self.last_accessed = time.time()
        self.access_count = 0
        self._slot: Dict[str, int] = {}              # node id -> integer slot (also for edge-only ids)
        self._names: List[str] = []
        self._table = EdgeTable(device)
        self._edge_meta: Dict[Tuple[int, int], Tuple[str, float]] = {}   # (edge_type, last_updated)
        self._pending_keys: Dict[Tuple[int, int], int] = {}             # (src,dst) of the not-yet-uploaded edges
        self._pending: List[Tuple[int, int, float, int]] = []            # edges not yet uploaded
        self._csr = None                                                  # (offsets, adj, n_nodes): adjacency index of the table

    # -- nodes
    def _slot_of(self, node_id: str) -> int:
        s = self._slot.get(node_id)
        if s is None:
            s = self._slot[node_id] = len(self._names)
            self._names.append(node_id)
        return s

    def add_node(self, node: Node) -> None:
        node.shard_key = self.shard_key                   # memory_shard.py:39
        self.nodes[node.id] = node
        self._slot_of(node.id)

    # -- edges
    def _flush(self):
        if self._pending:
            s, d, w, c = zip(*self._pending)
            self._table.append(np.int32(s), np.int32(d), np.float32(w), np.int32(c))
            self._pending.clear()
            self._pending_keys.clear()
            self._csr = None

    def _find_rows(self, keys: List[Tuple[int, int]]) -> List[int]:
        """Row of each (src slot, dst slot) key in the UPLOADED edge table, or -1: one lz_edge_find launch on the
        table's CSR (edges still pending on the host are not in it; callers check `_pending_keys` first)."""
        if not keys or self._table.n == 0:
            return [-1] * len(keys)
        if self._csr is None:
            off, adj = kernels.csr_build(self._table.src, self._table.dst, self._table.n, len(self._names))
            self._csr = (off, adj, len(self._names))
        off, adj, n_nodes = self._csr                  # slots created after the build have no uploaded edges: -1
        dev = self._table.device
        qs = torch.tensor([k[0] for k in keys], dtype=torch.int32, device=dev)
        qd = torch.tensor([k[1] for k in keys], dtype=torch.int32, device=dev)
        return kernels.edge_find(off, adj, self._table.dst, n_nodes, qs, qd).cpu().tolist()

    def _reinforce(self, rows: List[int], times: List[int]) -> None:
        dev = self._table.device
        kernels.edge_reinforce(self._table.weight, self._table.cooc, self._table.n,
                               torch.tensor(rows, dtype=torch.int32, device=dev),
                               torch.tensor(times, dtype=torch.int32, device=dev))

    def add_edges(self, edges: List[Edge]) -> None:
        """A run of add_edge calls (memory_shard.py:42-52) as one batch with the same outcome as calling them
        in order: the first occurrence of a key that is not in the table inserts it, every other occurrence
        reinforces (+0.1 capped at 1, co_occurrence += 1).  Lookup and reinforcement run on the device
        (lz_edge_find on the CSR, lz_edge_reinforce)."""
        if not edges:
            return
        self._flush()
        first: Dict[Tuple[int, int], Edge] = {}
        count: Dict[Tuple[int, int], int] = {}
        for e in edges:
            key = (self._slot_of(e.source), self._slot_of(e.target))
            if key not in first:
                first[key] = e
            count[key] = count.get(key, 0) + 1
        keys = list(first)
        rows = self._find_rows(keys)
        hit_rows, hit_times, fresh = [], [], []
        for key, row in zip(keys, rows):
            if row >= 0:
                hit_rows.append(row); hit_times.append(count[key])
            else:
                fresh.append(key)
        base = self._table.n
        for key in fresh:
            e = first[key]
            self._pending.append((key[0], key[1], float(e.weight), int(e.co_occurrence)))
            self._edge_meta[key] = (e.edge_type, e.last_updated)
        self._flush()
        for j, key in enumerate(fresh):
            if count[key] > 1:
                hit_rows.append(base + j); hit_times.append(count[key] - 1)
        if hit_rows:
            self._reinforce(hit_rows, hit_times)

    def add_edge(self, edge: Edge) -> None:
        """memory_shard.py:42-52: an existing (source, target) key is reinforced (+0.1 capped at 1,
        co_occurrence += 1), otherwise the edge is appended (uploaded with the next sweep or lookup batch)."""
        key = (self._slot_of(edge.source), self._slot_of(edge.target))
        if key in self._pending_keys:
            self._flush()                                # the edge to reinforce must be in the device table
        row = self._find_rows([key])[0]
        if row >= 0:
            self._reinforce([row], [1])
        else:
            self._pending_keys[key] = len(self._pending)
            self._pending.append((key[0], key[1], float(edge.weight), int(edge.co_occurrence)))
            self._edge_meta[key] = (edge.edge_type, edge.last_updated)

    def has_edges(self, pairs: List[Tuple[str, str]]) -> List[bool]:
        """Whether each (source, target) key exists (directional, as the reference's dict lookup)."""
        keys = [(self._slot.get(a, -1), self._slot.get(b, -1)) for a, b in pairs]
        out = [k in self._pending_keys for k in keys]
        todo = [i for i, k in enumerate(keys) if not out[i] and k[0] >= 0 and k[1] >= 0]
        for i, row in zip(todo, self._find_rows([keys[i] for i in todo])):
            out[i] = row >= 0
        return out

    def has_edge(self, source: str, target: str) -> bool:
        return self.has_edges([(source, target)])[0]

    def remove_node(self, node_id: str) -> None:
        """Drops a node and every edge touching it (memory_system.py:559-571).  Host-side rebuild of the
        edge table: eviction is off the hot path."""
        self.nodes.pop(node_id, None)
        slot = self._slot.get(node_id)
        if slot is None:
            return
        self._flush()
        s, d, w, c = self._table.to_numpy()
        keep = (s != slot) & (d != slot)
        if not keep.all():
            self._table = EdgeTable.from_arrays(s[keep], d[keep], w[keep], c[keep], device=self._table.device) \
                if keep.any() else EdgeTable(self._table.device)
            self._csr = None

    @property
    def edges(self) -> Dict[Tuple[str, str], Edge]:
        """Materialises the edge dict from HBM (a device->host copy; not for hot loops)."""
        self._flush()
        s, d, w, c = self._table.to_numpy()
        out = {}
        for a, b, ww, cc in zip(s.tolist(), d.tolist(), w.tolist(), c.tolist()):
            et, lu = self._edge_meta.get((a, b), ("relates_to", 0.0))
            out[(self._names[a], self._names[b])] = Edge(self._names[a], self._names[b], ww, et, cc, lu)
        return out

    def _adjacency(self):
        """The CSR over the current edge table (lz_csr_build), rebuilt after inserts / prunes / removals;
        weights are read live by the lookup, so decay alone keeps it valid."""
        self._flush()
        n_nodes = len(self._names)
        if self._csr is None or self._csr[2] != n_nodes:
            off, adj = kernels.csr_build(self._table.src, self._table.dst, self._table.n, n_nodes)
            self._csr = (off, adj, n_nodes)
        return self._csr

    def get_neighbors_batch(self, node_ids: List[str], min_weight: float = 0.3) -> List[List[str]]:
        """get_neighbors for many nodes in one launch (one warp per node on the device CSR)."""
        if not node_ids:
            return []
        off, adj, n_nodes = self._adjacency()
        slots = torch.tensor([self._slot.get(i, -1) for i in node_ids], dtype=torch.int32, device=self._table.device)
        nbr, cnt = kernels.neighbors(off, adj, self._table.src, self._table.dst, self._table.weight, n_nodes, slots, min_weight)
        nbr, cnt = nbr.cpu().tolist(), cnt.cpu().tolist()
        return [[self._names[x] for x in row[:c]] for row, c in zip(nbr, cnt)]

    def get_neighbors(self, node_id: str, min_weight: float = 0.3) -> List[str]:
        """memory_shard.py:54-62: the other endpoints of the node's edges with weight >= min_weight, in edge
        order — a lookup on the device CSR (lz_neighbors) instead of the reference's walk over every edge."""
        return self.get_neighbors_batch([node_id], min_weight)[0]

    # -- the hot path
    def _salience_sweep(self, decay_rate: float):
        if not self.nodes:
            return
        nodes = list(self.nodes.values())
        s = torch.tensor([n.salience for n in nodes], dtype=torch.float32).to(self._table.device)
        kernels.salience_decay(s, s.numel(), decay_rate, 0.2)
        for n, v in zip(nodes, s.cpu().tolist()):
            n.salience = v

    def apply_temporal_decay(self, decay_rate: float = 0.01) -> None:
        """memory_shard.py:64-77."""
        self._flush()
        self._table.decay(1.0 - decay_rate)
        self._salience_sweep(decay_rate)

    def prune_weak_edges(self, threshold: float = 0.5) -> int:
        """memory_shard.py:79-84; returns the number of edges removed."""
        self._flush()
        pruned = self._table.decay_prune(1.0, threshold)
        if pruned:
            self._csr = None
        return pruned

    def decay_and_prune(self, decay_rate: float = 0.01, threshold: float = 0.5) -> int:
        """apply_temporal_decay followed by prune_weak_edges as ONE sweep over the edge table
        (core/memory_system.py:624-630 calls them back to back)."""
        self._flush()
        pruned = self._table.decay_prune(1.0 - decay_rate, threshold)
        self._salience_sweep(decay_rate)
        if pruned:
            self._csr = None
        return pruned

    def size(self) -> Tuple[int, int]:
        return len(self.nodes), self._table.n + len(self._pending)

    def connected_components(self) -> List[set]:
        """Components of this shard's graph (edges undirected, any weight), on the device
        (lz_connected_components).  For a shard whose edges stay inside the shard this is what
        BufferGraph.get_connected_components (core/buffer_graph.py:99-120) finds for its nodes."""
        self._flush()
        n = len(self._names)
        if n == 0:
            return []
        labels = kernels.connected_components(self._table.src, self._table.dst, self._table.n, n).cpu().numpy()
        groups: Dict[int, set] = {}
        for slot, lab in enumerate(labels.tolist()):
            groups.setdefault(lab, set()).add(self._names[slot])
        return [groups[k] for k in sorted(groups)]


class BufferGraph:
    """Decay / prune fan-out of the reference's BufferGraph (core/buffer_graph.py:87-97)."""

    def __init__(self, shards: Dict[str, MemoryShard], super_nodes: Optional[Dict[str, Node]] = None):
        self.shards = shards
        self.super_nodes = super_nodes if super_nodes is not None else {}
        self._device = None

    def apply_temporal_decay(self, decay_rate: float = 0.01):
        for shard in self.shards.values():
            shard.apply_temporal_decay(decay_rate)

    def prune_weak_edges(self, threshold: float = 0.5) -> int:
        return sum(shard.prune_weak_edges(threshold) for shard in self.shards.values())

    def decay_and_prune(self, decay_rate: float = 0.01, threshold: float = 0.5) -> int:
        return sum(shard.decay_and_prune(decay_rate, threshold) for shard in self.shards.values())

    def size(self) -> Tuple[int, int]:
        n = sum(s.size()[0] for s in self.shards.values()) + len(self.super_nodes)
        return n, sum(s.size()[1] for s in self.shards.values())
