"""B200Store — a device-resident drop-in for the reference's `LanceDBStore`.

Mirrors /root/reference/src/lazzaro/core/vector_store.py (the `Store` protocol of
core/interfaces.py:55-102, same method names, argument meaning and "return [] / None rather than
raise" behaviour) plus the legacy `add/search/delete` surface the reference's tests still call
(tests/test_vector_store.py:15,43,46,66).  It is injected the same way the reference allows:
`MemorySystem(store=B200Store(...))` (core/memory_system.py:83,104-107).

What lives where
  HBM   rows float32 [capacity, dim_p] (insertion order, zero-padded to a multiple of 8 columns),
        norms float64 [capacity], tags int32 [capacity] (user index, -1 = deleted row)
  host  everything that is string-valued: ids, user ids, contents, JSON blobs, edges, profiles —
        the reference keeps those in Arrow tables; they never touch the hot path.

search_nodes = one call of the exact flat-scan kernel (lz_search_exact) over the user's row span
with the tag pre-filter; there is no CPU path: without the CUDA library every search raises.
"""
from __future__ import annotations

import json
import threading
import time
from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib, kernels


def _round_up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


class _Shadow:
    """The fp16 copy of the rows that the tensor-core pass reads, for ONE metric (cosine: normalised rows; l2: raw
    rows + the norm-term columns), kept up to date lazily, with what the certification margin needs: the largest
    rounding-error norm of any shadow row (deterministic margin) and the largest row norm (L2 margins)."""

    def __init__(self, metric: int):
        self.metric = metric
        self.t: Optional[torch.Tensor] = None        # fp16 [capacity, dim_p (+ 8)]
        self.err: Optional[torch.Tensor] = None      # float32 [capacity] per-row rounding-error norm
        self.n = 0                                    # rows that are up to date
        self.max_err = 0.0
        self.max_norm = 0.0

    def reset(self):
        self.n, self.max_err, self.max_norm = 0, 0.0, 0.0

    def ensure(self, store: "B200Store"):
        cap = store._rows.shape[0]
        if self.t is None or self.t.shape[0] != cap:
            t = torch.empty((cap, kernels.shadow_dim(store._dim_p, self.metric)), dtype=torch.float16, device=store.device)
            err = torch.empty(cap, dtype=torch.float32, device=store.device)
            if self.t is not None and self.n:
                t[:self.n] = self.t[:self.n]
                err[:self.n] = self.err[:self.n]
            self.t, self.err = t, err
        if self.n < store._n:
            a, b = self.n, store._n
            kernels.rows_to_shadow(store._rows[a:b], store._dim_p, store._norms[a:b], out=self.t[a:b], metric=self.metric,
                                   row_err=self.err[a:b])
            # row side of the margins: the largest row norm (L2) and the largest rounding-error norm of any shadow row
            # (deleted rows included — upper bounds are all that is needed); one small read-back per refresh
            self.max_norm = max(self.max_norm, float(store._norms[a:b].max().item()))
            self.max_err = max(self.max_err, float(self.err[a:b].max().item()))
            self.n = b


class B200Store:
    def __init__(self, db_dir: str = "db", dim: Optional[int] = None, metric: str = "cosine",
                 device: Optional[str] = None, capacity: int = 4096, batched_search: str = "auto"):
        _lib.require_cuda()
        _lib.load()
        if metric not in _lib.METRICS:
            raise ValueError(f"metric must be 'cosine' or 'l2', got {metric!r}")
        self.db_dir = db_dir
        self.metric = metric
        self._metric = _lib.METRICS[metric]
        self.device = torch.device(device or f"cuda:{torch.cuda.current_device()}")
        self.dim: Optional[int] = None
        self._dim_p = 0
        self._cap0 = max(16, int(capacity))
        self._n = 0                                   # rows appended so far (live + deleted)
        self._rows = self._norms = self._tags = None
        self._lock = threading.RLock()
        self._version = 1
        # host-side string world
        self._users: Dict[str, int] = {}
        self._user_names: List[str] = []
        self._span: List[List[int]] = []              # per user [lo, hi) over rows ever added
        self._live: List[int] = []                    # per user live row count
        self._row_id: Dict[int, str] = {}             # explicit ids
        self._row_meta: Dict[int, Dict[str, Any]] = {}
        self._row_user = np.empty(0, np.int32)        # host copy of tags (for get/delete bookkeeping)
        self._id_rows: Dict[tuple, List[int]] = {}    # (user idx, id) -> rows
        self._blocks: List[tuple] = []                # implicit-id blocks (lo, hi, prefix, first)
        self._edges: List[Dict[str, Any]] = []
        self._profiles: Dict[str, Dict[str, Any]] = {}
        self._pin: Dict[str, torch.Tensor] = {}
        self._io: Dict[tuple, tuple] = {}
        self._graphs: Dict[tuple, object] = {}
        self._gtags: Dict[tuple, torch.Tensor] = {}
        self.use_graphs = True                 # CUDA-graph replay of the (launch-bound) exact path for host calls
        self.use_fused_one = True              # single host queries: ONE launch reading the query from pinned host memory and
                                               # writing the answer back to it (lz_search_one), the host polls a flag
        self._one: Dict[int, dict] = {}        # per k: pinned query / answer block, device scratch, sequence number
        self.fused_one_max_rows = 65_536       # beyond this a 32-CTA scan is slower than the full-grid scan + merge
        self._scratch: Optional[torch.Tensor] = None
        self._gen = 0                          # bumped whenever rows / norms / tags are reallocated or repacked
        if batched_search not in ("auto", "never", "always"):
            raise ValueError("batched_search must be 'auto', 'never' or 'always'")
        self.batched_search = batched_search
        self.min_batch = 8                    # 'auto': query batches this large go to the tensor-core path
        self.min_rows_small_batch = 500_000   # ... and smaller batches too when the scan is this long: the coarse pass reads
                                              # the fp16 shadow, half the bytes of the float32 rows (10M x 768: one query
                                              # 3.3 ms against 4.7 ms for the exact scan, 2-7 queries 3.4 against 5-9 ms)
        self.shadow_for_single_queries = True # False: single queries always take the exact scan (no shadow needed for them)
        self._shadow_oom = False              # the shadow did not fit beside the rows: small batches stay on the exact scan
        self.margin_sigmas = 4.0
        # certification margin of the tensor-core path: "worst_case" = deterministic bound from the measured bf16
        # rounding-error norms of rows and query (valid for any data; the default, because north_star asks for
        # bit-exact ids); "statistical" = margin_sigmas standard deviations under independent rounding errors
        # (narrower, fewer fallbacks, but not a proof — see DESIGN.md §4)
        self.margin = "worst_case"
        # fp16 shadows by metric, built on first use: the store's own metric, and cosine on an l2 store when the
        # consolidation code asks for cosine neighbours (memory_system.py:858-867 ranks by cosine whatever the store does)
        self._shadows: Dict[int, _Shadow] = {}
        self._bscratch: Dict[int, torch.Tensor] = {}      # tensor-core scratch per lane (searches in flight on different streams)
        self.last_certified: Optional[torch.Tensor] = None
        self._nodes_table = self                      # memory_system.py:1434-1439 reaches for this
        # search policy: "exact" (the parity mode: every row of the user is scored) or "two_level" (approximate:
        # centroid routing, then only the `nprobe` best shards — set_shards() first; reported with recall, never
        # as parity; see recall_report)
        self.search_policy = "exact"
        self.nprobe: Optional[int] = None
        self._centroids = None
        if dim is not None:
            self._set_dim(int(dim))

    # ------------------------------------------------------------------ storage management
    def _set_dim(self, dim: int):
        self.dim = dim
        self._dim_p = _round_up(dim, 8)
        self._alloc(self._cap0)

    def _alloc(self, cap: int):
        rows = torch.zeros((cap, self._dim_p), dtype=torch.float32, device=self.device)
        norms = torch.zeros(cap, dtype=torch.float64, device=self.device)
        tags = torch.full((cap,), -1, dtype=torch.int32, device=self.device)
        if self._rows is not None and self._n:
            rows[:self._n] = self._rows[:self._n]
            norms[:self._n] = self._norms[:self._n]
            tags[:self._n] = self._tags[:self._n]
        self._rows, self._norms, self._tags = rows, norms, tags
        self._gen += 1
        self._graphs.clear()                   # captured graphs hold the old columns' addresses

    def _shadow_for(self, metric: int) -> _Shadow:
        """The up-to-date fp16 shadow for `metric` (created on first use)."""
        sh = self._shadows.get(metric)
        if sh is None:
            sh = self._shadows[metric] = _Shadow(metric)
        sh.ensure(self)
        return sh

    @property
    def _shadow(self) -> Optional[torch.Tensor]:
        sh = self._shadows.get(self._metric)
        return None if sh is None else sh.t

    def _reserve(self, extra: int):
        need = self._n + extra
        if need > self._rows.shape[0]:
            self._alloc(max(need, 2 * self._rows.shape[0]))
        if need > self._row_user.shape[0]:
            grown = np.full(max(need, 2 * self._row_user.shape[0], 16), -1, np.int32)
            grown[:self._n] = self._row_user[:self._n]
            self._row_user = grown

    def _pinned(self, name: str, shape, dtype) -> torch.Tensor:
        t = self._pin.get(name)
        n = int(np.prod(shape))
        if t is None or t.numel() < n or t.dtype != dtype:
            t = torch.empty(max(n, 1), dtype=dtype, pin_memory=True)
            self._pin[name] = t
        return t[:n].view(*shape)

    def _user(self, user_id: str, create: bool = False) -> int:
        u = self._users.get(user_id, -1)
        if u < 0 and create:
            u = len(self._user_names)
            self._users[user_id] = u
            self._user_names.append(user_id)
            self._span.append([1 << 62, 0])
            self._live.append(0)
        return u

    def _id_of(self, row: int) -> str:
        s = self._row_id.get(row)
        if s is not None:
            return s
        for lo, hi, prefix, first in self._blocks:
            if lo <= row < hi:
                return f"{prefix}{first + row - lo}"
        return str(row)

    def _ids_of(self, ids: np.ndarray, counts: np.ndarray) -> List[List[str]]:
        """Id strings of a [B, k] block of row indices (first counts[b] entries of each line).  The bulk layout —
        one adopted block of rows, no per-row id overrides — formats without the per-row lookups of _id_of."""
        lists = np.asarray(ids).tolist()
        cnt = np.asarray(counts).tolist()
        if not self._row_id and len(self._blocks) == 1:
            lo, hi, prefix, first = self._blocks[0]
            off = first - lo
            return [[f"{prefix}{r + off}" if lo <= r < hi else str(r) for r in row[:c]] for row, c in zip(lists, cnt)]
        return [[self._id_of(r) for r in row[:c]] for row, c in zip(lists, cnt)]

    @property
    def n_rows(self) -> int:
        return self._n

    # ------------------------------------------------------------------ nodes (vector_store.py:89-148)
    def add_nodes(self, nodes: List[Dict[str, Any]], user_id: str = "default"):
        if not nodes:                                               # vector_store.py:90-91
            return
        emb = np.asarray([n["embedding"] for n in nodes], dtype=np.float32)
        if emb.ndim != 2:
            raise ValueError("every node needs an 'embedding' of the same length")
        with self._lock:
            if self.dim is None:
                self._set_dim(int(emb.shape[1]))
            if emb.shape[1] != self.dim:
                raise ValueError(f"embedding has {emb.shape[1]} dims, store has {self.dim}")
            n0 = self._append_rows_host(emb, user_id)
            u = self._users[user_id]
            for i, node in enumerate(nodes):
                row = n0 + i
                self._row_id[row] = node["id"]
                self._id_rows.setdefault((u, node["id"]), []).append(row)
                self._row_meta[row] = {                             # vector_store.py:95-110
                    "content": node["content"],
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
                }
            self._version += 1

    def _append_rows_host(self, emb: np.ndarray, user_id: str) -> int:
        n = emb.shape[0]
        self._reserve(n)
        n0 = self._n
        stage = self._pinned("rows", (n, self._dim_p), torch.float32)
        stage.zero_()
        stage[:, :self.dim] = torch.from_numpy(emb)
        with torch.cuda.device(self.device):
            self._rows[n0:n0 + n].copy_(stage, non_blocking=True)
            torch.cuda.current_stream().synchronize()               # staging buffer is reused
        self._finish_append(n0, n, user_id)
        return n0

    def _finish_append(self, n0: int, n: int, user_id: str):
        u = self._user(user_id, create=True)
        kernels.row_norms(self._rows[n0:n0 + n], self._dim_p, out=self._norms[n0:n0 + n])
        self._tags[n0:n0 + n] = u
        self._row_user[n0:n0 + n] = u
        self._n = n0 + n
        sp = self._span[u]
        sp[0], sp[1] = min(sp[0], n0), max(sp[1], n0 + n)
        self._live[u] += n

    def add_rows(self, rows: torch.Tensor, user_id: str = "default", id_prefix: str = "row_",
                 first_id: Optional[int] = None) -> int:
        """Bulk ingest of float32 rows already in HBM (or a numpy array): no per-row Python objects;
        ids are implicit, f"{id_prefix}{first_id + i}".  Returns the first row index.  This is the
        ingest used at benchmark scale; `add_nodes` is the reference-shaped call."""
        with self._lock:
            if isinstance(rows, np.ndarray):
                rows = torch.from_numpy(np.ascontiguousarray(rows, dtype=np.float32))
            if self.dim is None:
                self._set_dim(int(rows.shape[1]))
            if rows.shape[1] != self.dim:
                raise ValueError(f"rows have {rows.shape[1]} dims, store has {self.dim}")
            n = rows.shape[0]
            self._reserve(n)
            n0 = self._n
            self._rows[n0:n0 + n, :self.dim].copy_(rows.to(torch.float32), non_blocking=True)
            if self._dim_p != self.dim:
                self._rows[n0:n0 + n, self.dim:].zero_()
            self._finish_append(n0, n, user_id)
            self._blocks.append((n0, n0 + n, id_prefix, n0 if first_id is None else first_id))
            self._version += 1
            return n0

    def adopt_rows(self, rows: torch.Tensor, user_id: str = "default", id_prefix: str = "row_"):
        """Take ownership of a [n, dim] float32 CUDA tensor as the store's row matrix without copying
        (dim % 8 == 0, store must be empty).  For benchmark-size tables that should exist once in HBM."""
        with self._lock:
            assert self._n == 0 and rows.is_cuda and rows.dtype == torch.float32 and rows.is_contiguous()
            assert rows.shape[1] % 8 == 0, "adopt_rows needs dim % 8 == 0"
            self.dim = self._dim_p = int(rows.shape[1])
            self.device = rows.device
            n = rows.shape[0]
            self._rows = rows
            self._gen += 1
            self._graphs.clear()
            self._norms = torch.empty(n, dtype=torch.float64, device=self.device)
            self._tags = torch.empty(n, dtype=torch.int32, device=self.device)
            self._row_user = np.full(n, -1, np.int32)
            self._finish_append(0, n, user_id)
            self._blocks.append((0, n, id_prefix, 0))
            self._version += 1

    def adopt_tenants(self, rows: torch.Tensor, rows_per_user: int, first_user: int = 0, user_prefix: str = "user_"):
        """Take ownership of a [n_users * rows_per_user, dim] float32 CUDA tensor holding equal-size tenants back to back
        (store must be empty, dim % 8 == 0): user f"{user_prefix}{first_user + u}" owns rows [u*rows_per_user,
        (u+1)*rows_per_user), ids are implicit f"{user}/{i}".  The multi-tenant benchmark layout (BASELINE configs[4]):
        no per-row Python objects, one tag fill on the device."""
        with self._lock:
            assert self._n == 0 and rows.is_cuda and rows.dtype == torch.float32 and rows.is_contiguous()
            assert rows.shape[1] % 8 == 0 and rows.shape[0] % rows_per_user == 0
            n, n_users = rows.shape[0], rows.shape[0] // rows_per_user
            self.dim = self._dim_p = int(rows.shape[1])
            self.device = rows.device
            self._rows = rows
            self._gen += 1
            self._graphs.clear()
            self._norms = kernels.row_norms(rows, self._dim_p)
            self._tags = (torch.arange(n, device=self.device, dtype=torch.int64) // rows_per_user).to(torch.int32)
            self._row_user = np.repeat(np.arange(n_users, dtype=np.int32), rows_per_user)
            self._n = n
            for u in range(n_users):
                name = f"{user_prefix}{first_user + u}"
                self._users[name] = u
                self._user_names.append(name)
                self._span.append([u * rows_per_user, (u + 1) * rows_per_user])
                self._live.append(rows_per_user)
                self._blocks.append((u * rows_per_user, (u + 1) * rows_per_user, name + "/", 0))
            self._version += 1

    def user_spans(self):
        """(lo int64 [n_users], hi int64 [n_users]) row span of every user, by user index (numpy, rebuilt on demand)."""
        sp = np.asarray(self._span, dtype=np.int64).reshape(-1, 2)
        return sp[:, 0].copy(), np.maximum(sp[:, 1], sp[:, 0])

    def get_nodes(self, user_id: str = "default") -> List[Dict[str, Any]]:
        with self._lock:
            u = self._user(user_id)
            if u < 0 or self._n == 0:                                # vector_store.py:116-117
                return []
            rows = np.nonzero(self._row_user[:self._n] == u)[0]
            if rows.size == 0:
                return []
            idx = torch.from_numpy(rows).to(self.device)
            vecs = self._rows.index_select(0, idx)[:, :self.dim].cpu().numpy()
            out = []
            for r, v in zip(rows.tolist(), vecs):
                m = self._row_meta.get(r, {})
                d = {"id": self._id_of(r), "user_id": user_id, "content": m.get("content", ""),
                     "vector": v.tolist(), "type": m.get("type", "semantic"),
                     "timestamp": m.get("timestamp", 0.0), "access_count": m.get("access_count", 0),
                     "last_accessed": m.get("last_accessed", 0.0), "salience": m.get("salience", 0.5),
                     "is_super_node": m.get("is_super_node", False),
                     "child_ids": json.loads(m.get("child_ids", "[]")),          # vector_store.py:119-129
                     "parent_id": m.get("parent_id", ""), "shard_key": m.get("shard_key", "default"),
                     "metadata": json.loads(m.get("metadata", "{}"))}
                out.append(d)
            return out

    def search_nodes(self, query_emb: Sequence[float], user_id: str = "default", limit: int = 5) -> List[str]:
        """vector_store.py:132-140: ids of the `limit` best rows of `user_id`, best first."""
        with self._lock:
            u = self._user(user_id)
            if self._n == 0 or u < 0 or self._live[u] == 0 or limit <= 0:   # vector_store.py:133-134
                return []
            q = np.asarray(query_emb, dtype=np.float32).reshape(1, -1)
            if self.search_policy == "two_level":
                return self._search_policy_two_level(q, u, limit)[0]
            ids, _, counts = self._search_host(q, u, limit)
            return [self._id_of(r) for r in ids[0, :counts[0]].tolist()]

    def _search_policy_two_level(self, q: np.ndarray, u: int, limit: int):
        if self.search_policy not in ("exact", "two_level"):
            raise ValueError(f"search_policy must be 'exact' or 'two_level', got {self.search_policy!r}")
        if self._centroids is None:
            raise RuntimeError("search_policy='two_level' needs set_shards() first")
        if len(self._user_names) != 1:
            raise RuntimeError("search_policy='two_level' serves single-tenant stores; per-user spans use the exact path")
        return self.search_two_level(q, limit=limit, nprobe=self.nprobe)

    def recall_report(self, queries, limit: int = 10, nprobes: Sequence[int] = (1, 2, 4, 8, 16, 32)):
        """Approximate-mode report: for each nprobe the recall@limit of the two-level search
        against the exact flat search on the same queries, the fraction of rows its member scans touch, and its
        time per batch (CUDA events).  Returns a list of dicts; nothing here is a parity claim."""
        with self._lock:
            q = np.ascontiguousarray(queries, dtype=np.float32)
            exact = self.search_batch(q, self._user_names[0], limit)
            S = int(self._centroids.shape[0])
            sizes = self._shard_off_host[1:] - self._shard_off_host[:-1]
            out = []
            for p in nprobes:
                p = max(1, min(int(p), S))
                self.search_two_level(q, limit, nprobe=p)                                  # warm-up
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                rows, _, cnt = self.search_two_level(q, limit, nprobe=p, return_rows=True)
                e1.record()
                torch.cuda.synchronize(self.device)
                rows, cnt = rows.cpu().numpy(), cnt.cpu().numpy()
                got = [[self._id_of(int(r)) for r in rows[b, :cnt[b]]] for b in range(q.shape[0])]
                rec = float(np.mean([len(set(a) & set(b)) / max(1, len(b)) for a, b in zip(got, exact)]))
                out.append({"nprobe": p, "shards": S, f"recall_at_{limit}": rec,
                            "rows_scanned_fraction_upper": float(np.sort(sizes)[::-1][:p].sum() / max(1, sizes.sum())),
                            "ms_per_batch": e0.elapsed_time(e1), "batch": int(q.shape[0])})
            return out

    def search_batch(self, queries, user_id: str = "default", limit: int = 5, return_scores: bool = False):
        """Batched form of `search_nodes` (our addition): queries float32 [B, dim] on the host.
        Returns a list of id lists (and the float64 score lists if asked)."""
        with self._lock:
            q = np.ascontiguousarray(queries, dtype=np.float32)
            u = self._user(user_id)
            if self._n == 0 or u < 0 or self._live[u] == 0 or limit <= 0:
                empty = [[] for _ in range(q.shape[0])]
                return (empty, empty) if return_scores else empty
            ids, scores, counts = self._search_host(q, u, limit)
            out = self._ids_of(ids, counts)
            if return_scores:
                return out, [scores[b, :counts[b]].tolist() for b in range(q.shape[0])]
            return out

    def _check_limit(self, limit: int, u: int) -> int:
        k = min(int(limit), self._live[u])
        if k > _lib.MAX_K:
            raise ValueError(f"limit={limit} exceeds the kernel's top-k bound of {_lib.MAX_K}")
        return max(k, 1)

    def _io_buffers(self, B: int, k: int):
        """Per-(B, k) reusable staging: pinned + device query buffers and ONE packed result buffer
        (ids int64 | scores float64 | counts int32) on each side, so a search is one H2D copy, the kernel
        launches and one D2H copy.  At the reference's own scale (10k rows) the call overhead is the cost."""
        key = (B, k, self._dim_p)
        io = self._io.get(key)
        if io is None:
            if len(self._io) > 32:
                self._io.clear()
            nb = B * k * 16 + B * 4
            q_pin = torch.zeros((B, self._dim_p), dtype=torch.float32).pin_memory()
            q_dev = torch.zeros((B, self._dim_p), dtype=torch.float32, device=self.device)
            o_dev = torch.empty(nb, dtype=torch.uint8, device=self.device)
            o_pin = torch.empty(nb, dtype=torch.uint8).pin_memory()

            def views(t):
                return (t[:B * k * 8].view(torch.int64).view(B, k), t[B * k * 8:B * k * 16].view(torch.float64).view(B, k),
                        t[B * k * 16:].view(torch.int32))
            io = self._io[key] = (q_pin, q_dev, o_dev, o_pin, views(o_dev), tuple(v.numpy() for v in views(o_pin)))
        return io

    def _search_one_host(self, q: np.ndarray, u: int, k: int):
        """One host query through lz_search_one: the kernel reads the query from pinned host memory, scans the user's row
        span in at most 32 CTAs, the last CTA merges and writes ids / scores / count into pinned host memory followed by
        a sequence flag that this thread polls — one launch, no copies enqueued, no stream synchronisation."""
        st = self._one.get(k)
        lib = _lib.load()
        if st is None:
            q_pin = torch.zeros(self._dim_p, dtype=torch.float32).pin_memory()
            out = torch.zeros(k * 16 + 8, dtype=torch.uint8).pin_memory()
            nbytes = int(lib.lz_search_one_scratch_bytes(k))
            scratch = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)      # zero: the completion ticket
            o = out.numpy()
            st = self._one[k] = {"q_pin": q_pin, "q": q_pin.numpy(), "out": out, "scratch": scratch, "nbytes": nbytes, "seq": 0,
                                 "ids": o[:k * 8].view(np.int64), "sc": o[k * 8:k * 16].view(np.float64),
                                 "cnt": o[k * 16:k * 16 + 4].view(np.int32), "flag": o[k * 16 + 4:].view(np.int32)}
        st["q"][:self.dim] = q[0]
        st["seq"] = seq = (st["seq"] % 0x7ffffff0) + 1
        lo, hi = self._span[u]
        single_user = len(self._user_names) == 1 and self._live[u] == self._n
        key = (self._gen, u, lo, hi, single_user)
        if st.get("key") != key:                            # the call's fixed arguments, rebuilt when the layout changes
            st["key"] = key
            st["args"] = (self._rows.data_ptr(), self._rows.stride(0), self._dim_p, self._norms.data_ptr(),
                          None if single_user else self._tags.data_ptr(), -1 if single_user else u, lo, hi,
                          st["q_pin"].data_ptr(), k, self._metric, 0, st["scratch"].data_ptr(), st["nbytes"], st["out"].data_ptr())
        idx = self.device.index
        guard = torch.cuda.device(self.device) if torch.cuda.current_device() != idx else None
        if guard is not None:
            guard.__enter__()
        try:
            stream = torch.cuda.current_stream()
            _lib.check(lib.lz_search_one(*st["args"], seq, stream.cuda_stream), "lz_search_one")
            flag = st["flag"]
            for _ in range(200_000):                       # the answer usually lands within ~20 us
                if flag[0] == seq:
                    break
            else:
                stream.synchronize()
                if flag[0] != seq:
                    st["scratch"][:256].zero_()             # an aborted launch may have left the completion ticket half way
                    raise _lib.LazzaroB200Error("lz_search_one finished without publishing its answer")
        finally:
            if guard is not None:
                guard.__exit__(None, None, None)
        return st["ids"][None, :], st["sc"][None, :], st["cnt"]

    def _search_host(self, q: np.ndarray, u: int, limit: int):
        """host queries [B, dim] -> (row ids int64 [B,k], scores float64 [B,k], counts int32 [B]) as numpy
        views of a pinned buffer (valid until the next search), one stream synchronisation at the end."""
        if q.shape[1] != self.dim:
            raise ValueError(f"query has {q.shape[1]} dims, store has {self.dim}")
        k = self._check_limit(limit, u)
        B = q.shape[0]
        if B == 1 and self.use_fused_one and self._span[u][1] - self._span[u][0] <= self.fused_one_max_rows:
            return self._search_one_host(q, u, k)           # small spans: launches and copies cost more than the scan
        q_pin, q_dev, o_dev, o_pin, dev_views, host_views = self._io_buffers(B, k)
        with torch.cuda.device(self.device):
            q_pin[:, :self.dim] = torch.from_numpy(q)
            lo, hi = self._span[u]
            if self.use_graphs and not self._takes_batched_path(B, k, hi - lo):
                # the exact path is launch-bound on small stores (the reference's own scale): replay the
                # H2D copy, scan, merge and D2H copy as one CUDA graph, re-captured when the layout changes
                key = (B, k, u, lo, hi, self._gen, len(self._user_names) == 1 and self._live[u] == self._n)
                entry = self._graphs.get(key)
                if entry is None:
                    if len(self._graphs) > 64:
                        self._graphs.clear()
                    # A captured graph replays raw addresses.  Everything it touches besides the store's columns (whose
                    # reallocation bumps _gen) is owned by the graph's own entry and lives exactly as long as the graph:
                    # its scratch (never the shared, growable self._scratch), its group tags and its I/O buffers.
                    io = (q_pin, q_dev, o_dev, o_pin, dev_views, host_views)
                    own = {"scratch": torch.empty(self._exact_scratch_bytes(B, k, hi - lo), dtype=torch.uint8, device=self.device),
                           "gtag": None}
                    torch.cuda.current_stream().synchronize()
                    self.search_device(q_dev, u, k, out=dev_views, _own=own)          # warm-up outside capture
                    torch.cuda.current_stream().synchronize()
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph):
                        q_dev.copy_(q_pin, non_blocking=True)
                        self.search_device(q_dev, u, k, out=dev_views, _own=own)
                        o_pin.copy_(o_dev, non_blocking=True)
                    entry = self._graphs[key] = (graph, own, io)
                entry[0].replay()
            else:
                q_dev.copy_(q_pin, non_blocking=True)
                self.search_device(q_dev, u, k, out=dev_views)
                o_pin.copy_(o_dev, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        return host_views

    @staticmethod
    def _check_margin(margin: str) -> str:
        if margin not in ("worst_case", "statistical"):
            raise ValueError(f"margin must be 'worst_case' or 'statistical', got {margin!r}")
        return margin

    def _takes_batched_path(self, B: int, k: int, span: int) -> bool:
        return (k <= _lib.BATCHED_MAX_K and
                (self.batched_search == "always" or
                 (self.batched_search == "auto" and
                  (B >= self.min_batch or (span >= self.min_rows_small_batch and not self._shadow_oom and
                                           (B >= 2 or self.shadow_for_single_queries))))))

    def _exact_groups(self, B: int):
        gs = 4 if B >= 4 else (2 if B >= 2 else 1)
        return gs, (B + gs - 1) // gs

    def _exact_scratch_bytes(self, B: int, k: int, span: int) -> int:
        _, n_groups = self._exact_groups(B)
        n_chunks = kernels.default_chunks(span, n_groups, self.device)
        return max(16, int(_lib.load().lz_search_exact_scratch_bytes(B, k, n_chunks)))

    def search_device(self, queries: torch.Tensor, user: int | str = 0, k: int = 10, id_base: int = 0,
                      row_range=None, out=None, _own=None, metric: Optional[str] = None, margin: Optional[str] = None,
                      lane: int = 0):
        """Device-in / device-out search: queries float32 [B, dim_p] in HBM -> (ids, scores, counts)
        tensors, nothing synchronised.  Small batches: the exact fp64 scan over the user's row span; larger
        ones: the tensor-core path (same answer).
        `metric`  rank by "cosine" / "l2" instead of the store's own metric (a second fp16 shadow is built on first use)
        `margin`  certification margin of the tensor-core path for this call (default: the store's `margin`)
        `lane`    which of the store's scratch buffers the tensor-core path uses: searches kept in flight on different
                  streams must use different lanes
        `_own`    (internal) scratch / group tags owned by a captured CUDA graph instead of the store's shared ones."""
        u = self._user(user) if isinstance(user, str) else user
        lo, hi = self._span[u]
        if row_range is not None:                      # e.g. one semantic shard's contiguous rows
            lo, hi = max(lo, int(row_range[0])), min(hi, int(row_range[1]))
            hi = max(hi, lo)
        single_user = len(self._user_names) == 1 and self._live[u] == self._n
        tags = None if single_user else self._tags
        B = queries.shape[0]
        mcode = self._metric if metric is None else _lib.METRICS[metric]
        batched = self._takes_batched_path(B, k, hi - lo)
        if batched and B < self.min_batch and self.batched_search == "auto":
            # for small batches the shadow is an accelerator, not a requirement: when it does not fit beside the rows
            # (+50 % of their bytes) they keep the exact scan — on the GPU, same answer
            try:
                self._shadow_for(mcode)
            except torch.cuda.OutOfMemoryError:
                self._shadows.pop(mcode, None)
                self._shadow_oom = True
                batched = False
        if batched:
            sh = self._shadow_for(mcode)
            need = int(_lib.load().lz_search_batched_scratch_bytes(hi - lo, B, k, self._dim_p))
            if lane not in self._bscratch or self._bscratch[lane].numel() < need:
                self._bscratch[lane] = torch.empty(need, dtype=torch.uint8, device=self.device)
            det = sh.max_err if self._check_margin(margin or self.margin) == "worst_case" else None
            ids, sc, cnt, cert = kernels.search_batched(sh.t, self._rows, self._dim_p, self._norms, tags,
                                                        -1 if tags is None else u, queries, k, row_range=(lo, hi),
                                                        margin_sigmas=self.margin_sigmas, id_base=id_base,
                                                        scratch=self._bscratch[lane], n_rows_total=self._n,
                                                        metric=mcode, max_row_norm=sh.max_norm, out=out,
                                                        det_row_err=det)
            self.last_certified = cert
            return ids, sc, cnt
        gs, n_groups = self._exact_groups(B)
        gtag = None
        if tags is not None:
            if _own is not None:
                if _own["gtag"] is None:
                    _own["gtag"] = torch.full((n_groups,), u, dtype=torch.int32, device=self.device)
                gtag = _own["gtag"]
            else:
                gtag = self._gtags.get((n_groups, u))
                if gtag is None:
                    if len(self._gtags) > 256:
                        self._gtags.clear()
                    gtag = self._gtags[(n_groups, u)] = torch.full((n_groups,), u, dtype=torch.int32, device=self.device)
        n_chunks = kernels.default_chunks(hi - lo, n_groups, self.device)
        need = int(_lib.load().lz_search_exact_scratch_bytes(B, k, n_chunks))
        if _own is not None:
            scratch = _own["scratch"]
            assert scratch.numel() >= need
        elif lane != 0:
            key = -lane                                    # exact-path scratch of the extra lanes lives beside the others
            if key not in self._bscratch or self._bscratch[key].numel() < need:
                self._bscratch[key] = torch.empty(need, dtype=torch.uint8, device=self.device)
            scratch = self._bscratch[key]
        else:
            if self._scratch is None or self._scratch.numel() < need:
                self._scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
            scratch = self._scratch
        return kernels.search_exact(self._rows, self._dim_p, self._norms, tags, queries, k, mcode,
                                    row_range=(lo, hi), group_tag=gtag, group_size=gs, n_chunks=n_chunks,
                                    scratch=scratch, id_base=id_base, out=out)

    # ------------------------------------------------------------------ multi-tenant and two-level search
    def search_users(self, queries, user_ids: Sequence[str], limit: int = 5):
        """One query per (possibly different) user in a single launch: query i is answered from the rows
        of user_ids[i] only (the per-user pre-filter of vector_store.py:137, batched).  Each query scans
        just its user's row span, so 100k small tenants cost what their rows cost, not the table.
        Returns a list of id lists."""
        with self._lock:
            q = np.ascontiguousarray(queries, dtype=np.float32)
            B = q.shape[0]
            assert len(user_ids) == B
            us = np.array([self._user(u) for u in user_ids], dtype=np.int64)
            known = us >= 0
            lo = np.array([self._span[u][0] if u >= 0 else 0 for u in us], dtype=np.int64)
            hi = np.array([self._span[u][1] if u >= 0 else 0 for u in us], dtype=np.int64)
            hi = np.maximum(hi, lo)
            live = np.array([self._live[u] if u >= 0 else 0 for u in us])
            if self._n == 0 or not known.any() or limit <= 0:
                return [[] for _ in range(B)]
            k = int(min(limit, max(1, live.max())))
            if k > _lib.MAX_K:
                raise ValueError(f"limit={limit} exceeds the kernel's top-k bound of {_lib.MAX_K}")
            ids, sc, cnt = self.search_ranges_device(self._to_device(q), lo, hi, us.astype(np.int32), k)
            torch.cuda.current_stream(self.device).synchronize()
            ids, cnt = ids.cpu().numpy(), cnt.cpu().numpy()
            return [[self._id_of(int(r)) for r in ids[b, :cnt[b]]] if known[b] else [] for b in range(B)]

    def _to_device(self, q: np.ndarray) -> torch.Tensor:
        if q.shape[1] != self.dim:
            raise ValueError(f"query has {q.shape[1]} dims, store has {self.dim}")
        out = torch.zeros((q.shape[0], self._dim_p), dtype=torch.float32, device=self.device)
        out[:, :self.dim] = torch.from_numpy(q).to(self.device, non_blocking=True)
        return out

    def search_ranges_device(self, queries: torch.Tensor, lo, hi, tag, k: int):
        """Exact search of query i over rows [lo[i], hi[i]) restricted to tag[i] (< 0: any live row)."""
        dev = self.device
        as_dev = lambda a, dt: a if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a), dtype=dt).to(dev, non_blocking=True)
        if isinstance(lo, torch.Tensor):
            span = int(getattr(lo, "_lz_max_span", 0)) or int((hi - lo).max().item())
        else:
            span = int(np.max(np.asarray(hi) - np.asarray(lo))) if len(lo) else 0
        n_chunks = kernels.default_chunks(max(span, 1), queries.shape[0], dev)
        return kernels.search_exact(self._rows, self._dim_p, self._norms, self._tags, queries, k, self._metric,
                                    row_range=(0, self._n), group_lo=as_dev(lo, torch.int64),
                                    group_hi=as_dev(hi, torch.int64), group_tag=as_dev(tag, torch.int32),
                                    group_size=1, n_chunks=n_chunks)

    def set_shards(self, shard_offsets: Sequence[int]):
        """Declare the semantic-shard layout of the rows: shard s is the contiguous row range
        shard_offsets[s]:shard_offsets[s+1] (rows are stored shard by shard).  Computes one centroid per
        shard = mean of its member rows (memory_system.py:916-917) and the centroid table's norms and fp16 shadow
        for the two-level search."""
        with self._lock:
            off = torch.as_tensor(np.asarray(shard_offsets, dtype=np.int64)).to(self.device)
            assert int(off[-1]) <= self._n and int(off[0]) >= 0
            cen = kernels.segment_mean(self._rows, self._dim_p, off).to(torch.float32).contiguous()
            self._shard_off, self._shard_off_host = off, np.asarray(shard_offsets, dtype=np.int64)
            self._centroids, self._centroid_norms = cen, kernels.row_norms(cen)
            err = torch.empty(cen.shape[0], dtype=torch.float32, device=self.device)
            self._centroid_shadow = kernels.rows_to_shadow(cen, self._dim_p, self._centroid_norms, row_err=err)
            self._centroid_err = float(err.max().item())
            self._tl_scratch = None

    def search_two_level_device(self, queries: torch.Tensor, k: int, nprobe: Optional[int] = None, margin: Optional[str] = None,
                                want_probe: bool = False):
        """Two-level search, device in / device out, nothing synchronised: queries float32 [B, dim_p] ->
        (ids, scores, counts[, probe]).  Level 1 always ranks the shards by cosine against their centroids
        (memory_system.py:466-470); level 2 ranks the member rows of the `nprobe` best shards by the store's metric.
        Batches take the tensor-core path (lz_search_two_level: centroid GEMM, then a grouped member GEMM over the
        probed shards); fewer than 8 queries take the exact float64 kernels (centroid scan, then ONE multi-range member
        scan, lz_search_exact_ranges).  Both return the exact top-k over the union of the probed shards."""
        S = int(self._centroids.shape[0])
        P = S if nprobe is None else max(1, min(int(nprobe), S))
        B = queries.shape[0]
        if self._takes_batched_path(B, k, self._n):
            sh = self._shadow_for(self._metric)
            worst = self._check_margin(margin or self.margin) == "worst_case"
            ids, sc, cnt, cert, probe, self._tl_scratch = kernels.search_two_level(
                sh.t, self._rows, self._dim_p, self._norms, self._n, self._shard_off, self._centroid_shadow, self._centroids,
                self._centroid_norms, queries, k, P, metric=self._metric, margin_sigmas=self.margin_sigmas,
                max_row_norm=sh.max_norm, det_row_err=sh.max_err if worst else None,
                cen_det_row_err=self._centroid_err if worst else None, scratch=self._tl_scratch, want_probe=want_probe)
            self.last_certified = cert
        else:
            probe, _, _ = kernels.search_exact(self._centroids, self._dim_p, self._centroid_norms, None, queries, P, _lib.COSINE)
            pc = probe.clamp(min=0)
            lo = torch.where(probe >= 0, self._shard_off[:-1][pc], torch.zeros_like(probe)).contiguous()
            hi = torch.where(probe >= 0, self._shard_off[1:][pc], torch.zeros_like(probe)).contiguous()
            ids, sc, cnt = kernels.search_exact_ranges(self._rows, self._dim_p, self._norms, queries, lo, hi, k, self._metric,
                                                       n_rows_total=self._n)
        return (ids, sc, cnt, probe) if want_probe else (ids, sc, cnt)

    def search_two_level(self, queries, limit: int = 5, nprobe: Optional[int] = None, return_rows: bool = False):
        """Two-level cosine search (north_star (1); the reference's level 1 is memory_system.py:464-470):
        score the queries against the shard centroids, then search only the `nprobe` best shards' member
        rows.  nprobe = None or >= the number of shards visits every shard and equals the flat
        search (the parity mode); smaller nprobe is approximate — callers report recall."""
        with self._lock:
            if len(self._user_names) != 1 or self._live[0] != self._n:
                raise RuntimeError("two-level search serves single-tenant stores without deleted rows")
            q = np.ascontiguousarray(queries, dtype=np.float32)
            qd = self._to_device(q)
            B = q.shape[0]
            k = self._check_limit(limit, 0)
            ids, sc, cnt = self.search_two_level_device(qd, k, nprobe)
            torch.cuda.current_stream(self.device).synchronize()
            if return_rows:
                return ids, sc, cnt
            ids, cnt = ids.cpu().numpy(), cnt.cpu().numpy()
            return self._ids_of(ids, cnt)

    def delete_nodes(self, node_ids: List[str], user_id: str = "default"):
        with self._lock:
            u = self._user(user_id)
            if u < 0:
                return
            if node_ids is None or len(node_ids) == 0:               # vector_store.py:143-145
                rows = np.nonzero(self._row_user[:self._n] == u)[0]
            else:
                found: List[int] = []
                for nid in node_ids:
                    found.extend(self._id_rows.pop((u, nid), []))
                    found.extend(self._implicit_rows(nid, u))
                rows = np.asarray(sorted(set(found)), dtype=np.int64)
                rows = rows[self._row_user[rows] == u] if rows.size else rows
            if rows.size:
                self._tags.index_fill_(0, torch.from_numpy(rows.astype(np.int64)).to(self.device), -1)
                self._row_user[rows] = -1
                self._live[u] -= int(rows.size)
                for r in rows.tolist():
                    rid = self._row_id.pop(r, None)
                    self._row_meta.pop(r, None)
                    if rid is not None:
                        lst = self._id_rows.get((u, rid))
                        if lst and r in lst:
                            lst.remove(r)
                            if not lst:
                                del self._id_rows[(u, rid)]
            self._version += 1
            dead = self._n - sum(self._live)
            if dead > max(4096, sum(self._live)):
                self._compact()

    def _compact(self):
        """Drops tombstoned rows (the reference rewrites a user's whole graph on every save,
        memory_system.py:1283-1287, so deleted rows pile up).  Live rows keep their relative order, so
        the insertion-order tie-break is unchanged.  Device side is a gather (torch indexing = plumbing)."""
        live = np.nonzero(self._row_user[:self._n] >= 0)[0]
        idx = torch.from_numpy(live).to(self.device)
        m = int(live.size)
        for name in ("_rows", "_norms", "_tags"):
            t = getattr(self, name)
            t[:m] = t.index_select(0, idx)
        for sh in self._shadows.values():
            sh.reset()                                 # rebuilt lazily by the next batched search
        remap = {int(old): new for new, old in enumerate(live.tolist())}
        self._row_id = {remap[r]: v for r, v in self._row_id.items() if r in remap}
        self._row_meta = {remap[r]: v for r, v in self._row_meta.items() if r in remap}
        self._id_rows = {k: [remap[r] for r in v if r in remap] for k, v in self._id_rows.items()}
        blocks = []
        for lo, hi, prefix, first in self._blocks:      # implicit-id blocks survive only if wholly live
            seg = live[(live >= lo) & (live < hi)]
            if seg.size == hi - lo:
                blocks.append((remap[lo], remap[lo] + (hi - lo), prefix, first))
            else:
                for r in seg.tolist():
                    self._row_id[remap[r]] = f"{prefix}{first + r - lo}"
                    self._id_rows.setdefault((int(self._row_user[r]), f"{prefix}{first + r - lo}"), []).append(remap[r])
        self._blocks = blocks
        self._gen += 1
        self._graphs.clear()
        self._row_user[:m] = self._row_user[live]
        self._row_user[m:self._n] = -1
        self._tags[m:self._n] = -1
        self._n = m
        for u in range(len(self._span)):
            rows = np.nonzero(self._row_user[:m] == u)[0]
            self._span[u] = [int(rows[0]), int(rows[-1]) + 1] if rows.size else [1 << 62, 0]

    def _implicit_rows(self, nid: str, u: int) -> List[int]:
        out = []
        for lo, hi, prefix, first in self._blocks:
            if nid.startswith(prefix) and nid[len(prefix):].isdigit():
                r = lo + int(nid[len(prefix):]) - first
                if lo <= r < hi and r not in self._row_id:
                    out.append(r)
        return out

    def get_latest_version(self) -> int:                             # vector_store.py:150-156
        return self._version

    # ------------------------------------------------------------------ edges (vector_store.py:159-212)
    def add_edges(self, edges: List[Dict[str, Any]], user_id: str = "default"):
        if not edges:
            return
        with self._lock:
            for edge in edges:
                src = edge.get("source") or edge.get("source_id")
                tgt = edge.get("target") or edge.get("target_id")
                self._edges.append({
                    "id": edge.get("id", f"{src}_{tgt}"), "user_id": user_id,
                    "source_id": src, "target_id": tgt,
                    "weight": float(np.float32(edge.get("weight", 1.0))),      # float32 column
                    "edge_type": edge.get("edge_type") or edge.get("type", "relates_to"),
                    "co_occurrence": edge.get("co_occurrence", 1),
                    "last_updated": edge.get("last_updated", 0.0),
                    "metadata": json.dumps(edge.get("metadata", {})),
                })

    def delete_edges(self, source_id: Optional[str] = None, user_id: str = "default"):
        with self._lock:
            self._edges = [e for e in self._edges
                           if not (e["user_id"] == user_id and (not source_id or e["source_id"] == source_id))]

    def get_edges(self, source_id: Optional[str] = None, user_id: str = "default") -> List[Dict[str, Any]]:
        with self._lock:
            out = []
            for e in self._edges:
                if e["user_id"] != user_id or (source_id and e["source_id"] != source_id):
                    continue
                d = dict(e)
                d["type"] = d["edge_type"]                            # vector_store.py:204-205
                d["metadata"] = json.loads(d["metadata"])
                out.append(d)
            return out

    # ------------------------------------------------------------------ profiles (vector_store.py:216-233)
    def save_profile(self, profile_data: Dict[str, Any], user_id: str = "default"):
        with self._lock:
            self._profiles[user_id] = {"data": json.dumps(profile_data), "updated_at": time.time()}

    def load_profile(self, user_id: str = "default") -> Optional[Dict[str, Any]]:
        with self._lock:
            p = self._profiles.get(user_id)
            return json.loads(p["data"]) if p else None

    def close(self):                                                  # vector_store.py:235-244
        with self._lock:
            self._rows = self._norms = self._tags = self._scratch = None
            self._bscratch.clear()
            self._one.clear()
            self._shadows.clear()
            self._pin.clear()
            self._io.clear()
            self._graphs.clear()
            self._n = 0

    # ------------------------------------------------------------------ legacy surface
    # tests/test_vector_store.py:15,43,46,66 and tests/test_lancedb_integration.py:67,79,84
    def add(self, nodes):
        return self.add_nodes(nodes)

    def search(self, query_emb=None, limit: int = 5):
        if query_emb is None:                                         # table-style use, memory_system.py:1438
            return _UserScan(self)
        return self.search_nodes(query_emb, limit=limit)

    def delete(self, ids):
        return self.delete_nodes(ids)


class _UserScan:
    """`store._nodes_table.search().select(["user_id"]).to_list()` (memory_system.py:1434-1439)."""

    def __init__(self, store: B200Store):
        self._s = store

    def select(self, cols):
        return self

    def to_list(self):
        s = self._s
        return [{"user_id": s._user_names[u]} for u in s._row_user[:s._n].tolist() if u >= 0]


# The reference's older name for the same thing (tests/test_vector_store.py:8).
LanceDBVectorStore = B200Store
