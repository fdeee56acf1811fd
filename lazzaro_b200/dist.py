"""Multi-GPU search: rows partitioned across the GPUs of one box, per-GPU top-k merged after an
all-gather.

One process per GPU (torchrun); rank r holds whole semantic shards — a contiguous slab of global row
ids [row_offset, row_offset + n_local).  A search broadcasts nothing (every rank is handed the same
query batch), runs the local exact scan with `id_base = row_offset` so the ids it emits are already
global, all-gathers the (ids int64, scores float64) lists — B*k*16 bytes per rank over NVLink — and
merges them on every rank with lz_merge_topk under the same (score desc, id asc) order, which makes
the merge associative and the result identical to a single-GPU search.  Decay, prune and shard
assignment are shard-local and need no collective.

The reference has no distributed code at all; this module is the data-parallel layout
north_star asks for.  `local_search` / `merge` are injectable so the gather/merge bookkeeping can be
tested on CPU with gloo (tests/test_dist_gloo.py) — the product path always uses the CUDA kernels.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist


def partition_shards(shard_rows: Sequence[int], world: int) -> List[List[int]]:
    """Greedy bin-packing of semantic shards onto GPUs by row count (largest first); returns the
    shard indices owned by each rank.  Each shard lives wholly on one GPU."""
    order = sorted(range(len(shard_rows)), key=lambda s: (-shard_rows[s], s))
    load = [0] * world
    owned: List[List[int]] = [[] for _ in range(world)]
    for s in order:
        r = min(range(world), key=lambda i: (load[i], i))
        owned[r].append(s)
        load[r] += shard_rows[s]
    return [sorted(o) for o in owned]


def even_slabs(n_rows: int, world: int) -> List[Tuple[int, int]]:
    """[lo, hi) global row ranges of equal size (the last rank takes the remainder)."""
    per = n_rows // world
    return [(r * per, (r + 1) * per if r < world - 1 else n_rows) for r in range(world)]


class ShardedSearch:
    def __init__(self, store, row_offset: int = 0, world: Optional[int] = None, group=None,
                 local_search: Optional[Callable] = None, merge: Optional[Callable] = None):
        self.store = store
        self.row_offset = int(row_offset)
        self.group = group
        self.world = world if world is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        self._local = local_search or self._local_cuda
        self._merge = merge or self._merge_cuda
        self._packed_merge = merge is None            # product path: lz_merge_topk_strided straight from the gathered block
        self._gather = self._part = self._cnt = None
        self._lanes = {}                              # lane -> {"stream", "done", buffers}: searches kept in flight

    # -- the product path
    def _local_cuda(self, queries: torch.Tensor, k: int, out=None, lane: int = 0):
        ids, sc, cnt = self.store.search_device(queries, 0, k, id_base=self.row_offset, out=out, lane=lane)
        return ids, sc, cnt

    def _merge_cuda(self, ids: torch.Tensor, scores: torch.Tensor):
        from . import kernels
        return kernels.merge_topk(ids, scores, self.store._metric)

    def search_device(self, queries: torch.Tensor, k: int, lane: Optional[int] = None):
        """queries float32 [B, dim_p] on this rank's GPU (same batch on every rank) -> global
        (ids int64 [B,k], scores float64 [B,k], counts int32 [B]) on every rank.

        lane = None: everything is enqueued on the current stream.  lane = 0 / 1 / ...: the search runs on that lane's
        own stream with its own scratch and exchange buffers and returns at once; consecutive searches on DIFFERENT lanes
        overlap on the GPU — the exact pass, all-gather and merge of one search run while the next search's coarse pass
        (which holds every SM's shared memory but leaves room for the small exact-pass CTAs) is under way, which hides
        the per-search fixed costs that do not shrink when the rows are split over more GPUs.  Results of a lane are
        valid after join() (or after that lane's next call has been ordered behind them on its stream).

        The exchange is ONE all-gather per search: each rank's local search writes its ids and its scores' bits
        into one int64 block [2, B, k], the blocks are gathered into [world, 2, B, k], and the merge kernel reads
        the gathered block in place (lz_merge_topk_strided)."""
        if lane is not None:
            return self._search_on_lane(queries, k, int(lane))
        if self.world == 1:
            return self._local(queries, k)
        B, dev = queries.shape[0], queries.device
        if self._gather is None or self._gather.shape[2:] != (B, k) or self._gather.device != dev:
            self._part = torch.empty((2, B, k), dtype=torch.int64, device=dev)
            self._gather = torch.empty((self.world, 2, B, k), dtype=torch.int64, device=dev)
            self._cnt = torch.empty((B,), dtype=torch.int32, device=dev)
        part = self._part
        if self._local == self._local_cuda:
            self._local(queries, k, out=(part[0], part[1].view(torch.float64), self._cnt))    # written in place
        else:                                                # injected local search (CPU tests of the bookkeeping)
            ids, sc, _ = self._local(queries, k)
            part[0].copy_(ids)
            part[1].copy_(sc.contiguous().view(torch.int64))
        self._all_gather(self._gather, part)
        if self._packed_merge:
            from . import kernels
            return kernels.merge_topk_packed(self._gather, self.store._metric)
        return self._merge(self._gather[:, 0].contiguous(), self._gather[:, 1].contiguous().view(torch.float64))

    def _search_on_lane(self, queries: torch.Tensor, k: int, lane: int):
        dev = queries.device
        L = self._lanes.get(lane)
        if L is None:
            L = self._lanes[lane] = {"stream": torch.cuda.Stream(device=dev), "done": None, "shape": None}
        B = queries.shape[0]
        ready = torch.cuda.Event()
        ready.record()                                       # the queries (and anything before) on the caller's stream
        with torch.cuda.stream(L["stream"]):
            L["stream"].wait_event(ready)
            if L["shape"] != (B, k):
                L["part"] = torch.empty((2, B, k), dtype=torch.int64, device=dev)
                L["gather"] = torch.empty((self.world, 2, B, k), dtype=torch.int64, device=dev)
                L["cnt"] = torch.empty((B,), dtype=torch.int32, device=dev)
                L["shape"] = (B, k)
            part = L["part"]
            out = self._local_cuda(queries, k, out=(part[0], part[1].view(torch.float64), L["cnt"]), lane=lane)
            if self.world > 1:
                self._all_gather(L["gather"], part)
                from . import kernels
                out = kernels.merge_topk_packed(L["gather"], self.store._metric)
            L["done"] = torch.cuda.Event()
            L["done"].record()
        return out

    def join(self):
        """The current stream waits for every lane's last search."""
        for L in self._lanes.values():
            if L["done"] is not None:
                torch.cuda.current_stream().wait_event(L["done"])

    def _all_gather(self, out: torch.Tensor, part: torch.Tensor):
        if dist.get_backend(self.group) == "gloo":           # CPU tests of the bookkeeping
            dist.all_gather(list(out.unbind(0)), part.contiguous(), group=self.group)
        else:
            dist.all_gather_into_tensor(out, part.contiguous(), group=self.group)

    def search_host(self, queries: np.ndarray, k: int):
        """Host queries in, host (ids, scores, counts) out: pinned H2D, search, D2H, one sync."""
        st = self.store
        q = np.ascontiguousarray(queries, dtype=np.float32)
        B = q.shape[0]
        with torch.cuda.device(st.device):
            stage = st._pinned("q", (B, st._dim_p), torch.float32)
            if st._dim_p != st.dim:
                stage.zero_()
            stage[:, :st.dim] = torch.from_numpy(q)
            qd = torch.empty((B, st._dim_p), dtype=torch.float32, device=st.device)
            qd.copy_(stage, non_blocking=True)
            ids, sc, cnt = self.search_device(qd, k)
            h_ids = st._pinned("ids", (B, k), torch.int64)
            h_sc = st._pinned("sc", (B, k), torch.float64)
            h_cnt = st._pinned("cnt", (B,), torch.int32)
            h_ids.copy_(ids, non_blocking=True)
            h_sc.copy_(sc, non_blocking=True)
            h_cnt.copy_(cnt, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        return h_ids.numpy().copy(), h_sc.numpy().copy(), h_cnt.numpy().copy()

    # -- pipelined host interface: the next batch's staging + H2D copy overlap the current batch's search
    def submit_host(self, queries: np.ndarray, k: int, slot: int = 0) -> "_PendingSearch":
        """Starts a search of host queries without waiting for it: stage into pinned memory, H2D on a copy
        stream, the search and the D2H of the result on the slot's lane.  `slot` (0..3) picks a set of staging
        buffers — resubmit a slot only after its previous result() was taken; consecutive submissions alternate between
        the two lanes.
        Same answer as search_host; a loop that keeps one or two searches in flight hides the copies and the host
        work behind them (two absorb the jitter of the other ranks' hosts when the search has a collective in it)."""
        st = self.store
        q = np.ascontiguousarray(queries, dtype=np.float32)
        B = q.shape[0]
        with torch.cuda.device(st.device):
            if not hasattr(self, "_copy_stream"):
                self._copy_stream = torch.cuda.Stream(device=st.device)
                self._qdev = {}
            stage = st._pinned(f"q{slot}", (B, st._dim_p), torch.float32)
            if st._dim_p != st.dim:
                stage.zero_()
            stage[:, :st.dim] = torch.from_numpy(q)
            key = (slot, B)
            if key not in self._qdev:
                self._qdev[key] = torch.empty((B, st._dim_p), dtype=torch.float32, device=st.device)
            qd = self._qdev[key]
            main = torch.cuda.current_stream()
            # no wait on `main` here: the slot's previous search finished before its result() returned, and waiting
            # on the stream would queue this copy behind the search that is running now
            with torch.cuda.stream(self._copy_stream):
                qd.copy_(stage, non_blocking=True)
                ready = torch.cuda.Event()
                ready.record()
            main.wait_event(ready)
            # the search itself runs on the slot's lane (own stream, own scratch): with two slots alternating, the exact
            # pass, exchange and read-back of one search overlap the coarse pass of the next
            lane = self._host_seq = (getattr(self, "_host_seq", 1) + 1) & 1      # lanes alternate by submission order
            ids, sc, cnt = self.search_device(qd, k, lane=lane)
            h_ids = st._pinned(f"ids{slot}", (B, k), torch.int64)
            h_sc = st._pinned(f"sc{slot}", (B, k), torch.float64)
            h_cnt = st._pinned(f"cnt{slot}", (B,), torch.int32)
            with torch.cuda.stream(self._lanes[lane]["stream"]):
                h_ids.copy_(ids, non_blocking=True)
                h_sc.copy_(sc, non_blocking=True)
                h_cnt.copy_(cnt, non_blocking=True)
                done = torch.cuda.Event()
                done.record()
        return _PendingSearch(done, h_ids, h_sc, h_cnt)


class _PendingSearch:
    def __init__(self, done, h_ids, h_sc, h_cnt):
        self._done, self._bufs = done, (h_ids, h_sc, h_cnt)

    def result(self):
        """Waits for the search and returns host (ids int64 [B,k], scores float64 [B,k], counts int32 [B])."""
        self._done.synchronize()
        return tuple(b.numpy().copy() for b in self._bufs)


PARITY_CASES = ((1, 10), (3, 100), (64, 10), (300, 7))     # (batch, k): exact-scan path and tensor-core path


def parity_check(n: int = 300_000, dim: int = 128, cases=PARITY_CASES, group=None):
    """Run on every rank of an initialised NCCL job (one rank per GPU): builds the SAME n x dim rows on each rank,
    gives rank r the r-th slab, and checks that the sharded search (local search with id_base, one packed
    all-gather, strided merge) returns bit for bit what a single-GPU lz_search_exact over all rows returns —
    ids, float64 scores and counts, through the device API and through the pipelined host API.  Queries are
    drawn near rows of the whole matrix, so true neighbours sit on every rank.
    Returns a list of {"batch", "k", "ok"} (identical on every rank: the verdicts are all-reduced)."""
    from . import _lib, kernels as K, synth
    from .store import B200Store
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = torch.device("cuda", torch.cuda.current_device())
    rows, _ = synth.clustered_rows_device(n, dim, 32, dev)          # same seed on every rank: same rows
    lo, hi = even_slabs(n, world)[rank]
    st = B200Store(device=str(dev))
    st.adopt_rows(rows[lo:hi].contiguous(), id_prefix="r")
    ss = ShardedSearch(st, row_offset=lo, world=world, group=group)
    norms = K.row_norms(rows)
    out = []
    # every case as the store would route it at this size, then the small-batch cases once more through the fp16 shadow
    # (what single queries and batches below 8 take on spans of 500k rows or more): same bits either way
    runs = [(B, k, None) for B, k in cases] + [(B, k, 1) for B, k in cases if B < st.min_batch and k <= _lib.BATCHED_MAX_K]
    for B, k, small_rows in runs:
        st.min_rows_small_batch = small_rows if small_rows is not None else 500_000
        q = synth.queries_near_rows_device(rows, B, seed=synth.SEED_QUERY + B)
        ids, sc, cnt = ss.search_device(q, k)
        want_ids, want_sc, want_cnt = K.search_exact(rows, dim, norms, None, q, k, _lib.COSINE)
        good = bool((ids == want_ids).all()) and bool((sc == want_sc).all()) and bool((cnt == want_cnt).all())
        h = ss.submit_host(q.cpu().numpy(), k).result()
        good = good and np.array_equal(h[0], want_ids.cpu().numpy())
        if B >= 2 * world:                                         # the answers really are spread over the ranks
            owners = torch.bincount((want_ids[want_ids >= 0] // max(1, n // world)).clamp(max=world - 1), minlength=world)
            good = good and bool((owners > 0).all())
        flag = torch.tensor([int(good)], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        out.append({"batch": B, "k": k, "path": "shadow" if (small_rows is not None or B >= st.min_batch) else "exact scan",
                    "ok": bool(int(flag))})
    st.close()
    return out


class TenantSearch:
    """Multi-tenant search across the GPUs of one box (BASELINE configs[4]; SURVEY §8e "partition by user").

    All of a user's rows live on ONE rank, so the per-user pre-filter of the reference (vector_store.py:136-139:
    `.where(user_id = ...)` before the top-k) makes a query a single-GPU operation: the host routes it to the rank that
    owns the user and that rank scans just the user's row span.  No collective is on the data path; throughput scales
    with the number of ranks because the tenants (and their queries) do.  Users are dealt to ranks in contiguous blocks
    of `users_per_rank` global user indices; this rank's store holds its block as local user indices 0..users_per_rank-1."""

    def __init__(self, store, users_per_rank: int, rank: int = 0, world: int = 1):
        self.store, self.users_per_rank, self.rank, self.world = store, int(users_per_rank), int(rank), int(world)
        self._lo, self._hi = store.user_spans()

    def owner(self, global_user: np.ndarray) -> np.ndarray:
        """Rank that owns each global user index."""
        return np.minimum(np.asarray(global_user, dtype=np.int64) // self.users_per_rank, self.world - 1)

    def route(self, global_user: np.ndarray) -> np.ndarray:
        """Positions (into the caller's batch) of the queries this rank must answer."""
        return np.nonzero(self.owner(global_user) == self.rank)[0]

    def prepare(self, local_user: np.ndarray):
        """Device-resident per-query row spans and tags for a batch of local user indices (reusable across searches)."""
        dev = self.store.device
        lo_h, hi_h = self._lo[local_user], self._hi[local_user]
        lo = torch.as_tensor(lo_h).to(dev)
        hi = torch.as_tensor(hi_h).to(dev)
        lo._lz_max_span = int((hi_h - lo_h).max()) if len(lo_h) else 0
        return lo, hi, torch.as_tensor(np.asarray(local_user, dtype=np.int32)).to(dev)

    def search_device(self, queries: torch.Tensor, local_user, k: int):
        """queries float32 [b, dim_p] on this GPU, local_user int [b] (or what prepare() returned for it) ->
        (row ids int64 [b,k], scores, counts), unsynchronised."""
        lo, hi, tag = local_user if isinstance(local_user, tuple) else self.prepare(local_user)
        return self.store.search_ranges_device(queries, lo, hi, tag, k)

    def submit_host(self, queries: np.ndarray, global_user: np.ndarray, k: int, slot: int = 0) -> "_PendingTenantSearch":
        """Starts the routed part of a host batch without waiting for it: the rows of `queries` this rank must answer are
        gathered into pinned memory (torch's threaded index_select), copied H2D on a copy stream, searched, and the ids and
        counts copied back; result() waits.  `slot` (0..3) picks a set of staging buffers — resubmit a slot only after
        its previous result() was taken.  A loop that keeps one or two batches in flight hides the host-side gather
        (12 MB for 4096 queries of 768 floats) and the copies behind the scan of the previous batch."""
        st = self.store
        mine = self.route(global_user)
        if mine.size == 0:
            return _PendingTenantSearch(None, mine, np.empty((0, k), np.int64), np.empty(0, np.int32))
        m = int(mine.size)
        with torch.cuda.device(st.device):
            if not hasattr(self, "_copy_stream"):
                self._copy_stream = torch.cuda.Stream(device=st.device)
                self._qdev = {}
            stage = st._pinned(f"tq{slot}", (m, st._dim_p), torch.float32)
            sn = stage.numpy()
            src = np.asarray(queries)
            if st._dim_p != st.dim:
                sn[:, st.dim:] = 0.0
                sn[:, :st.dim] = src[mine]
            elif src.dtype == np.float32 and src.ndim == 2 and src.shape[1] == st.dim and src.flags.c_contiguous:
                stage.copy_(torch.index_select(torch.from_numpy(src), 0, torch.from_numpy(mine)))   # threaded gather + copy
            else:
                sn[:] = src[mine]
            key = (slot, m)
            if key not in self._qdev:
                self._qdev[key] = torch.empty((m, st._dim_p), dtype=torch.float32, device=st.device)
            qd = self._qdev[key]
            main = torch.cuda.current_stream()
            with torch.cuda.stream(self._copy_stream):
                qd.copy_(stage, non_blocking=True)
                ready = torch.cuda.Event()
                ready.record()
            main.wait_event(ready)
            local = (np.asarray(global_user, dtype=np.int64)[mine] - self.rank * self.users_per_rank)
            ids, sc, cnt = self.search_device(qd, local, k)
            h_ids = st._pinned(f"tids{slot}", tuple(ids.shape), torch.int64)
            h_cnt = st._pinned(f"tcnt{slot}", tuple(cnt.shape), torch.int32)
            h_ids.copy_(ids, non_blocking=True)
            h_cnt.copy_(cnt, non_blocking=True)
            done = torch.cuda.Event()
            done.record()
        return _PendingTenantSearch(done, mine, h_ids, h_cnt)

    def search_host(self, queries: np.ndarray, global_user: np.ndarray, k: int):
        """The routed part of a host batch: (positions answered here, row ids int64 [m,k] as numpy, counts).  Every rank
        calls this with the same batch; the union of the positions is the whole batch."""
        return self.submit_host(queries, global_user, k, slot=0).result()


class _PendingTenantSearch:
    def __init__(self, done, mine, h_ids, h_cnt):
        self._done, self._mine, self._ids, self._cnt = done, mine, h_ids, h_cnt

    def result(self):
        """Waits for the search: (positions answered here, row ids int64 [m,k], counts int32 [m]) as numpy."""
        if self._done is None:
            return self._mine, self._ids, self._cnt
        self._done.synchronize()
        return self._mine, self._ids.numpy().copy(), self._cnt.numpy().copy()


class NcclExchange:
    """The exchange step through the C-ABI (lz_allgather_merge) instead of torch.distributed: what a non-Python binding of
    include/lazzaro_b200.h would do.  Owns an ncclComm_t created with the NCCL library already loaded in this process
    (the unique id travels through the initialised torch.distributed group, which is used for nothing else here)."""

    def __init__(self, group=None):
        import ctypes as C
        self._C = C
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self._nccl = C.CDLL("libnccl.so.2")

        class _Uid(C.Structure):
            _fields_ = [("internal", C.c_char * 128)]
        uid = _Uid()
        if self.rank == 0:
            rc = self._nccl.ncclGetUniqueId(C.byref(uid))
            assert rc == 0, f"ncclGetUniqueId failed ({rc})"
        box = [C.string_at(C.byref(uid), 128) if self.rank == 0 else None]      # all 128 bytes (the field access stops at a NUL)
        dist.broadcast_object_list(box, src=0, group=group)
        C.memmove(C.byref(uid), box[0], 128)
        self.comm = C.c_void_p()
        self._nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, _Uid, C.c_int]
        rc = self._nccl.ncclCommInitRank(C.byref(self.comm), self.world, uid, self.rank)
        assert rc == 0, f"ncclCommInitRank failed ({rc})"

    def allgather_merge(self, part: torch.Tensor, metric: int):
        """part int64 [2, B, k] (ids | score bits) on this GPU -> (ids, scores, counts) merged over all ranks."""
        from . import _lib
        from ._lib import check, ptr, stream_ptr
        dev = part.device
        _, B, k = part.shape
        gathered = torch.empty((self.world, 2, B, k), dtype=torch.int64, device=dev)
        ids = torch.empty((B, k), dtype=torch.int64, device=dev)
        sc = torch.empty((B, k), dtype=torch.float64, device=dev)
        cnt = torch.empty((B,), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            check(_lib.load().lz_allgather_merge(self.comm, self.world, ptr(part), ptr(gathered), B, k, int(metric), ptr(ids), ptr(sc),
                                                 ptr(cnt), stream_ptr(dev)), "lz_allgather_merge")
        self._keep = gathered                          # alive until the stream has consumed it
        return ids, sc, cnt

    def close(self):
        if self.comm:
            self._nccl.ncclCommDestroy.argtypes = [self._C.c_void_p]
            self._nccl.ncclCommDestroy(self.comm)
            self.comm = None
