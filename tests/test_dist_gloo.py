"""CPU, world_size 2, gloo: the gather / merge bookkeeping of the multi-GPU search (lazzaro_b200/dist.py).
The local search and the merge are injected numpy stand-ins (the oracle) — on a GPU box the same class
runs the CUDA kernels; here only the partitioning, id offsets, all-gather layout and merge order are
under test."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _worker(rank, world, port, n, dim, B, k, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from lazzaro_b200 import synth
    from lazzaro_b200.dist import ShardedSearch, even_slabs
    from oracle import flat_search
    rows, _ = synth.clustered_rows(n, dim, 8, seed=5)
    rows[100] = rows[n - 50]                       # a tie that straddles the two slabs: lower id must win
    qs = synth.queries_near_rows(rows, B, seed=6)
    qs[0] = rows[100]
    lo, hi = even_slabs(n, world)[rank]

    def local(queries, kk):                        # stand-in for the exact scan with id_base = slab offset
        ids, sc = flat_search.flat_topk_batch(rows[lo:hi], queries.numpy(), kk)
        ids = np.where(ids >= 0, ids + lo, ids)
        return torch.from_numpy(ids), torch.from_numpy(sc), torch.from_numpy((ids >= 0).sum(1).astype(np.int32))

    def merge(all_ids, all_sc):                    # stand-in for lz_merge_topk
        L, Bq, kk = all_ids.shape
        oi = np.full((Bq, kk), -1, np.int64); os_ = np.full((Bq, kk), np.nan)
        for b in range(Bq):
            i, s = flat_search.merge_topk(list(all_ids[:, b].numpy()), list(all_sc[:, b].numpy()), kk)
            oi[b, :i.size] = i; os_[b, :s.size] = s
        return torch.from_numpy(oi), torch.from_numpy(os_), torch.from_numpy((oi >= 0).sum(1).astype(np.int32))

    ss = ShardedSearch(store=None, row_offset=lo, world=world, local_search=local, merge=merge)
    ids, sc, cnt = ss.search_device(torch.from_numpy(qs), k)
    want_ids, want_sc = flat_search.flat_topk_batch(rows, qs, k)
    ok = np.array_equal(ids.numpy(), want_ids) and np.allclose(sc.numpy(), want_sc, rtol=1e-12)
    ok = ok and ids[0, 0].item() == 100 and ids[0, 1].item() == n - 50
    out[rank] = bool(ok)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_search_equals_single_search():
    world = 2
    port = 29500 + (os.getpid() % 2000)
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, port, 3000, 32, 5, 7, out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}


def test_partitioning_helpers():
    from lazzaro_b200.dist import even_slabs, partition_shards
    assert even_slabs(10, 3) == [(0, 3), (3, 6), (6, 10)]
    owned = partition_shards([5, 9, 1, 7, 3, 3], 2)
    assert sorted(sum(owned, [])) == list(range(6))
    loads = [sum([5, 9, 1, 7, 3, 3][s] for s in o) for o in owned]
    assert abs(loads[0] - loads[1]) <= 2                 # greedy largest-first keeps the slabs even
    assert partition_shards([4, 4], 4)[2:] == [[], []]


def test_tenant_routing_covers_every_query_once():
    """TenantSearch.owner / route (configs[4]: users dealt to ranks in contiguous blocks): pure host bookkeeping."""
    from lazzaro_b200.dist import TenantSearch

    class _S:                                  # stands in for a store: only user_spans() is read at construction
        def user_spans(self):
            return np.arange(5) * 10, np.arange(1, 6) * 10

    users = np.random.default_rng(0).integers(0, 20, 1000)
    parts = [TenantSearch(_S(), 5, r, 4).route(users) for r in range(4)]
    assert sorted(np.concatenate(parts).tolist()) == list(range(1000))
    for r, idx in enumerate(parts):
        assert (users[idx] // 5 == r).all()
    assert TenantSearch(_S(), 5, 0, 4).owner(np.array([0, 4, 5, 19, 25])).tolist() == [0, 0, 1, 3, 3]
