"""Seeded synthetic memories, queries and edge tables.

There is no network, so every embedding in tests and benchmarks is synthetic: unit-norm cluster
centres (one per semantic shard) with members `normalise(centre + sigma * g)`, `g ~ N(0, I/D)`,
which gives same-shard cosines near 0.89 with top-10 gaps of 1e-3..1e-4 — tight enough to exercise
a bf16 coarse pass.  Row ids are insertion indices, so "ties by id" equals the reference's
"ties by insertion order" (memory_system.py:470,827,865 use stable sorts).

The numpy generators are used at the sizes the CPU oracle can check; the torch generators build the
same distribution directly in HBM for the full-size benchmark configurations.
"""
from __future__ import annotations

import numpy as np

SEED_DATA, SEED_QUERY, SEED_EDGES = 1234, 4321, 777
SIGMA = 0.35


def _unit(x):
    return x / np.linalg.norm(x, axis=-1, keepdims=True)


def clustered_rows(n: int, dim: int, n_shards: int, seed: int = SEED_DATA, sigma: float = SIGMA):
    """(rows float32 [n, dim], shard int32 [n]); rows are stored shard by shard (each shard a
    contiguous row range), shards as equal as n allows."""
    rng = np.random.default_rng(seed)
    centres = _unit(rng.standard_normal((n_shards, dim)))
    shard = (np.arange(n, dtype=np.int64) * n_shards // max(n, 1)).astype(np.int32)
    rows = np.empty((n, dim), dtype=np.float32)
    step = 1 << 16
    for s in range(0, n, step):
        g = rng.standard_normal((min(step, n - s), dim)) / np.sqrt(dim)
        rows[s:s + step] = _unit(centres[shard[s:s + step]] + sigma * g).astype(np.float32)
    return rows, shard


def queries_near_rows(rows: np.ndarray, n_queries: int, seed: int = SEED_QUERY, noise: float = 0.1):
    """normalise(random member row + noise * g'), float32 [n_queries, dim]."""
    rng = np.random.default_rng(seed)
    pick = rng.integers(0, rows.shape[0], size=n_queries)
    g = rng.standard_normal((n_queries, rows.shape[1])) / np.sqrt(rows.shape[1])
    return _unit(rows[pick].astype(np.float64) + noise * g).astype(np.float32)


def near_duplicates(rows: np.ndarray, n: int, seed: int = SEED_QUERY + 1, noise: float = 0.25):
    """(dups float32 [n, dim], source row index) with cosine ≈ 0.97 to their source row."""
    rng = np.random.default_rng(seed)
    pick = rng.integers(0, rows.shape[0], size=n)
    g = rng.standard_normal((n, rows.shape[1])) / np.sqrt(rows.shape[1])
    return _unit(rows[pick].astype(np.float64) + noise * g).astype(np.float32), pick


def edge_table(n_edges: int, n_nodes: int, seed: int = SEED_EDGES):
    """Edge SoA: src/dst int32 uniform over nodes, weight float32 ~ U[0,1), co_occurrence = 1."""
    rng = np.random.default_rng(seed)
    src = rng.integers(0, n_nodes, size=n_edges, dtype=np.int32)
    dst = rng.integers(0, n_nodes, size=n_edges, dtype=np.int32)
    w = rng.random(n_edges, dtype=np.float32)
    cooc = np.ones(n_edges, dtype=np.int32)
    return src, dst, w, cooc


# ---- device-side generators (full-size benchmark configurations) ---------------------------------

def clustered_rows_device(n: int, dim: int, n_shards: int, device, seed: int = SEED_DATA,
                          sigma: float = SIGMA, out=None, chunk: int = 1 << 18):
    """Same distribution as `clustered_rows`, generated in HBM with torch's CUDA generator (a
    different random stream from the numpy one — full-size runs are checked by properties and by
    the exact fp64 scan, not against numpy fixtures).  Returns (rows float32 [n, dim] on device,
    shard int32 [n] on device)."""
    import torch
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    centres = torch.randn(n_shards, dim, generator=gen, device=device, dtype=torch.float32)
    centres = centres / centres.norm(dim=1, keepdim=True)
    rows = out if out is not None else torch.empty(n, dim, device=device, dtype=torch.float32)
    shard = (torch.arange(n, device=device, dtype=torch.int64) * n_shards // max(n, 1)).to(torch.int32)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        g = torch.randn(e - s, dim, generator=gen, device=device, dtype=torch.float32) * (sigma / dim ** 0.5)
        g += centres[shard[s:e].long()]
        g /= g.norm(dim=1, keepdim=True)
        rows[s:e] = g
    return rows, shard


def queries_near_rows_device(rows, n_queries: int, seed: int = SEED_QUERY, noise: float = 0.1):
    import torch
    gen = torch.Generator(device=rows.device)
    gen.manual_seed(seed)
    pick = torch.randint(0, rows.shape[0], (n_queries,), generator=gen, device=rows.device)
    g = torch.randn(n_queries, rows.shape[1], generator=gen, device=rows.device, dtype=torch.float32)
    q = rows[pick] + g * (noise / rows.shape[1] ** 0.5)
    return q / q.norm(dim=1, keepdim=True)
