"""Seeded synthetic memories, queries and edge tables.

There is no network, so every embedding in tests and benchmarks is synthetic: unit-norm cluster
centres (one per semantic shard) with members `normalise(centre + sigma * g)`, `g ~ N(0, I/D)`,
which gives same-shard cosines near 0.89 with top-10 gaps of 1e-3..1e-4 — tight enough to exercise
a 16-bit coarse pass.  Row ids are insertion indices, so "ties by id" equals the reference's
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


# ---- adversarial fixture: fp16 rounding errors that line up ------------------------------------------

def correlated_rounding_case(n_high: int = 16, n_low_groups: int = 600, seed: int = 99):
    """A 768-d case on which the STATISTICAL certification margin of the tensor-core search is wrong.

    Returns (rows float32 [n, 768], query float32 [768], b_row, info).  Every row but one has unit norm and
    components that are small dyadic rationals (a/1024, |a| < 64), so its fp16 shadow and its coarse score are
    EXACT:
      * `n_high` rows (first group-aligned block) with cosine H = 30808/32768 = 0.940186 to the query,
      * 8 * `n_low_groups` rows with cosine L = 30803/32768 = 0.940033,
    and row `b_row` (alone in the last group, padded with copies of a far-away row) has exact cosine ~0.94033 > H
    but every component that meets the query sits just BELOW an fp16 rounding midpoint low in its binade, so all
    those rounding errors have the same sign and its coarse score is ~0.93995 < L.  With more L groups than the
    candidate buffer keeps, B is dropped; the k-th exact score H exceeds the drop threshold L by 1.5e-4, more than
    the statistical margin (1.1e-4 for this query) but less than B's true error (3.8e-4).  The deterministic
    margin (measured rounding-error norms) refuses to certify and the fallback finds B.
    """
    rng = np.random.default_rng(seed)
    dim, S1, S2 = 768, 192, 256                       # query: 2/32 on S1 dims, 1/32 on the next S2 dims, 0 elsewhere
    q = np.zeros(dim)
    q[:S1], q[S1:S1 + S2] = 2 / 32, 1 / 32            # |q|^2 = 192/256 + 256/1024 = 1
    free = np.arange(S1 + S2, dim)
    UNIT = 1024

    def dyadic_row(numer):
        # a/1024 with integer a: sum a^2 = 1024^2 and 2*sum_S1 a + sum_S2 a = numer (cosine = numer/32768)
        for _ in range(1000):
            a = np.zeros(dim, dtype=np.int64)
            a[:S1], a[S1:S1 + S2] = 60, 30
            add = numer - (2 * int(a[:S1].sum()) + int(a[S1:S1 + S2].sum()))
            n1 = int(rng.integers(0, min(S1, add // 2) + 1)) if add > 0 else 0   # S1 entries raised by one (2 each)
            n2 = add - 2 * n1                                                      # S2 entries raised by one (1 each)
            if add < 0 or n2 < 0 or n2 > S2:
                continue
            a[rng.choice(S1, n1, replace=False)] += 1
            a[S1 + rng.choice(S2, n2, replace=False)] += 1
            vals = np.full(free.size, 19, dtype=np.int64)
            rest = UNIT * UNIT - int((a * a).sum()) - int((vals * vals).sum())   # settle with 19 -> 20 (+39) / 19 -> 18 (-37)
            done = False
            for y in range(0, 39):
                if rest + 37 * y >= 0 and (rest + 37 * y) % 39 == 0 and (rest + 37 * y) // 39 + y <= free.size:
                    x = (rest + 37 * y) // 39
                    idx = rng.permutation(free.size)
                    vals[idx[:x]] = 20
                    vals[idx[x:x + y]] = 18
                    done = True
                    break
            if not done:
                continue
            a[free] = vals * rng.choice([-1, 1], free.size)
            assert int((a * a).sum()) == UNIT * UNIT and 2 * a[:S1].sum() + a[S1:S1 + S2].sum() == numer
            return a / float(UNIT)
        raise RuntimeError("no dyadic row found")

    H, L = 30808, 30803
    n_high = (n_high + 7) // 8 * 8
    rows = [dyadic_row(H) for _ in range(n_high)] + [dyadic_row(L) for _ in range(8 * n_low_groups)]
    # B: (m + 0.45) ulp with m = 1100 (low in the binade [1024, 2048), where half an ulp is the largest relative
    # error) on every component the query touches: each rounds DOWN by 4.1e-4 of its value
    b = np.zeros(dim)
    b[:S1] = 1100.45 * 2.0 ** -14                     # binade [2^-4, 2^-3): ulp 2^-14
    b[S1:S1 + S2] = 1100.45 * 2.0 ** -16              # binade [2^-6, 2^-5): ulp 2^-16
    rest = 1.0 - float((b ** 2).sum())
    b[free] = np.sqrt(rest / free.size) * rng.choice([-1, 1], free.size)    # |b| = 1; these never meet the query
    far = -q                                                               # filler for B's group: cosine -1
    b_row = len(rows)
    rows += [b] + [far] * 7
    rows = np.asarray(rows, dtype=np.float32)
    info = {"H": H / 32768, "L": L / 32768, "b_cosine": float(b @ q / np.linalg.norm(b)),
            "n_high": n_high, "n_low": 8 * n_low_groups}
    return rows, q.astype(np.float32), b_row, info
