"""GPU parity: two-level semantic-shard search (north_star (1)) — centroid routing, then the member rows of the probed
shards, on the tensor cores (lz_search_two_level) and on the exact float64 kernels (lz_search_exact_ranges) — against
the oracle's restatement (oracle/flat_search.py two_level_topk): ids and probed shards bit-exact for every nprobe,
scores to 1e-12; nprobe = all shards equals the flat search."""
import numpy as np
import pytest
import torch

from oracle import flat_search

pytestmark = pytest.mark.gpu


def ragged_offsets(n, S, seed):
    rng = np.random.default_rng(seed)
    cuts = np.sort(rng.choice(np.arange(1, n), S - 1, replace=False))
    return np.concatenate([[0], cuts, [n]]).astype(np.int64)


def build(n, dim, S, seed, metric="cosine", ragged=False, scale=False):
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    rows, shard = synth.clustered_rows(n, dim, S, seed=seed)
    if scale:
        rows = rows * np.random.default_rng(seed + 1).uniform(0.5, 2.0, n).astype(np.float32)[:, None]
    off = ragged_offsets(n, S, seed + 2) if ragged else np.searchsorted(shard, np.arange(S + 1)).astype(np.int64)
    st = B200Store(metric=metric)
    st.add_rows(rows, id_prefix="r")
    st.set_shards(off)
    return st, rows, off


@pytest.mark.parametrize("n,dim,S,B,k,metric,ragged", [
    (40000, 128, 32, 64, 10, "cosine", False),
    (60000, 768, 64, 256, 10, "cosine", False),        # BASELINE configs[1] in small: 64 shards, batch 256, top-10
    (30000, 64, 17, 40, 5, "cosine", True),            # ragged shards, some smaller than a 256-row tile
    (50000, 96, 40, 130, 20, "l2", True),
    (9000, 72, 3, 9, 100, "cosine", True),
])
def test_tensor_core_two_level_matches_oracle(n, dim, S, B, k, metric, ragged):
    from lazzaro_b200 import synth
    st, rows, off = build(n, dim, S, seed=n + S, metric=metric, ragged=ragged, scale=(metric == "l2"))
    qs = synth.queries_near_rows(rows, B, seed=B)
    if metric == "l2":
        qs = qs * np.float32(1.2)
    qd = st._to_device(qs)
    sub = np.linspace(0, B - 1, min(B, 12)).astype(int)               # the numpy oracle checks a spread of the queries ...
    for nprobe in sorted({1, 3, max(1, S // 2), S}):
        ids, sc, cnt, probe = st.search_two_level_device(qd, k, nprobe, want_probe=True)
        torch.cuda.synchronize()
        want_ids, want_sc, want_probe = flat_search.two_level_topk(rows, off, qs[sub], k, nprobe, metric)
        np.testing.assert_array_equal(probe.cpu().numpy()[sub], want_probe)
        np.testing.assert_array_equal(ids.cpu().numpy()[sub], want_ids)
        got_sc = sc.cpu().numpy()[sub]
        ok = want_ids >= 0
        np.testing.assert_allclose(got_sc[ok], want_sc[ok], rtol=1e-12, atol=1e-300)
        np.testing.assert_array_equal(cnt.cpu().numpy()[sub], ok.sum(axis=1))
        # ... and every query is checked bit for bit against the exact float64 kernels over the same probed shards
        pc = probe.clamp(min=0)
        lo = torch.where(probe >= 0, st._shard_off[:-1][pc], torch.zeros_like(probe)).contiguous()
        hi = torch.where(probe >= 0, st._shard_off[1:][pc], torch.zeros_like(probe)).contiguous()
        from lazzaro_b200 import kernels as K
        eids, esc, ecnt = K.search_exact_ranges(st._rows, st._dim_p, st._norms, qd, lo, hi, k, st._metric, n_rows_total=st._n)
        assert torch.equal(ids, eids) and torch.equal(cnt, ecnt) and torch.equal(sc[ids >= 0], esc[eids >= 0])
        if nprobe == S:                                               # every shard visited: the flat search, bit for bit
            fids, fsc, _ = st.search_device(qd, 0, k)
            assert torch.equal(ids, fids) and torch.equal(sc[ids >= 0], fsc[fids >= 0])
        assert st.last_certified.float().mean().item() > 0.5


def test_small_batches_take_the_exact_kernels_and_agree():
    from lazzaro_b200 import synth
    st, rows, off = build(30000, 64, 20, seed=7, ragged=True)
    qs = synth.queries_near_rows(rows, 5, seed=8)
    qd = st._to_device(qs)
    for nprobe in (1, 4, 20):
        ids, sc, cnt, probe = st.search_two_level_device(qd, 10, nprobe, want_probe=True)
        want_ids, want_sc, want_probe = flat_search.two_level_topk(rows, off, qs, 10, nprobe)
        torch.cuda.synchronize()
        np.testing.assert_array_equal(probe.cpu().numpy(), want_probe)
        np.testing.assert_array_equal(ids.cpu().numpy(), want_ids)
        np.testing.assert_allclose(sc.cpu().numpy(), want_sc, rtol=1e-12)
        st.batched_search = "always"                                   # the same five queries through the tensor cores
        ids2, sc2, _ = st.search_two_level_device(qd, 10, nprobe)
        st.batched_search = "auto"
        assert torch.equal(ids, ids2) and torch.equal(sc, sc2)


def test_queries_concentrated_on_one_shard_and_empty_shards():
    """All queries probe the same shard (several q-tiles of one shard, the others unused); empty shards; duplicated rows
    across the probe boundary must not be scored twice."""
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    rows, shard = synth.clustered_rows(20000, 64, 8, seed=3)
    off = np.searchsorted(shard, np.arange(9)).astype(np.int64)
    off = np.concatenate([off[:4], [off[4], off[4]], off[5:]])          # shard 4 is empty (and one more shard overall)
    st = B200Store()
    st.add_rows(rows, id_prefix="r")
    st.set_shards(off)
    qs = synth.queries_near_rows(rows[off[1]:off[2]], 300, seed=4)      # every query near shard 1
    qd = st._to_device(qs)
    sub = np.arange(0, 300, 25)
    for nprobe in (1, 2, 9):
        ids, sc, cnt = st.search_two_level_device(qd, 10, nprobe)
        want_ids, want_sc, _ = flat_search.two_level_topk(rows, off, qs[sub], 10, nprobe)
        torch.cuda.synchronize()
        np.testing.assert_array_equal(ids.cpu().numpy()[sub], want_ids)
        if nprobe == 9:
            fids, fsc, _ = st.search_device(qd, 0, 10)
            assert torch.equal(ids, fids)


def test_adversarial_rounding_case_through_two_level():
    """The correlated-rounding fixture (tests/test_gpu_certify.py) split into shards: the deterministic margin must fall
    back here too, and the fallback scans exactly the probed shards."""
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    rows, q, b_row, info = synth.correlated_rounding_case()
    n = rows.shape[0]
    off = np.array([0, 1600, 3200, n], dtype=np.int64)
    st = B200Store(dim=768)
    st.add_rows(rows, id_prefix="r")
    st.set_shards(off)
    st.batched_search = "always"
    qs = np.concatenate([q[None], synth.queries_near_rows(rows, 8, seed=5)])
    for nprobe in (3, 2):
        ids, sc, cnt, probe = st.search_two_level_device(st._to_device(qs), 10, nprobe, want_probe=True)
        want_ids, want_sc, want_probe = flat_search.two_level_topk(rows, off, qs, 10, nprobe)
        torch.cuda.synchronize()
        np.testing.assert_array_equal(probe.cpu().numpy(), want_probe)
        np.testing.assert_array_equal(ids.cpu().numpy(), want_ids)
