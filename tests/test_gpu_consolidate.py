"""GPU parity: consolidation steps (dedup, linking, shard assignment, centroids) against outputs of the
reference's own code (tests/golden) and the numpy oracle."""
import json

import numpy as np
import pytest
import torch

from oracle import graph_ops

pytestmark = pytest.mark.gpu


def make_store(rows, ids, **kw):
    from lazzaro_b200.store import B200Store
    st = B200Store(**kw)
    st.add_nodes([{"id": i, "content": i, "embedding": r.tolist()} for i, r in zip(ids, rows)])
    return st


def test_dedup_matches_reference_consolidation(golden):
    # memory_system.py:719-742 as run by the reference's _async_consolidate (tests/golden/make_golden.py)
    from lazzaro_b200 import consolidate
    g = golden
    n_graph = g["dd_rows"].shape[0]
    st = make_store(g["dd_rows"], [f"g{i}" for i in range(n_graph)])
    names, sims, dup = consolidate.dedup_against_store(st, g["dd_new"])
    access = np.zeros(n_graph, np.int64)
    sal = np.full(n_graph, 0.5)
    for i in np.nonzero(dup)[0]:                       # the reference's merge rule (memory_system.py:735-741)
        j = int(names[i][1:])
        access[j] += 1
        sal[j] = max(sal[j], g["dd_sal_new"][i])
    np.testing.assert_array_equal(access, g["dd_access"])
    np.testing.assert_array_equal(sal, g["dd_sal_after"])
    assert n_graph + int((~dup).sum()) == int(g["dd_nodes_after"])
    best, osim, odup = graph_ops.dedup_matches(g["dd_new"], g["dd_rows"])
    assert [int(x[1:]) for x in names] == best.tolist() and dup.tolist() == odup.tolist()
    np.testing.assert_allclose(sims, osim, rtol=1e-12)


def test_cross_linking_matches_reference(golden):
    # memory_system.py:838-891
    from lazzaro_b200 import consolidate
    g = golden
    st = make_store(g["ln_rows"], [f"e{i}" for i in range(g["ln_rows"].shape[0])])
    links = consolidate.link_candidates(st, g["ln_new"])
    got = [(f"x{i}", tid, w) for i, lst in enumerate(links) for tid, s, w in lst]
    ref = json.loads(str(g["ln_cross"]))
    assert [(a, b) for a, b, _, _ in ref] == [(a, b) for a, b, _ in got]
    np.testing.assert_allclose([w for _, _, w, _ in ref], [w for _, _, w in got], rtol=1e-12)


def test_within_shard_linking_matches_reference(golden):
    # memory_system.py:816-836: the same scan restricted to the shard's rows (stored contiguously)
    from lazzaro_b200 import consolidate
    g = golden
    rows, shard, new, new_shard = g["ln_rows"], g["ln_shard"], g["ln_new"], g["ln_new_shard"]
    st = make_store(rows, [f"e{i}" for i in range(rows.shape[0])])
    ref = {(a, b): w for a, b, w, _ in json.loads(str(g["ln_within"])) if b.startswith("e")}
    got = {}
    for s in np.unique(new_shard):
        members = np.nonzero(shard == s)[0]
        assert members[-1] - members[0] + 1 == members.size           # contiguous by construction
        ids = np.nonzero(new_shard == s)[0]
        links = consolidate.link_candidates(st, new[ids], row_range=(int(members[0]), int(members[-1]) + 1))
        for a, lst in zip(ids, links):
            for tid, sim, w in lst:
                got[(f"x{a}", tid)] = w
    assert set(got) == set(ref)
    for key, w in ref.items():
        assert got[key] == pytest.approx(w, rel=1e-12)


def test_centroid_matches_reference(golden):
    # memory_system.py:916-917
    from lazzaro_b200 import consolidate
    rows = torch.as_tensor(golden["ce_rows"]).cuda()
    seg = torch.tensor([0, rows.shape[0]], dtype=torch.int64, device="cuda")
    got = consolidate.centroid_means(rows, seg_offsets=seg).cpu().numpy()[0]
    np.testing.assert_allclose(got, golden["ce_centroid"], rtol=1e-13, atol=1e-16)


@pytest.mark.parametrize("n,dim,K", [(5000, 64, 16), (20000, 384, 200), (7, 16, 3), (30011, 128, 1024), (4099, 32, 2048)])
def test_assign_and_update_match_oracle(n, dim, K):
    # >= 4096 rows against <= 2048 centroids takes the bucketed top-1 kernel (rows staged per group of 8
    # centroids); fewer rows, the warp-per-row kernel.  Rows arrive shuffled so every bucket sees many groups.
    from lazzaro_b200 import consolidate, synth
    rows_np, _ = synth.clustered_rows(n, dim, K, seed=n)
    rows_np = rows_np[np.random.default_rng(n).permutation(n)]
    cen_np, _ = synth.clustered_rows(K, dim, K, seed=n + 1, sigma=2.0)
    cen_np *= np.linspace(0.5, 2.0, K, dtype=np.float32)[:, None]      # un-normalised centroids
    if K > 2:
        cen_np[2] = cen_np[1]                                           # tie inside a group: lowest index wins
    if K > 16:
        cen_np[13] = cen_np[4]                                          # tie across two groups of 8
    rows, cen = torch.as_tensor(rows_np).cuda(), torch.as_tensor(cen_np).cuda()
    idx, cos = consolidate.assign_to_centroids(rows, cen)
    want_idx, want_cos = graph_ops.assign_to_centroids(rows_np, cen_np)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_idx)
    np.testing.assert_allclose(cos.cpu().numpy(), want_cos, rtol=1e-12)
    assert not (idx == 2).any() or K <= 2
    assert not (idx == 13).any() or K <= 16
    means = consolidate.centroid_means(rows, assign=idx, n_centroids=K).cpu().numpy()
    np.testing.assert_allclose(means, graph_ops.centroid_means(rows_np, want_idx, K), rtol=1e-12, atol=1e-15)


def test_cfg3_size_dedup_and_reassignment():
    """BASELINE configs[2]: 10k-node episodic buffer against a 1M-node graph + shard reassignment with
    k = 1024 centroids — checked by planted duplicates, by the exact scan on a query subset, and by
    properties of the assignment (each row's own-shard centroid wins; update = mean of members)."""
    from lazzaro_b200 import consolidate, kernels as K, synth, _lib
    from lazzaro_b200.store import B200Store
    n, dim, n_new, kc = 1_000_000, 768, 10_000, 1024
    rows, shard = synth.clustered_rows_device(n, dim, kc, "cuda")
    st = B200Store()
    st.adopt_rows(rows, id_prefix="g")
    gen = torch.Generator(device="cuda"); gen.manual_seed(99)
    src = torch.randint(0, n, (n_new,), generator=gen, device="cuda")
    noise = torch.randn(n_new, dim, generator=gen, device="cuda") / dim ** 0.5
    new = rows[src] + 0.25 * noise                                     # cos ~ 0.97 to the source row
    new[1000:] = torch.randn(n_new - 1000, dim, generator=gen, device="cuda")   # 9000 fresh memories
    new = new / new.norm(dim=1, keepdim=True)
    names, sims, dup = consolidate.dedup_against_store(st, new.cpu().numpy())
    assert dup[:1000].all() and not dup[1000:].any()
    assert [int(x[1:]) for x in names[:1000]] == src[:1000].tolist()
    sub = torch.arange(0, n_new, 157, device="cuda")
    eids, esc, _ = K.search_exact(rows, dim, st._norms, None, new[sub].contiguous(), 1, _lib.COSINE)
    assert [int(names[i][1:]) for i in sub.tolist()] == eids[:, 0].tolist()
    np.testing.assert_array_equal(sims[sub.cpu().numpy()], esc[:, 0].cpu().numpy())
    # shard reassignment: centroids = shard means (segments are contiguous), then assign every row
    seg = torch.searchsorted(shard.long().contiguous(), torch.arange(kc + 1, device="cuda"))
    cen64 = consolidate.centroid_means(rows, seg_offsets=seg)
    ref = rows[: seg[1]].double().mean(0)
    assert torch.allclose(cen64[0], ref, rtol=1e-12, atol=1e-15)
    idx, cos = consolidate.assign_to_centroids(rows, cen64.float().contiguous())
    assert (idx == shard.long()).float().mean().item() > 0.999          # clusters are well separated
    assert (cos > 0.9).all()
    means = consolidate.centroid_means(rows, assign=idx, n_centroids=kc)
    assert torch.allclose(means[5], rows[idx == 5].double().mean(0), rtol=1e-11, atol=1e-14)


@pytest.mark.parametrize("margin", ["worst_case", "statistical"])
@pytest.mark.parametrize("order", ["clustered", "shuffled"])
def test_assign_with_nearly_coincident_centroids(margin, order):
    """k-means with k above the number of clusters: ~10 centroids per true cluster, nearly coincident (means of equal
    slabs of the rows), so a row sees many centroids inside the rescoring window — the multi-candidate kernel.  Rows
    stored cluster by cluster (the store's layout: a block shares its candidates) and shuffled (every block needs far
    more table rows than its cache holds: staged in rounds).  Exact ties (duplicated centroids) go to the lowest index."""
    from lazzaro_b200 import consolidate, synth
    n, dim, clusters, K = 24000, 96, 12, 120
    rows_np, _ = synth.clustered_rows(n, dim, clusters, seed=5)
    seg = np.linspace(0, n, K + 1).astype(np.int64)
    cen_np = np.stack([rows_np[a:b].astype(np.float64).mean(0) for a, b in zip(seg[:-1], seg[1:])]).astype(np.float32)
    cen_np[7] = cen_np[3]                                                # exact duplicates: index 3 must win
    cen_np[55] = cen_np[51]
    if order == "shuffled":
        rows_np = rows_np[np.random.default_rng(1).permutation(n)]
    rows, cen = torch.as_tensor(rows_np).cuda(), torch.as_tensor(cen_np).cuda()
    idx, cos = consolidate.assign_to_centroids(rows, cen, margin=margin)
    want_idx, want_cos = graph_ops.assign_to_centroids(rows_np, cen_np)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_idx)
    np.testing.assert_allclose(cos.cpu().numpy(), want_cos, rtol=1e-12)
    assert not (idx == 7).any() and not (idx == 55).any()
    # the window really holds several centroids per row on this data
    sims = graph_ops.cosine_matrix(rows_np[:200], cen_np)
    assert np.median((sims >= sims.max(1, keepdims=True) - 1e-3).sum(1)) >= 2


@pytest.mark.parametrize("n,K", [(1, 1), (5000, 7), (100_003, 1024), (70_000, 16384), (4096, 3)])
def test_group_by_label_is_a_stable_counting_sort(n, K):
    from lazzaro_b200 import kernels as Kn
    g = torch.Generator(device="cuda"); g.manual_seed(n)
    labels = torch.randint(0, K, (n,), generator=g, device="cuda", dtype=torch.int64)
    if n > 10:
        labels[3] = -1                     # out-of-range labels are left out
        labels[7] = K
    order, seg = Kn.group_by_label(labels, K)
    ok = (labels >= 0) & (labels < K)
    idx = torch.nonzero(ok).flatten()
    want = idx[torch.argsort(labels[idx], stable=True)]
    assert torch.equal(order[: want.numel()], want)
    counts = torch.bincount(labels[idx], minlength=K)
    assert torch.equal(seg[1:] - seg[:-1], counts) and int(seg[0]) == 0


def test_assign_rows_resident_in_a_store_matches_the_plain_call():
    """The store's own fp16 shadow and rounding-error norms as the query side (no conversion pass): same assignment."""
    from lazzaro_b200 import consolidate, synth
    from lazzaro_b200.store import B200Store
    rows_np, _ = synth.clustered_rows(20000, 128, 10, seed=2)
    rows_np *= np.linspace(0.5, 2.0, 20000, dtype=np.float32)[:, None]
    cen_np = np.stack([rows_np[a:a + 200].astype(np.float64).mean(0) for a in range(0, 20000, 200)]).astype(np.float32)
    st = B200Store(metric="l2")                                       # an l2 store builds its cosine shadow on demand
    st.add_rows(rows_np)
    cen = torch.as_tensor(cen_np).cuda()
    idx, cos = consolidate.assign_store_rows(st, cen)
    want_idx, want_cos = graph_ops.assign_to_centroids(rows_np, cen_np)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_idx)
    np.testing.assert_allclose(cos.cpu().numpy(), want_cos, rtol=1e-12)
    idx2, cos2 = consolidate.assign_store_rows(st, cen, row_range=(5000, 12000))
    assert torch.equal(idx2, idx[5000:12000]) and torch.equal(cos2, cos[5000:12000])
