"""GPU parity: the tensor-core batched search (lz_search_batched) against the exact scan and the
numpy oracle. ids bit-exact; scores bit-identical to the exact scan (same float64 summation order)."""
import numpy as np
import pytest
import torch

from oracle import flat_search

pytestmark = pytest.mark.gpu


def setup(rows_np):
    from lazzaro_b200 import kernels as K
    rows = torch.as_tensor(np.ascontiguousarray(rows_np, np.float32)).cuda()
    norms = K.row_norms(rows)
    shadow = K.rows_to_shadow(rows, rows.shape[1], norms)
    return K, rows, norms, shadow


def test_shadow_is_normalised_bf16():
    from lazzaro_b200 import synth
    rows_np, _ = synth.clustered_rows(1000, 64, 4, seed=1)
    rows_np *= np.linspace(0.5, 3.0, 1000, dtype=np.float32)[:, None]
    rows_np[7] = 0
    K, rows, norms, shadow = setup(rows_np)
    want = torch.as_tensor(rows_np / np.maximum(np.linalg.norm(rows_np, axis=1, keepdims=True), 1e-30)).to(torch.float16)
    got = shadow.cpu()
    assert (got[7] == 0).all()
    assert torch.allclose(got.float(), want.float(), atol=2 ** -8, rtol=2 ** -7)


@pytest.mark.parametrize("n,dim,B,k", [(5000, 64, 3, 5), (20000, 384, 130, 10), (33333, 768, 256, 10),
                                       (4096, 128, 128, 100), (70000, 200, 64, 20), (300, 72, 9, 10)])
def test_batched_matches_exact_scan_and_oracle(n, dim, B, k):
    from lazzaro_b200 import synth, _lib
    rows_np, _ = synth.clustered_rows(n, dim, 8, seed=n)
    qs_np = synth.queries_near_rows(rows_np, B, seed=B)
    K, rows, norms, shadow = setup(rows_np)
    qs = torch.as_tensor(qs_np).cuda()
    ids, sc, cnt, cert = K.search_batched(shadow, rows, dim, norms, None, -1, qs, k)
    eids, esc, ecnt = K.search_exact(rows, dim, norms, None, qs, k, _lib.COSINE)
    torch.cuda.synchronize()
    assert torch.equal(ids, eids) and torch.equal(cnt, ecnt)
    assert torch.equal(sc[ids >= 0], esc[eids >= 0])                    # bit-identical float64 scores
    sub = slice(0, min(B, 16))
    want_ids, want_sc = flat_search.flat_topk_batch(rows_np, qs_np[sub], k)
    np.testing.assert_array_equal(ids[sub].cpu().numpy(), want_ids)
    kk = min(k, n)
    np.testing.assert_allclose(sc[sub, :kk].cpu().numpy(), want_sc[:, :kk], rtol=1e-12)
    assert cert.float().mean().item() > 0.5, "the tensor-core pass should certify most queries on this data"


def test_batched_with_filter_range_and_uncertifiable_queries():
    from lazzaro_b200 import synth, _lib
    rows_np, _ = synth.clustered_rows(30000, 128, 8, seed=77)
    rows_np[100:400] = rows_np[50]                      # 300 exact duplicates: coarse ties, cannot be certified
    qs_np = synth.queries_near_rows(rows_np, 40, seed=78)
    qs_np[0] = rows_np[50]
    qs_np[1] = 0                                        # zero query
    tags_np = (np.arange(30000) % 3).astype(np.int32)
    tags_np[::11] = -1
    K, rows, norms, shadow = setup(rows_np)
    qs, tags = torch.as_tensor(qs_np).cuda(), torch.as_tensor(tags_np).cuda()
    for want, rng in [(-1, (0, 30000)), (1, (0, 30000)), (2, (777, 25001)), (0, (5000, 5003))]:
        ids, sc, cnt, cert = K.search_batched(shadow, rows, 128, norms, tags, want, qs, 10, row_range=rng)
        torch.cuda.synchronize()
        mask = (tags_np >= 0) if want < 0 else (tags_np == want)
        mask = mask & (np.arange(30000) >= rng[0]) & (np.arange(30000) < rng[1])
        want_ids, want_sc = flat_search.flat_topk_batch(rows_np, qs_np, 10, mask=mask)
        np.testing.assert_array_equal(ids.cpu().numpy(), want_ids)
        np.testing.assert_array_equal(cnt.cpu().numpy(), (want_ids >= 0).sum(1))
    assert cert[0].item() == 0 or rng != (0, 30000)      # the duplicate-heavy query went through the fallback


def test_batched_cfg2_size_against_exact_scan():
    """BASELINE configs[1]: 1M x 768, 64 shards, batch 256, top-10 — tensor-core path vs exact scan."""
    from lazzaro_b200 import synth, kernels as K, _lib
    n, dim, B, k = 1_000_000, 768, 256, 10
    rows, _ = synth.clustered_rows_device(n, dim, 64, "cuda")
    qs = synth.queries_near_rows_device(rows, B)
    norms = K.row_norms(rows)
    shadow = K.rows_to_shadow(rows, dim, norms)
    ids, sc, cnt, cert = K.search_batched(shadow, rows, dim, norms, None, -1, qs, k)
    eids, esc, ecnt = K.search_exact(rows, dim, norms, None, qs, k, _lib.COSINE)
    torch.cuda.synchronize()
    assert torch.equal(ids, eids) and torch.equal(sc, esc) and torch.equal(cnt, ecnt)
    assert cert.float().mean().item() > 0.9


@pytest.mark.parametrize("n,dim,B,k", [(20000, 384, 130, 10), (33333, 768, 64, 10), (4096, 128, 128, 100), (3000, 72, 9, 3)])
def test_batched_l2_matches_exact_scan_and_oracle(n, dim, B, k):
    """LanceDB's default metric on the tensor cores: the L2 ranking comes out of the same contraction
    (x.q - |x|^2/2 with the norm term carried as three extra bf16 columns)."""
    from lazzaro_b200 import synth, _lib, kernels as K
    rows_np, _ = synth.clustered_rows(n, dim, 8, seed=n + 1)
    rows_np *= np.random.default_rng(n).uniform(0.5, 2.0, size=(n, 1)).astype(np.float32)   # un-normalised rows
    qs_np = synth.queries_near_rows(rows_np, B, seed=B + 1) * np.float32(1.3)
    rows, qs = torch.as_tensor(rows_np).cuda(), torch.as_tensor(qs_np).cuda()
    norms = K.row_norms(rows)
    shadow = K.rows_to_shadow(rows, dim, norms, metric=_lib.L2)
    assert shadow.shape[1] == dim + 8
    ids, sc, cnt, cert = K.search_batched(shadow, rows, dim, norms, None, -1, qs, k, metric=_lib.L2)
    eids, esc, ecnt = K.search_exact(rows, dim, norms, None, qs, k, _lib.L2)
    torch.cuda.synchronize()
    assert torch.equal(ids, eids) and torch.equal(cnt, ecnt) and torch.equal(sc, esc)
    want_ids, want_sc = flat_search.flat_topk_batch(rows_np, qs_np[:8], k, "l2")
    np.testing.assert_array_equal(ids[:8].cpu().numpy(), want_ids)
    np.testing.assert_allclose(sc[:8].cpu().numpy(), want_sc, rtol=1e-12)
    assert cert.float().mean().item() > 0.5


def test_store_l2_batched_equals_single_queries():
    from lazzaro_b200 import synth
    from lazzaro_b200.store import B200Store
    rows_np, _ = synth.clustered_rows(12000, 96, 6, seed=3)
    rows_np *= np.linspace(0.5, 1.5, 12000, dtype=np.float32)[:, None]
    qs_np = synth.queries_near_rows(rows_np, 40, seed=4)
    st = B200Store(metric="l2")
    st.add_rows(rows_np, id_prefix="n")
    got = st.search_batch(qs_np, limit=7)
    assert st.last_certified is not None                                    # went through the tensor-core path
    st.batched_search = "never"
    assert got == [st.search_nodes(q.tolist(), limit=7) for q in qs_np]


def test_cfg4_size_10m_rows_properties():
    """BASELINE configs[3] on one GPU: 10M x 768, batch 1024, top-100 and top-10 — size-independent properties:
    planted rows are their own best match, results are sorted, every query certifies, the tensor-core
    answer equals the exact float64 scan on a query subset bit for bit, and a sharded search (two halves
    merged on the device) equals the whole."""
    from lazzaro_b200 import synth, kernels as K, _lib
    n, dim, B = 10_000_000, 768, 1024
    rows, _ = synth.clustered_rows_device(n, dim, 1024, "cuda")
    qs = synth.queries_near_rows_device(rows, B)
    pick = torch.tensor([0, 123_456, 5_000_000, n - 1], device="cuda")
    qs[:4] = rows[pick]
    norms = K.row_norms(rows)
    shadow = K.rows_to_shadow(rows, dim, norms)
    for k in (100, 10):
        ids, sc, cnt, cert = K.search_batched(shadow, rows, dim, norms, None, -1, qs, k)
        torch.cuda.synchronize()
        assert ids[:4, 0].tolist() == pick.tolist()
        assert (cnt == k).all() and (sc[:, 1:] <= sc[:, :-1]).all()
        assert cert.float().mean().item() == 1.0
        sub = qs[:8].contiguous()
        eids, esc, _ = K.search_exact(rows, dim, norms, None, sub, k, _lib.COSINE)
        assert torch.equal(ids[:8], eids) and torch.equal(sc[:8], esc)
    half = n // 2
    parts = [K.search_batched(shadow, rows, dim, norms, None, -1, qs, 10, row_range=r)[:2] for r in ((0, half), (half, n))]
    mids, msc, _ = K.merge_topk(torch.stack([p[0] for p in parts]), torch.stack([p[1] for p in parts]), _lib.COSINE)
    assert torch.equal(mids, ids) and torch.equal(msc, sc)


def test_searches_in_flight_on_two_lanes_give_the_same_answers():
    """ShardedSearch lanes: consecutive searches on alternating lanes (own stream, own scratch) overlap on the GPU and
    must return what the one-stream search returns — different query batches per step, so a scratch mix-up shows."""
    from lazzaro_b200 import synth
    from lazzaro_b200.dist import ShardedSearch
    from lazzaro_b200.store import B200Store
    rows_np, _ = synth.clustered_rows(60000, 128, 8, seed=5)
    st = B200Store()
    st.add_rows(rows_np)
    ss = ShardedSearch(st, world=1)
    batches = [torch.as_tensor(synth.queries_near_rows(rows_np, 96, seed=100 + i)).cuda() for i in range(6)]
    want = [tuple(t.clone() for t in ss.search_device(q, 10)) for q in batches]
    torch.cuda.synchronize()
    got = []
    for i, q in enumerate(batches):
        out = ss.search_device(q, 10, lane=i & 1)
        if i >= 1:                                       # a lane's buffers are reused two steps later: copy the previous result
            pass
        got.append(out)
        if i % 2 == 1:
            ss.join()
            torch.cuda.current_stream().synchronize()
            for j in (i - 1, i):
                assert torch.equal(got[j][0], want[j][0]) and torch.equal(got[j][1], want[j][1]) and torch.equal(got[j][2], want[j][2])
