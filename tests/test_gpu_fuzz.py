"""GPU: seeded random sweep of shapes, k, metrics, filters and ranges — exact scan and tensor-core path
against the numpy oracle (ids bit-exact, scores to 1e-12)."""
import numpy as np
import pytest
import torch

from oracle import flat_search

pytestmark = pytest.mark.gpu


def _case(seed):
    rng = np.random.default_rng(seed)
    dim = int(rng.choice([8, 16, 24, 64, 72, 128, 200, 384, 520]))
    n = int(rng.choice([1, 7, 255, 256, 257, 1000, 4097, 20011]))
    B = int(rng.choice([1, 2, 5, 8, 9, 127, 129, 300]))
    k = int(rng.choice([1, 2, 3, 4, 10, 17, 64, 100]))
    return rng, dim, n, B, k


@pytest.mark.parametrize("seed", range(24))
def test_random_configurations(seed):
    from lazzaro_b200 import _lib, kernels as K, synth
    rng, dim, n, B, k = _case(seed)
    rows_np, _ = synth.clustered_rows(n, dim, max(1, min(8, n)), seed=seed, sigma=float(rng.choice([0.2, 0.6, 2.0])))
    rows_np *= rng.uniform(0.1, 5.0, size=(n, 1)).astype(np.float32)            # un-normalised rows
    if n > 3:
        rows_np[n // 2] = 0.0
        rows_np[n // 3] = rows_np[0]
    qs_np = synth.queries_near_rows(rows_np + 1e-3, B, seed=seed + 100, noise=float(rng.choice([0.05, 0.5, 3.0])))
    tags_np = rng.integers(-1, 3, size=n).astype(np.int32) if rng.random() < 0.5 else None
    want_tag = int(rng.integers(-1, 3)) if tags_np is not None else -1
    lo = int(rng.integers(0, max(1, n // 2))) if rng.random() < 0.5 else 0
    hi = int(rng.integers(lo, n + 1)) if rng.random() < 0.5 else n
    mask = np.zeros(n, bool); mask[lo:hi] = True
    if tags_np is not None:
        mask &= (tags_np >= 0) if want_tag < 0 else (tags_np == want_tag)
    rows, qs = torch.as_tensor(rows_np).cuda(), torch.as_tensor(qs_np).cuda()
    tags = None if tags_np is None else torch.as_tensor(tags_np).cuda()
    norms = K.row_norms(rows)
    kk = min(k, max(1, int(mask.sum())))
    for metric in ("cosine", "l2"):
        want_ids, want_sc = flat_search.flat_topk_batch(rows_np, qs_np, kk, metric, mask)
        gs = 1 if tags is not None else None
        gtag = None if tags is None else torch.full((B,), want_tag, dtype=torch.int32, device="cuda")
        ids, sc, cnt = K.search_exact(rows, dim, norms, tags, qs, kk, _lib.METRICS[metric], row_range=(lo, hi),
                                      group_tag=gtag, group_size=gs)
        np.testing.assert_array_equal(ids.cpu().numpy(), want_ids)
        ok = want_ids >= 0
        np.testing.assert_allclose(sc.cpu().numpy()[ok], want_sc[ok], rtol=1e-12, atol=1e-14)
        assert cnt.cpu().numpy().tolist() == ok.sum(1).tolist()
    if kk <= 128:
        for metric in ("cosine", "l2"):
            shadow = K.rows_to_bf16(rows, dim, norms, metric=_lib.METRICS[metric])
            want_ids, want_sc = flat_search.flat_topk_batch(rows_np, qs_np, kk, metric, mask)
            ids, sc, cnt, cert = K.search_batched(shadow, rows, dim, norms, tags, want_tag, qs, kk, row_range=(lo, hi),
                                                  metric=_lib.METRICS[metric])
            ids_np, ok = ids.cpu().numpy(), want_ids >= 0
            np.testing.assert_array_equal(ids_np, want_ids)
            np.testing.assert_allclose(sc.cpu().numpy()[ok], want_sc[ok], rtol=1e-12, atol=1e-14)
