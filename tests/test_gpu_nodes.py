"""GPU parity: node-side sweeps of SURVEY §8f — importance scoring + lowest-k eviction
(memory_system.py:535-578) against the reference's recorded decisions and the oracle."""
import numpy as np
import pytest
import torch

from oracle import graph_ops

pytestmark = pytest.mark.gpu


def _evict(sal, acc, last, now, k, skip=None, want_all=False):
    from lazzaro_b200 import kernels as K
    t = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a).astype(dt)).cuda()
    return K.evict_lowest(t(sal, sal.dtype), t(acc, np.int32), t(last, np.float64), now, k,
                          skip=None if skip is None else t(skip, np.uint8), want_all=want_all)


def test_eviction_matches_reference(golden_next):
    g = golden_next
    for c in range(3):
        sal, acc, last, sup = g[f"ev{c}_sal"], g[f"ev{c}_acc"], g[f"ev{c}_last"], g[f"ev{c}_super"]
        excess = len(sal) - int(g[f"ev{c}_limit"])
        idx, imp = _evict(sal, acc, last, float(g["ev_now"]), excess, skip=sup)
        assert g[f"ev{c}_order"][idx.cpu().numpy()].tolist() == g[f"ev{c}_deleted"].tolist()
        _, want_imp = graph_ops.evict_lowest(sal, acc, last, float(g["ev_now"]), excess, skip=sup)
        np.testing.assert_array_equal(imp.cpu().numpy(), want_imp)              # bit-exact float64


@pytest.mark.parametrize("n,k,f64", [(1, 1, True), (7, 10, False), (5000, 64, True), (300_000, 1024, False), (2_000_003, 257, False)])
def test_eviction_matches_oracle(n, k, f64):
    rng = np.random.default_rng(n)
    now = 1.76e9
    sal = rng.uniform(0, 1, n).round(2 if n > 1000 else 6)                      # many exact ties at scale
    sal = sal if f64 else sal.astype(np.float32)
    acc = rng.integers(0, 30, n).astype(np.int32)
    last = now - rng.choice([0.0, 60.0, 86400.0, 10 * 86400.0, 900 * 86400.0], n)
    skip = rng.random(n) < 0.1
    sal[n // 2] = np.nan                                                        # never evicted
    idx, imp, all_imp = _evict(sal, acc, last, now, k, skip=skip, want_all=True)
    want_idx, want_imp = graph_ops.evict_lowest(sal, acc, last, now, k, skip=skip)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_idx)
    np.testing.assert_array_equal(imp.cpu().numpy(), want_imp)
    ref_all = graph_ops.importance(sal, acc, (now - last) / 86400)
    np.testing.assert_array_equal(all_imp.cpu().numpy(), ref_all)
    idx2, _ = _evict(sal, acc, last, now, k)                                    # no skip mask
    np.testing.assert_array_equal(idx2.cpu().numpy(), graph_ops.evict_lowest(sal, acc, last, now, k)[0])


def test_eviction_edge_cases():
    from lazzaro_b200 import _lib
    e = np.zeros(0)
    idx, imp = _evict(e.astype(np.float32), e, e, 0.0, 5)
    assert idx.numel() == 0 and imp.numel() == 0
    one = np.ones(4)
    idx, _ = _evict(one.astype(np.float32), one, one, 10.0, 0)
    assert idx.numel() == 0
    idx, _ = _evict(one.astype(np.float32), one, one, 10.0, 3, skip=np.ones(4, bool))
    assert idx.numel() == 0
    with pytest.raises(_lib.LazzaroB200Error):
        _evict(np.ones(5000, np.float32), np.ones(5000), np.ones(5000), 10.0, 2000)


def test_cfg5_size_eviction_equals_device_sort():
    """12.5M nodes (one GPU's share of BASELINE configs[4]): the k least important nodes equal a stable float64 sort of
    the same expression evaluated operation by operation with torch on the device (each torch op rounds once, as the
    reference's Python expression does)."""
    from lazzaro_b200 import kernels as K
    n, now = 12_500_000, 1.76e9
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    sal = (torch.randint(0, 101, (n,), device="cuda", generator=g).float() / 100).contiguous()      # many exact ties
    acc = torch.randint(0, 15, (n,), device="cuda", generator=g, dtype=torch.int32)
    last = now - torch.randint(0, 400, (n,), device="cuda", generator=g).double() * 86400.0
    skip = (torch.rand(n, device="cuda", generator=g) < 0.02).to(torch.uint8)
    c = lambda v: torch.tensor(v, dtype=torch.float64, device="cuda")     # tensor divisors: torch turns `x / python_scalar`
    days = (now - last) / c(86400.0)                                       # into a multiplication by the reciprocal
    imp = sal.double() * 0.5 + torch.clamp(acc.double() / c(10.0), max=1.0) * 0.3 + torch.reciprocal(1.0 + days) * 0.2
    masked = torch.where(skip.bool(), torch.full_like(imp, float("inf")), imp)
    for k in (1, 64, 1024):
        idx, val, all_imp = K.evict_lowest(sal, acc, last, now, k, skip=skip, want_all=True)
        assert bool((all_imp == imp).all())
        order = torch.sort(masked, stable=True).indices[:k]
        assert idx.tolist() == order.tolist()
        assert bool((val == imp[order]).all())
