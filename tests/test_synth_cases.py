"""CPU: the adversarial certification fixture has the properties its docstring states (fp16 emulated with torch)."""
import numpy as np
import torch


def test_correlated_rounding_case_properties():
    from lazzaro_b200 import synth
    rows, q, b, info = synth.correlated_rounding_case(n_high=16, n_low_groups=40)
    x = rows.astype(np.float64)
    n = np.linalg.norm(x, axis=1)
    assert np.abs(n - 1).max() < 1e-7 and abs(float(np.linalg.norm(q.astype(np.float64))) - 1) < 1e-12
    cos = x @ q.astype(np.float64) / n
    assert set(np.unique(cos[:b]).tolist()) == {info["L"], info["H"]}
    assert cos[b] > info["H"] + 4e-5
    xb = torch.tensor((x / n[:, None]).astype(np.float32)).to(torch.float16).double().numpy()
    qb = torch.tensor(q).to(torch.float16).double().numpy()
    assert (xb[:b] == x[:b]).all() and (qb == q).all()            # exact shadows: zero coarse error
    coarse_b = float(xb[b] @ qb)
    assert coarse_b < info["L"] - 3e-5                            # B drops below every other row in the coarse pass
    stat_eps = 4 * 2.0 ** -11 * np.sqrt(2 / 3) * float(np.abs(q).max()) + 8e-6
    assert info["H"] > info["L"] + stat_eps                       # ... and the statistical margin would still certify
    assert cos[b] - coarse_b > 3 * stat_eps
