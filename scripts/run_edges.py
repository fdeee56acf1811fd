"""Decay + prune sweep timing: python scripts/run_edges.py [n_edges]   (LZ_PRUNE_TWO_PASS=1: the count/scan/scatter form)"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import kernels as K, _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
g = torch.Generator(device="cuda"); g.manual_seed(1)
cols = [torch.randint(0, 1 << 20, (n,), generator=g, device="cuda", dtype=torch.int32) for _ in range(2)] + \
       [torch.rand(n, generator=g, device="cuda"), torch.ones(n, device="cuda", dtype=torch.int32)]
outs = [torch.empty_like(c) for c in cols]
kept = torch.zeros(1, dtype=torch.int64, device="cuda")
scratch = torch.empty(int(_lib.load().lz_edge_decay_prune_scratch_bytes(n)), dtype=torch.uint8, device="cuda")
def timed(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
ms = timed(lambda: K.edge_decay_prune(*cols, n, 0.99, 0.5, out=tuple(outs), scratch=scratch, kept=kept))
kf = float(kept.item()) / n
gb = (16 + 16 * kf) * n / 1e9
print(f"out of place: {ms:.4f} ms  {n / ms / 1e6:.1f} G edges/s  {gb / ms * 1e3:.0f} GB/s  kept {kf:.4f}", flush=True)
# in place on a fresh copy each time is not possible without the copy; time the in-place kernel on already-pruned data (kept = 1)
work = [c.clone() for c in cols]
K.edge_decay_prune(*work, n, 0.99, 0.5, scratch=scratch, kept=kept)
m = int(kept.item())
ms = timed(lambda: K.edge_decay_prune(*work, m, 1.0, 0.0, scratch=scratch, kept=kept))
print(f"in place (all {m} kept): {ms:.4f} ms  {32 * m / 1e9 / ms * 1e3:.0f} GB/s", flush=True)
