"""Measurement only (LZ_PRUNE_FORM=4): the streamed kernel's data path with precomputed offsets; scratch is oversized."""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import kernels as K, _lib
lib = _lib.load(); dev = "cuda"
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
g = torch.Generator(device=dev); g.manual_seed(1)
cols = [torch.randint(0, 1 << 20, (n,), generator=g, device=dev, dtype=torch.int32) for _ in range(2)] + \
       [torch.rand(n, generator=g, device=dev), torch.randint(1, 9, (n,), generator=g, device=dev, dtype=torch.int32)]
outs = [torch.empty_like(c) for c in cols]
kept = torch.zeros(1, dtype=torch.int64, device=dev)
scratch = torch.zeros(4 * int(lib.lz_edge_decay_prune_scratch_bytes(n)), dtype=torch.uint8, device=dev)
f = lambda: K.edge_decay_prune(*cols, n, 0.99, 0.5, out=tuple(outs), scratch=scratch, kept=kept)
for _ in range(50): f()
torch.cuda.synchronize()
w64 = cols[2].double() * 0.99; keep = ~(w64 < 0.5)
m = int(kept.item())
print("kept", m, "ok" if m == int(keep.sum()) and torch.equal(outs[0][:m], cols[0][keep]) and torch.equal(outs[2][:m], w64.float()[keep]) else "MISMATCH")
best = 1e9
for _ in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): f()
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / 20)
kf = m / n
print(f"form4 st={os.environ.get('LZ_PRUNE_STAGES','-')} ctas={os.environ.get('LZ_PRUNE_CTAS','-')} n={n}: {best*1e3:.1f} us {(16+16*kf)*n/1e9/best*1e3:.0f} GB/s")
