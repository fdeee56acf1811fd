import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import _lib, kernels as K, synth
ne = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
src, dst, w, cooc = synth.edge_table(ne, 12_500_000)
cols = [torch.as_tensor(a).cuda() for a in (src, dst, w, cooc)]
outs = [torch.empty_like(c) for c in cols]
kept = torch.zeros(1, dtype=torch.int64, device="cuda")
scratch = torch.empty(int(_lib.load().lz_edge_decay_prune_scratch_bytes(ne)), dtype=torch.uint8, device="cuda")
f = lambda: K.edge_decay_prune(*cols, ne, 0.99, 0.5, out=tuple(outs), scratch=scratch, kept=kept)
for _ in range(3): f()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): f()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
kf = kept.item() / ne
print(f"{ne} edges: {ms:.4f} ms, {ne/ms/1e6:.1f} G edges/s, {(16+16*kf)*ne/ms/1e6:.0f} GB/s, kept {kf:.3f}")
