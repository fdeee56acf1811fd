"""In-place decay + prune timing: each timed call works on a fresh copy of the table (the copy is not timed).
python scripts/run_edges_inplace.py [n_edges]     (LZ_PRUNE_INPLACE_OLD=1: counts published by the compute warps)"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import kernels as K, _lib
lib = _lib.load(); dev = "cuda"
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
g = torch.Generator(device=dev); g.manual_seed(1)
cols = [torch.randint(0, 1 << 20, (n,), generator=g, device=dev, dtype=torch.int32) for _ in range(2)] + \
       [torch.rand(n, generator=g, device=dev), torch.arange(n, device=dev, dtype=torch.int32)]
scratch = torch.empty(int(lib.lz_edge_decay_prune_scratch_bytes(n)), dtype=torch.uint8, device=dev)
kept = torch.zeros(1, dtype=torch.int64, device=dev)
for thr in (0.5, 0.0, 0.9):
    w64 = cols[2].double() * 0.99; keep = ~(w64 < thr)
    want = [cols[0][keep], w64.float()[keep], cols[3][keep]]
    ts = []
    for it in range(8):
        work = [c.clone() for c in cols]
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        K.edge_decay_prune(*work, n, 0.99, thr, scratch=scratch, kept=kept)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    m = int(kept.item())
    ok = m == int(keep.sum()) and torch.equal(work[0][:m], want[0]) and torch.equal(work[2][:m], want[1]) and torch.equal(work[3][:m], want[2])
    ms = sorted(ts)[len(ts) // 2]
    print(f"in place n={n} thr={thr} kept={m/n:.3f}: {ms*1e3:.1f} us  {(16+16*m/n)*n/1e9/ms*1e3:.0f} GB/s  {'ok' if ok else 'MISMATCH'}", flush=True)
