"""Decay + prune on small tables: per-call time when the calls are issued from Python one by one vs replayed from a CUDA
graph of 20 calls (the host side of a call costs more than the sweep of a few million edges).
python scripts/run_edges_graph.py [n_edges ...]"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import kernels as K, _lib
lib = _lib.load(); dev = "cuda"
for n in ([int(a) for a in sys.argv[1:]] or [1_000_000, 6_250_000, 25_000_000]):
    g = torch.Generator(device=dev); g.manual_seed(1)
    cols = [torch.randint(0, 1 << 20, (n,), generator=g, device=dev, dtype=torch.int32) for _ in range(2)] + \
           [torch.rand(n, generator=g, device=dev), torch.ones(n, device=dev, dtype=torch.int32)]
    outs = [torch.empty_like(c) for c in cols]
    kept = torch.zeros(1, dtype=torch.int64, device=dev)
    scratch = torch.empty(int(lib.lz_edge_decay_prune_scratch_bytes(n)), dtype=torch.uint8, device=dev)
    f = lambda: K.edge_decay_prune(*cols, n, 0.99, 0.5, out=tuple(outs), scratch=scratch, kept=kept)
    for _ in range(300): f()
    torch.cuda.synchronize()
    def timed(fn, it):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / it
    direct = min(timed(lambda: [f() for _ in range(20)], 20) for _ in range(5))
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        f(); torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            for _ in range(20): f()
    torch.cuda.synchronize()
    for _ in range(5): gr.replay()
    torch.cuda.synchronize()
    graph = min(timed(gr.replay, 20) for _ in range(5))
    kf = float(kept.item()) / n
    gb = (16 + 16 * kf) * n / 1e9
    print(f"n={n}: direct {direct*1e3:.1f} us ({gb/direct*1e3:.0f} GB/s)   graph of 20 {graph*1e3:.1f} us ({gb/graph*1e3:.0f} GB/s)", flush=True)
