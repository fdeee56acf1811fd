"""Certified fraction and time of the tensor-core search under both margins (10M x 768 by default)."""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import kernels as K, synth, _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
dim = 768
rows, _ = synth.clustered_rows_device(n, dim, max(1, n // 9765), "cuda")
norms = K.row_norms(rows)
err = torch.empty(n, dtype=torch.float32, device="cuda")
shadow = K.rows_to_shadow(rows, dim, norms, row_err=err)
mx = float(err.max())
print("row_err max %.3e mean %.3e" % (mx, float(err.mean())))
for B, k, noise in ((1024, 10, 0.1), (1024, 100, 0.1), (256, 10, 0.1), (10000, 1, 0.25), (10000, 3, 0.25), (10000, 1, 1.0)):
    qs = synth.queries_near_rows_device(rows, B, noise=noise)
    for det in (None, mx):
        out = K.search_batched(shadow, rows, dim, norms, None, -1, qs, k, det_row_err=det)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            out = K.search_batched(shadow, rows, dim, norms, None, -1, qs, k, det_row_err=det)
        e1.record(); torch.cuda.synchronize()
        print(f"B={B} k={k} noise={noise} margin={'det' if det else 'stat'}: {e0.elapsed_time(e1)/3:.3f} ms  certified {float(out[3].float().mean()):.4f}", flush=True)
