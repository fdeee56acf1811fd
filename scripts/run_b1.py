"""Batch-1 exact search driver for profiling: python scripts/run_b1.py [rows] [k]"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import kernels as K, synth, _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 10
rows, _ = synth.clustered_rows_device(n, 768, 1024, "cuda")
norms = K.row_norms(rows)
q = synth.queries_near_rows_device(rows, 1)
for _ in range(3):
    out = K.search_exact(rows, 768, norms, None, q, k, _lib.COSINE)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    out = K.search_exact(rows, 768, norms, None, q, k, _lib.COSINE)
e1.record(); torch.cuda.synchronize()
print("ms", e0.elapsed_time(e1) / 10)
