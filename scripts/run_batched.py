"""Small driver for profiling the batched search: python scripts/run_batched.py [rows] [batch] [k] [iters]"""
import os, sys, time
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
from lazzaro_b200 import _lib
_lib.LIB_PATH = os.environ.get('LZ_LIB', _lib.LIB_PATH)
if 'LZ_LIB' in os.environ:                     # an older build for A/B timing: bind only the symbols it has
    import ctypes
    _probe = ctypes.CDLL(_lib.LIB_PATH)
    _lib.SIGNATURES = {k: v for k, v in _lib.SIGNATURES.items() if hasattr(_probe, k)}
import torch
from lazzaro_b200 import synth, kernels as K

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
metric = _lib.METRICS[sys.argv[5]] if len(sys.argv) > 5 else _lib.COSINE
dim = 768
rows, _ = synth.clustered_rows_device(n, dim, max(1, n // 9766), "cuda")
qs = synth.queries_near_rows_device(rows, B)
norms = K.row_norms(rows)
shadow = K.rows_to_shadow(rows, dim, norms, metric=metric)
mx = float(norms.max())
out = K.search_batched(shadow, rows, dim, norms, None, -1, qs, k, metric=metric, max_row_norm=mx)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    out = K.search_batched(shadow, rows, dim, norms, None, -1, qs, k, metric=metric, max_row_norm=mx)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(f"rows={n} B={B} k={k}: {ms:.3f} ms/batch, {2.0*B*n*dim/ms/1e9:.1f} TFLOP/s, certified={out[3].float().mean().item():.3f}")
