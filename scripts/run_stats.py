"""Per-role cycle statistics of gemm_topk_kernel from a -DLZ_STATS build (experiments only)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
from lazzaro_b200 import _lib
_lib.LIB_PATH = os.environ.get("LZ_LIB", _lib.LIB_PATH)
import torch
from lazzaro_b200 import synth, kernels as K
lib = _lib.load()
n, B, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rows, _ = synth.clustered_rows_device(n, 768, max(1, n // 9766), "cuda")
qs = synth.queries_near_rows_device(rows, B)
norms = K.row_norms(rows); shadow = K.rows_to_shadow(rows, 768, norms)
K.search_batched(shadow, rows, 768, norms, None, -1, qs, k)
buf = (C.c_ulonglong * 8)()
lib.lz_debug_stats.argtypes = [C.c_void_p, C.c_int]
lib.lz_debug_stats(buf, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); K.search_batched(shadow, rows, 768, norms, None, -1, qs, k); e1.record(); torch.cuda.synchronize()
lib.lz_debug_stats(buf, 0)
v = list(buf)
tiles = v[5] / 4
print(f"ms={e0.elapsed_time(e1):.2f} tiles(cta)={tiles:.0f}  epi busy/tile={v[0]/max(v[5],1):.0f} cyc  epi max={v[1]}  "
      f"slow tiles(>6000)={v[6]/max(v[5],1)*100:.1f}% excess/tile={v[7]/max(v[5],1):.0f}  "
      f"mma wait tempty/tile={v[2]/tiles:.0f}  mma wait full/tile={v[3]/tiles:.0f}  prod wait empty/tile={v[4]/tiles:.0f}")
