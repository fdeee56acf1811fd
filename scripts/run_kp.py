"""Top-k batch search through the store at a given k, with the candidate-buffer size overridable (LZ_KP):
python scripts/run_kp.py [rows] [batch] [k]"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import synth
from lazzaro_b200.store import B200Store
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
k = int(sys.argv[3]) if len(sys.argv) > 3 else 100
rows, _ = synth.clustered_rows_device(n, 768, max(1, n // 9766), "cuda")
st = B200Store(device="cuda:0"); st.adopt_rows(rows, id_prefix="node_")
q = synth.queries_near_rows_device(rows, B)
for _ in range(3): out = st.search_device(q, 0, k)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): out = st.search_device(q, 0, k)
e1.record(); torch.cuda.synchronize()
ref = st.search_device(q[:8].contiguous(), 0, k) if False else None
print(f"LZ_KP={os.environ.get('LZ_KP','-')} rows={n} B={B} k={k}: {e0.elapsed_time(e1)/5:.3f} ms/batch certified={st.last_certified.float().mean().item():.4f}", flush=True)
