"""Batch-1024 search time and certified fraction under the two certification margins, alternating: python scripts/run_margin2.py [rows]"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import synth
from lazzaro_b200.store import B200Store
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
rows, _ = synth.clustered_rows_device(n, 768, max(1, n // 9766), "cuda")
st = B200Store(); st.adopt_rows(rows)
qs = synth.queries_near_rows_device(rows, 1024)
for m in ("statistical", "worst_case", "statistical", "worst_case"):
    st.margin = m
    for _ in range(2):
        st.search_device(qs, 0, 10)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        st.search_device(qs, 0, 10)
    e1.record(); torch.cuda.synchronize()
    print(m, e0.elapsed_time(e1) / 5, flush=True)
