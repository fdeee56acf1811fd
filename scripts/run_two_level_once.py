"""One configuration of the two-level search for launch lists: python scripts/run_two_level_once.py rows shards batch k nprobe"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import synth
from lazzaro_b200.store import B200Store
n, S, B, k, nprobe = (int(a) for a in sys.argv[1:6])
rows, shard = synth.clustered_rows_device(n, 768, S, "cuda")
st = B200Store(); st.adopt_rows(rows, id_prefix="r")
off = torch.searchsorted(shard.long(), torch.arange(S + 1, device="cuda")).cpu().numpy()
st.set_shards(off)
qs = synth.queries_near_rows_device(rows, B)
for _ in range(4):
    out = st.search_two_level_device(qs, k, nprobe=nprobe)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    out = st.search_two_level_device(qs, k, nprobe=nprobe)
e1.record(); torch.cuda.synchronize()
print(f"nprobe={nprobe}: {e0.elapsed_time(e1)/10:.3f} ms")
