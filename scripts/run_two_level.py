"""BASELINE configs[1]: 1M x 768 in 64 shards, batch 256, top-10 — flat tensor-core search vs two-level at several nprobe.
    python scripts/run_two_level.py [rows] [shards] [batch] [k]"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import numpy as np
import torch
from lazzaro_b200 import synth
from lazzaro_b200.store import B200Store

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
S = int(sys.argv[2]) if len(sys.argv) > 2 else 64
B = int(sys.argv[3]) if len(sys.argv) > 3 else 256
k = int(sys.argv[4]) if len(sys.argv) > 4 else 10
dim = 768
rows, shard = synth.clustered_rows_device(n, dim, S, "cuda")
st = B200Store()
st.adopt_rows(rows, id_prefix="r")
off = torch.searchsorted(shard.long(), torch.arange(S + 1, device="cuda")).cpu().numpy()
st.set_shards(off)
qs = synth.queries_near_rows_device(rows, B)


def timed(fn, it=20):
    for _ in range(3):
        out = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it, out


ms, flat = timed(lambda: st.search_device(qs, 0, k))
print(f"flat: {ms:.3f} ms  {B / ms * 1e3:.0f} QPS  certified {float(st.last_certified.float().mean()):.3f}", flush=True)
fid = flat[0]
for nprobe in (1, 2, 4, 8, 16, S):
    if nprobe > S:
        continue
    ms, out = timed(lambda: st.search_two_level_device(qs, k, nprobe))
    ids = out[0]
    rec = float(((ids.unsqueeze(2) == fid.unsqueeze(1)).any(2)).float().mean())
    same = bool(torch.equal(ids, fid))
    print(f"two-level nprobe={nprobe}: {ms:.3f} ms  {B / ms * 1e3:.0f} QPS  recall@{k} {rec:.4f}  identical_to_flat {same}  "
          f"certified {float(st.last_certified.float().mean()):.3f}", flush=True)
