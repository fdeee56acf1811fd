"""Shard reassignment driver for profiling: python scripts/run_assign.py [slab]
default: centroids = the true shard means (well separated); slab: means of 1024 equal slabs of the rows, i.e. ~10
nearly coincident centroids per true cluster (what k-means with k > #clusters looks like)."""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import consolidate, synth
n, dim, kc = 1_000_000, 768, 1024
slab = len(sys.argv) > 1 and sys.argv[1] == "slab"
rows, shard = synth.clustered_rows_device(n, dim, 102 if slab else kc, "cuda")
seg = torch.linspace(0, n, kc + 1, device="cuda").long() if slab else \
    torch.searchsorted(shard.long().contiguous(), torch.arange(kc + 1, device="cuda"))
cen32 = consolidate.centroid_means(rows, seg_offsets=seg).float().contiguous()
for margin in ("worst_case", "statistical"):
    for _ in range(2):
        idx, cos = consolidate.assign_to_centroids(rows, cen32, margin=margin)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        idx, cos = consolidate.assign_to_centroids(rows, cen32, margin=margin)
    e1.record(); torch.cuda.synchronize()
    print(margin, "ms", e0.elapsed_time(e1) / 3, "own-slab fraction" if slab else "own-shard fraction",
          (idx == (torch.arange(n, device="cuda") * kc // n if slab else shard.long())).float().mean().item(), flush=True)

from lazzaro_b200.store import B200Store
st = B200Store(); st.adopt_rows(rows)
st._shadow_for(0)
for _ in range(2):
    idx2, cos2 = consolidate.assign_store_rows(st, cen32)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    idx2, cos2 = consolidate.assign_store_rows(st, cen32)
e1.record(); torch.cuda.synchronize()
print("resident rows (store shadow as the query side) ms", e0.elapsed_time(e1) / 3, "same answer", bool(torch.equal(idx2, idx)), flush=True)
for _ in range(2):
    m = consolidate.centroid_means(rows, assign=idx, n_centroids=kc)
torch.cuda.synchronize()
e0.record()
for _ in range(3):
    m = consolidate.centroid_means(rows, assign=idx, n_centroids=kc)
e1.record(); torch.cuda.synchronize()
print("centroid update from an arbitrary assignment ms", e0.elapsed_time(e1) / 3, flush=True)
