import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import consolidate, synth
n, dim, kc = 1_000_000, 768, 1024
rows, shard = synth.clustered_rows_device(n, dim, kc, "cuda")
seg = torch.searchsorted(shard.long().contiguous(), torch.arange(kc + 1, device="cuda"))
cen32 = consolidate.centroid_means(rows, seg_offsets=seg).float().contiguous()
for _ in range(2):
    idx, cos = consolidate.assign_to_centroids(rows, cen32)
torch.cuda.synchronize()
print((idx == shard.long()).float().mean().item())
