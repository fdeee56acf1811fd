"""Small run of every kernel, for compute-sanitizer (memcheck / racecheck): python scripts/sanitize_smoke.py"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import numpy as np, torch
from lazzaro_b200 import _lib, consolidate, kernels as K, synth
from lazzaro_b200.graph import EdgeTable
from lazzaro_b200.store import B200Store

rows_np, shard = synth.clustered_rows(6000, 200, 8, seed=1)
qs_np = synth.queries_near_rows(rows_np, 20, seed=2)
rows, qs = torch.as_tensor(rows_np).cuda(), torch.as_tensor(qs_np).cuda()
norms = K.row_norms(rows)
for metric in (_lib.COSINE, _lib.L2):
    for B, k in ((1, 10), (3, 7), (4, 100)):
        K.search_exact(rows, 200, norms, None, qs[:B].contiguous(), k, metric)
tags = torch.as_tensor((np.arange(6000) % 3).astype(np.int32)).cuda()
K.search_exact(rows, 200, norms, tags, qs[:5].contiguous(), 9, _lib.COSINE, group_tag=torch.full((5,), 1, dtype=torch.int32, device="cuda"), group_size=1)
shadow = K.rows_to_shadow(rows, 200, norms)
for B, k in ((20, 10), (9, 1), (20, 100), (3, 3)):
    K.search_batched(shadow, rows, 200, norms, None, -1, qs[:B].contiguous(), k)
K.search_batched(shadow, rows, 200, norms, tags, 2, qs, 10, row_range=(100, 5003))
ids, sc, cnt = K.search_exact(rows, 200, norms, None, qs, 5, _lib.COSINE)
K.merge_topk(torch.stack([ids, ids]), torch.stack([sc, sc]), _lib.COSINE)
K.score_pairs(rows, 200, norms, qs, torch.arange(20, dtype=torch.int32, device="cuda"), ids[:, 0].contiguous(), _lib.COSINE)
seg = torch.as_tensor(np.searchsorted(shard, np.arange(9))).cuda()
K.segment_mean(rows, 200, seg)
consolidate.assign_to_centroids(rows, K.segment_mean(rows, 200, seg).float().contiguous())
src, dst, w, cooc = synth.edge_table(70001, 500)
t = EdgeTable.from_arrays(src, dst, w, cooc, device="cuda"); t.decay(0.99); t.decay_prune(0.99, 0.5)
t2 = EdgeTable(device="cuda", capacity=70001, in_place=True); t2.append(src, dst, w, cooc); t2.decay_prune(0.99, 0.5)
s = torch.rand(1001, device="cuda"); K.salience_decay(s, 1001, 0.01, 0.2)
st = B200Store(); st.add_rows(rows_np[:3000], user_id="a"); st.add_rows(rows_np[3000:], user_id="b")
st.search_batch(qs_np, "b", 5); st.search_users(qs_np[:4], ["a", "b", "a", "zz"], 5); st.delete_nodes([], "a")
torch.cuda.synchronize()
print("sanitize smoke done")
