"""Multi-GPU parity check (torchrun, one rank per GPU): the sharded search (local search with id_base, ONE packed
all-gather, strided merge) returns exactly what a single GPU returns over all rows.  Prints OK per case on rank 0.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/check_dist.py
"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
os.environ.setdefault("NCCL_DEBUG", "WARN")
import numpy as np
import torch
import torch.distributed as dist
from lazzaro_b200 import kernels as K, synth, _lib
from lazzaro_b200.dist import ShardedSearch, even_slabs
from lazzaro_b200.store import B200Store

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, dim = 300_000, 128
rows, _ = synth.clustered_rows_device(n, dim, 32, "cuda")           # same seed on every rank: same rows
lo, hi = even_slabs(n, world)[rank]
st = B200Store()
st.adopt_rows(rows[lo:hi].contiguous(), id_prefix="r")
ss = ShardedSearch(st, row_offset=lo, world=world)
norms = K.row_norms(rows)
ok = True
for B, k in ((1, 10), (3, 100), (64, 10), (300, 7)):
    q = synth.queries_near_rows_device(rows, B)
    ids, sc, cnt = ss.search_device(q, k)
    want_ids, want_sc, want_cnt = K.search_exact(rows, dim, norms, None, q, k, _lib.COSINE)
    good = bool((ids == want_ids).all()) and bool((sc == want_sc).all()) and bool((cnt == want_cnt).all())
    h = ss.submit_host(q.cpu().numpy(), k).result()
    good = good and np.array_equal(h[0], want_ids.cpu().numpy())
    flag = torch.tensor([int(good)], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"B={B} k={k}:", "OK" if int(flag) else "MISMATCH", flush=True)
    ok = ok and bool(int(flag))
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
