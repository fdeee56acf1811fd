"""Multi-GPU check of lz_allgather_merge (torchrun, one rank per GPU): the C-ABI exchange (ncclAllGather + strided merge in
one call, own ncclComm_t) returns what the single-GPU exact search over all rows returns, bit for bit.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29535 scripts/check_cabi_exchange.py
"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
os.environ.setdefault("NCCL_DEBUG", "WARN")
import torch
import torch.distributed as dist
from lazzaro_b200 import kernels as K, synth, _lib
from lazzaro_b200.dist import NcclExchange, even_slabs
from lazzaro_b200.store import B200Store

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ex = NcclExchange()
n, dim = 200_000, 128
rows, _ = synth.clustered_rows_device(n, dim, 32, "cuda")
lo, hi = even_slabs(n, world)[rank]
st = B200Store(); st.adopt_rows(rows[lo:hi].contiguous())
norms = K.row_norms(rows)
ok = True
for B, k in ((1, 10), (64, 10), (40, 100)):
    q = synth.queries_near_rows_device(rows, B, seed=B)
    part = torch.empty((2, B, k), dtype=torch.int64, device="cuda")
    cnt = torch.empty((B,), dtype=torch.int32, device="cuda")
    st.search_device(q, 0, k, id_base=lo, out=(part[0], part[1].view(torch.float64), cnt))
    ids, sc, c = ex.allgather_merge(part, st._metric)
    want = K.search_exact(rows, dim, norms, None, q, k, _lib.COSINE)
    torch.cuda.synchronize()
    good = bool((ids == want[0]).all()) and bool((sc == want[1]).all()) and bool((c == want[2]).all())
    flag = torch.tensor([int(good)], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"B={B} k={k}:", "OK" if int(flag) else "MISMATCH", flush=True)
    ok = ok and bool(int(flag))
ex.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
