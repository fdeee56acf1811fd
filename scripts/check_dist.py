"""Multi-GPU parity check (torchrun, one rank per GPU): the sharded search (local search with id_base, ONE packed
all-gather, strided merge) returns exactly what a single GPU returns over all rows (lazzaro_b200.dist.parity_check,
the same check bench.py runs before its timed region when world > 1).  Prints OK per case on rank 0.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/check_dist.py
"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
os.environ.setdefault("NCCL_DEBUG", "WARN")
import torch
import torch.distributed as dist
from lazzaro_b200.dist import parity_check

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
res = parity_check()
if rank == 0:
    for r in res:
        print(f"B={r['batch']} k={r['k']} ({r.get('path', '-')}):", "OK" if r["ok"] else "MISMATCH", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if all(r["ok"] for r in res) else 1)
