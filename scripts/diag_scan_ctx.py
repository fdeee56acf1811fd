"""Times the single-query exact scan over this rank's rows at several points of a multi-process run (before the process
group exists, after it, after the parity check, after sharded searches) to see which step changes its speed."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from lazzaro_b200 import synth, _lib
from lazzaro_b200.store import B200Store
from lazzaro_b200.dist import ShardedSearch, parity_check

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
_lib.load()
n = int(os.environ.get("DIAG_ROWS", "1250000")); dim = 768
rows, _ = synth.clustered_rows_device(n, dim, 128, dev, seed=1 + rank)
store = B200Store(device=str(dev)); store.adopt_rows(rows, id_prefix="node_")
q = synth.queries_near_rows_device(rows, 1, seed=5)
res = {}

def t(tag, steps=20, reps=3):
    best = []
    for _ in range(reps):
        for _ in range(3):
            store.search_device(q, 0, 10)
        torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            store.search_device(q, 0, 10)
        e1.record(); torch.cuda.synchronize()
        best.append(e0.elapsed_time(e1) / steps)
    res[tag] = [round(n * dim * 4 / (b * 1e-3) / 1e9) for b in best]

t("start")
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
    t("after_init")
    dist.barrier(); t("after_barrier")
    parity_check(); t("after_parity")
sharded = ShardedSearch(store, row_offset=rank * n, world=world)
qq = synth.queries_near_rows_device(rows, 1024, seed=6)
for i in range(4):
    sharded.search_device(qq, 10, lane=i & 1)
sharded.join(); t("after_lanes")
for i in range(4):
    sharded.search_device(q, 10)
t("after_sharded_b1")
t("steps10", steps=10); t("steps100", steps=100)
print(json.dumps({"rank": rank, "GBps": res}), flush=True)
if world > 1:
    dist.destroy_process_group()
