"""Where the host-API search loses time against the device-API one at N ranks: torchrun ... scripts/diag_e2e.py"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from lazzaro_b200 import synth, _lib
from lazzaro_b200.store import B200Store
from lazzaro_b200.dist import ShardedSearch

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
_lib.load()
n = 10_000_000 // world; dim = 768; B = 1024; k = 10
rows, _ = synth.clustered_rows_device(n, dim, 128, dev, seed=1 + rank)
store = B200Store(device=str(dev)); store.adopt_rows(rows, id_prefix="node_")
sh = ShardedSearch(store, row_offset=rank * n, world=world)
q = synth.queries_near_rows_device(rows, B, seed=5)
if world > 1:
    dist.broadcast(q, 0)
qh = q.cpu().numpy()
res = {}

def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier(); torch.cuda.synchronize()

def wall(fn, steps=30):
    fn(6); sync()
    t0 = time.perf_counter(); fn(steps); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    t = torch.tensor([dt], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return round(float(t.item()) / steps * 1e3, 3)

def dev_lanes(nn):
    for i in range(nn):
        sh.search_device(q, k, lane=i & 1)
    sh.join()

def host(depth):
    def run(nn):
        pend = []
        for i in range(nn):
            pend.append(sh.submit_host(qh, k, slot=i % (depth + 1)))
            if len(pend) > depth:
                pend.pop(0).result()
        while pend:
            pend.pop(0).result()
    return run

def host_cpu_only(nn):      # how long the host side of one submit takes when nothing is waited for
    t0 = time.perf_counter()
    hs = [sh.submit_host(qh, k, slot=i % 4) for i in range(4)]
    dt = (time.perf_counter() - t0) / 4
    for h in hs: h.result()
    return dt

res["device_lanes_ms"] = wall(dev_lanes)
for d in (1, 2, 3):
    res[f"host_depth{d}_ms"] = wall(host(d))
res["submit_cpu_ms"] = round(host_cpu_only(4) * 1e3, 3)
t0 = time.perf_counter()
for _ in range(20):
    stage = store._pinned("q0", (B, store._dim_p), torch.float32); stage[:, :dim] = torch.from_numpy(qh)
res["stage_copy_ms"] = round((time.perf_counter() - t0) / 20 * 1e3, 3)
res["device_lanes_again_ms"] = wall(dev_lanes)
res["threads"] = torch.get_num_threads()
if rank == 0:
    print(json.dumps(res), flush=True)
if world > 1:
    dist.destroy_process_group()
