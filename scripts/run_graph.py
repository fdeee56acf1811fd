"""Small driver for profiling the node / adjacency sweeps: python scripts/run_graph.py [edges] [nodes]"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import kernels as K, synth
ne = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
nn = int(sys.argv[2]) if len(sys.argv) > 2 else 12_500_000
g = torch.Generator(device="cuda"); g.manual_seed(5)
sal = torch.rand(nn, device="cuda", generator=g)
acc = torch.randint(0, 30, (nn,), device="cuda", generator=g, dtype=torch.int32)
last = 1.76e9 - torch.rand(nn, device="cuda", generator=g, dtype=torch.float64) * 86400 * 400
for k in (64, 1024):
    for _ in range(2):
        K.evict_lowest(sal, acc, last, 1.76e9, k)
src, dst, w, _ = synth.edge_table(ne, nn)
src, dst, w = (torch.as_tensor(a).cuda() for a in (src, dst, w))
for _ in range(2):
    off, adj = K.csr_build(src, dst, ne, nn)
nodes = torch.randint(0, nn, (4096,), device="cuda", generator=g, dtype=torch.int32)
K.neighbors(off, adj, src, dst, w, nn, nodes, 0.3)
torch.cuda.synchronize()
print("ok")
