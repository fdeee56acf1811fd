"""Times the BASELINE.json configurations that are parity-test cases rather than bench lines (1 GPU),
with CUDA events, and prints one JSON line per measurement.  Numbers quoted in DESIGN.md come from here.

    python scripts/run_configs.py [cfg2] [cfg3] [cfg4] [cfg5]
"""
import json
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import numpy as np
import torch

from lazzaro_b200 import _lib, consolidate, kernels as K, synth
from lazzaro_b200.graph import EdgeTable
from lazzaro_b200.store import B200Store

PEAKS = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) \
    if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters, out


def emit(name, ms, **kw):
    print(json.dumps({"config": name, "ms": round(ms, 4), **{k: (round(v, 4) if isinstance(v, float) else v) for k, v in kw.items()}}), flush=True)


def cfg2():
    n, dim, B, k = 1_000_000, 768, 256, 10
    rows, shard = synth.clustered_rows_device(n, dim, 64, "cuda")
    st = B200Store(); st.adopt_rows(rows, id_prefix="node_")
    q = synth.queries_near_rows_device(rows, B)
    ms, _ = timed(lambda: st.search_device(q, 0, k), 20)
    emit("cfg2 flat batched 1Mx768 B=256 k=10", ms, qps=B / ms * 1e3, tflops=2.0 * B * n * dim / ms / 1e9,
         frac_tensor=2.0 * B * n * dim / ms / 1e9 / PEAKS["bf16_tflops"], frac_hbm_bf16=n * dim * 2 / ms / 1e6 / PEAKS["hbm_gbs"])
    off = np.searchsorted(shard.cpu().numpy(), np.arange(65))
    st.set_shards(off)
    qh = q.cpu().numpy()
    flat = st.search_batch(qh, limit=k)
    for nprobe in (64, 4, 1):
        ms, got = timed(lambda: st.search_two_level(qh, limit=k, nprobe=nprobe), 5)
        recall = float(np.mean([len(set(a) & set(b)) / k for a, b in zip(got, flat)]))
        emit(f"cfg2 two-level nprobe={nprobe} (host in/out)", ms, qps=B / ms * 1e3, recall_at_10=recall)


def cfg3():
    n, dim, n_new, kc = 1_000_000, 768, 10_000, 1024
    rows, shard = synth.clustered_rows_device(n, dim, kc, "cuda")
    st = B200Store(); st.adopt_rows(rows, id_prefix="g")
    new = synth.queries_near_rows_device(rows, n_new, noise=0.25)
    ms, _ = timed(lambda: st.search_device(new, 0, 1), 5)
    emit("cfg3 dedup 10k x 1M top-1", ms, tflops=2.0 * n_new * n * dim / ms / 1e9, frac_tensor=2.0 * n_new * n * dim / ms / 1e9 / PEAKS["bf16_tflops"])
    ms, _ = timed(lambda: st.search_device(new, 0, 3), 5)
    emit("cfg3 link 10k x 1M top-3", ms, tflops=2.0 * n_new * n * dim / ms / 1e9)
    seg = torch.searchsorted(shard.long().contiguous(), torch.arange(kc + 1, device="cuda"))
    ms, cen = timed(lambda: consolidate.centroid_means(rows, seg_offsets=seg), 5)
    emit("cfg3 centroid update (contiguous shards) 1M x 768 -> 1024", ms, gbs=n * dim * 4 / ms / 1e6, frac_hbm=n * dim * 4 / ms / 1e6 / PEAKS["hbm_gbs"])
    cen32 = cen.float().contiguous()
    ms, (idx, _) = timed(lambda: consolidate.assign_to_centroids(rows, cen32), 3)
    emit("cfg3 k-means assign 1M x 1024", ms, tflops=2.0 * n * kc * dim / ms / 1e9, frac_tensor=2.0 * n * kc * dim / ms / 1e9 / PEAKS["bf16_tflops"])
    ms, _ = timed(lambda: consolidate.centroid_means(rows, assign=idx, n_centroids=kc), 3)
    emit("cfg3 centroid update (arbitrary assignment)", ms, gbs=n * dim * 4 / ms / 1e6)


def cfg4():
    n, dim = 10_000_000, 768
    rows, _ = synth.clustered_rows_device(n, dim, 1024, "cuda")
    st = B200Store(); st.adopt_rows(rows, id_prefix="node_")
    for B, k in ((1, 100), (1024, 100), (1, 10), (1024, 10)):
        q = synth.queries_near_rows_device(rows, B)
        ms, _ = timed(lambda: st.search_device(q, 0, k), 5)
        if B == 1:
            emit(f"cfg4 1 GPU B={B} k={k}", ms, qps=B / ms * 1e3, gbs=n * dim * 4 / ms / 1e6, frac_hbm=n * dim * 4 / ms / 1e6 / PEAKS["hbm_gbs"])
        else:
            emit(f"cfg4 1 GPU B={B} k={k}", ms, qps=B / ms * 1e3, tflops=2.0 * B * n * dim / ms / 1e9,
                 frac_tensor=2.0 * B * n * dim / ms / 1e9 / PEAKS["bf16_tflops"], certified=float(st.last_certified.float().mean()))


def cfg4_two_level():
    # approximate two-level search at batch 1: centroid routing, then only the nprobe best shards are scanned
    n, dim, S = 10_000_000, 768, 1024
    rows, shard = synth.clustered_rows_device(n, dim, S, "cuda")
    st = B200Store(); st.adopt_rows(rows, id_prefix="node_")
    st.set_shards(np.searchsorted(shard.cpu().numpy(), np.arange(S + 1)))
    q = synth.queries_near_rows_device(rows, 64).cpu().numpy()
    flat = st.search_batch(q, limit=10)
    for nprobe in (1, 8, 32):
        got = [st.search_two_level(q[i:i + 1], limit=10, nprobe=nprobe)[0] for i in range(64)]
        recall = float(np.mean([len(set(a) & set(b)) / 10 for a, b in zip(got, flat)]))
        ms, _ = timed(lambda: st.search_two_level(q[:1], limit=10, nprobe=nprobe), 20)
        emit(f"cfg4 two-level batch 1, nprobe={nprobe} of {S} shards (host in/out)", ms, qps=1e3 / ms, recall_at_10=recall)
    ms, _ = timed(lambda: st.search_nodes(q[0].tolist(), limit=10), 10)
    emit("cfg4 flat batch 1 via search_nodes (host in/out)", ms, qps=1e3 / ms)


def cfg5():
    # one GPU's share of 100k users x 1k memories under weak scaling over 8 GPUs: 12 500 users, D = 768
    users, per, dim = 12_500, 1000, 768
    n = users * per
    rows, _ = synth.clustered_rows_device(n, dim, 1024, "cuda")
    st = B200Store(); st.adopt_rows(rows, id_prefix="m")
    B = 4096
    q = synth.queries_near_rows_device(rows, B)
    u = torch.randint(0, users, (B,), generator=torch.Generator().manual_seed(1)).numpy()
    lo, hi = u.astype(np.int64) * per, (u.astype(np.int64) + 1) * per
    tag = np.full(B, -1, np.int32)
    ms, _ = timed(lambda: st.search_ranges_device(q, lo, hi, tag, 10), 5)
    emit("cfg5 per-user search, 4096 queries x 1000-row tenants (12.5M x 768 resident)", ms, qps=B / ms * 1e3,
         gbs=B * per * dim * 4 / ms / 1e6, frac_hbm=B * per * dim * 4 / ms / 1e6 / PEAKS["hbm_gbs"])
    del st, rows
    torch.cuda.empty_cache()
    ne = 50_000_000
    src, dst, w, cooc = synth.edge_table(ne, 12_500_000)
    cols = [torch.as_tensor(a).cuda() for a in (src, dst, w, cooc)]
    outs = [torch.empty_like(c) for c in cols]
    kept = torch.zeros(1, dtype=torch.int64, device="cuda")
    scratch = torch.empty(int(_lib.load().lz_edge_decay_prune_scratch_bytes(ne)), dtype=torch.uint8, device="cuda")
    ms, _ = timed(lambda: K.edge_decay_prune(*cols, ne, 0.99, 0.5, out=tuple(outs), scratch=scratch, kept=kept), 10)
    kf = kept.item() / ne
    emit("cfg5 decay+prune 50M edges (out of place)", ms, edges_per_s=ne / ms * 1e3, kept_fraction=kf,
         gbs=(16 + 16 * kf) * ne / ms / 1e6, frac_hbm=(16 + 16 * kf) * ne / ms / 1e6 / PEAKS["hbm_gbs"])
    ms, _ = timed(lambda: K.edge_decay(cols[2], ne, 0.99), 10)
    emit("cfg5 decay only 50M edges", ms, edges_per_s=ne / ms * 1e3, gbs=8 * ne / ms / 1e6, frac_hbm=8 * ne / ms / 1e6 / PEAKS["hbm_gbs"])
    sal = torch.rand(12_500_000, device="cuda")
    ms, _ = timed(lambda: K.salience_decay(sal, sal.numel(), 0.01, 0.2), 10)
    emit("cfg5 salience decay 12.5M nodes", ms, gbs=8 * sal.numel() / ms / 1e6)


def cfg5_graph():
    # the node / adjacency sweeps at one GPU's cfg5 share: 12.5M nodes, 50M edges
    nn, ne = 12_500_000, 50_000_000
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    sal = torch.rand(nn, device="cuda", generator=g)
    acc = torch.randint(0, 30, (nn,), device="cuda", generator=g, dtype=torch.int32)
    last = 1.76e9 - torch.rand(nn, device="cuda", generator=g, dtype=torch.float64) * 86400 * 400
    skip = (torch.rand(nn, device="cuda", generator=g) < 0.01).to(torch.uint8)
    for k in (64, 1024):
        ms, _ = timed(lambda: K.evict_lowest(sal, acc, last, 1.76e9, k, skip=skip), 10)
        emit(f"cfg5 eviction sweep: importance + lowest-{k} of 12.5M nodes", ms, nodes_per_s=nn / ms * 1e3,
             gbs=17 * nn / ms / 1e6, frac_hbm=17 * nn / ms / 1e6 / PEAKS["hbm_gbs"])
    del sal, acc, last, skip
    src, dst, w, _ = synth.edge_table(ne, nn)
    src, dst, w = (torch.as_tensor(a).cuda() for a in (src, dst, w))
    ms, (off, adj) = timed(lambda: K.csr_build(src, dst, ne, nn), 5)
    passes = 3                                                      # bits(12.5M + 1) = 24
    moved = ne * (8 + 16) + nn * 12 + passes * 2 * ne * (4 + 16)    # entries written, degrees scanned, 3 x (hist read + pairs read/written)
    emit("cfg5 adjacency build: CSR over 50M edges / 12.5M nodes (stable radix sort, 3 passes)", ms,
         edges_per_s=ne / ms * 1e3, gbs_moved=moved / ms / 1e6, frac_hbm_moved=moved / ms / 1e6 / PEAKS["hbm_gbs"])
    for q in (1, 4096, 1_000_000):
        nodes = torch.randint(0, nn, (q,), device="cuda", generator=g, dtype=torch.int32)
        ms, _ = timed(lambda: K.neighbors(off, adj, src, dst, w, nn, nodes, 0.3, max_out=64), 10)
        emit(f"cfg5 get_neighbors on the CSR, {q} nodes per launch (mean degree 8)", ms, lookups_per_s=q / ms * 1e3)


if __name__ == "__main__":
    which = sys.argv[1:] or ["cfg2", "cfg3", "cfg4", "cfg5"]
    for name in which:
        globals()[name]()
        torch.cuda.empty_cache()
