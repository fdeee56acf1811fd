#!/usr/bin/env python
"""bench.py — the reference's headline metric on B200: top-10 search QPS at 10M x 768 memories.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N ...

One "step" = one top-k search of a batch of B queries over ALL rows (rows are partitioned across
the N ranks, so total work is fixed: "scaling": "strong").  Prints ONE JSON line on rank 0.

  value     whole-job QPS with rows AND queries already resident in HBM (device-timed, CUDA events
            on the launching stream, max over ranks)
  e2e       the same metric through the public Store call with HOST query buffers and host results
            (pinned H2D of the queries and D2H of ids/scores inside the timed region)
  roofline  batch >= 8: the tensor-core search against the measured bf16 peak, algorithmic flops =
            2 * batch * rows * dim; batch < 8: the exact scan against the measured HBM peak, algorithmic
            bytes = rows * dim * 4 per pass.  The default run reports batch 1024 as `value`
            and adds the batch-1 search (north_star's second target) under "batch1"
  cpu_baseline  the CPU port (oracle.flat_search.fast_topk_f32: float32 BLAS GEMM on all host cores
            + argpartition) on a bounded row sample, scaled linearly to the full row count

`--impl reference` times only that CPU port, as the reference arm: LanceDB, which the reference
calls for this path (core/vector_store.py:132-140), is an unpinned third-party package that is not
installable here, so the arm is the numpy restatement ("kind": "port").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "top-10 search QPS @10M x 768 memories"
UNIT = "queries/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=768)
    ap.add_argument("--shards", type=int, default=1024)
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--no-batch1", action="store_true", help="skip the secondary batch-1 (HBM-bound) measurement")
    ap.add_argument("--no-lanes", dest="lanes", action="store_false",
                    help="time the K steps strictly one after another on one stream instead of on two alternating lanes")
    ap.add_argument("--no-k100", action="store_true", help="skip the top-100 sub-result (BASELINE configs[3] as written)")
    ap.add_argument("--no-configs", action="store_true", help="skip the cfg1 / cfg2 / cfg5_search sub-results")
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--cpu-sample-rows", type=int, default=400_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--edges", type=int, default=50_000_000, help="edge-table size for the decay+prune sub-result (0 = skip)")
    ap.add_argument("--no-consolidation", action="store_true", help="skip the consolidation sub-result (BASELINE configs[2])")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return (d.get("hbm_gbs", 6650.0), d.get("bf16_tflops", 1590.0), "measured (MEASURED_PEAKS.json)",
                d.get("bf16_tflops_sustained", 1400.0))
    return 6650.0, 1590.0, "fallback (B200_PROFILING.md)", 1400.0


class ClockSampler:
    """Samples nvidia-smi SM clocks / throttle reasons while the timed region runs.  One sampler per job: rank 0 polls the
    GPUs of all ranks (every rank polling for itself puts N nvidia-smi processes on the driver's locks, which showed up as
    host-side launch stalls in the 8-GPU end-to-end number); the other ranks sample nothing."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, indices, period_ms: int = 20):
        self.indices, self.period, self.rows, self.proc, self.t0 = list(indices), period_ms, [], None, None

    def __enter__(self):
        if not self.indices:
            return self
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", ",".join(str(i) for i in self.indices), "-lms", str(self.period)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def wait_first(self, timeout: float = 10.0):
        """nvidia-smi takes a second or more to come up on a multi-GPU box: do not start the timed region before it reports."""
        t_end = time.time() + timeout
        while self.proc and not self.rows and time.time() < t_end:
            time.sleep(0.02)

    def mark(self):
        self.t0 = time.time()                        # samples from here on are "under load"

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [r for t, r in self.rows if self.t0 is None or t >= self.t0] or [r for _, r in self.rows]
        for r in rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm), "gpus_sampled": len(self.indices)}


def cpu_search_qps(args, steps, warmup):
    """The CPU port on a bounded sample; returns (QPS scaled to args.rows, description, cores)."""
    import numpy as np
    from lazzaro_b200 import synth
    from oracle import flat_search
    n = min(args.cpu_sample_rows, args.rows)
    rows, _ = synth.clustered_rows(n, args.dim, max(1, args.shards * n // args.rows), seed=synth.SEED_DATA)
    qs = synth.queries_near_rows(rows, max(args.batch, 1), seed=synth.SEED_QUERY)
    inv = (1.0 / np.linalg.norm(rows, axis=1)).astype(np.float32)
    for _ in range(max(1, min(warmup, 2))):
        flat_search.fast_topk_f32(rows, qs, args.k, "cosine", inv)
    t0 = time.perf_counter()
    for _ in range(steps):
        flat_search.fast_topk_f32(rows, qs, args.k, "cosine", inv)
    dt = (time.perf_counter() - t0) / steps
    qps_sample = qs.shape[0] / dt
    scaled = qps_sample * n / args.rows
    desc = (f"numpy float32 sgemm+argpartition flat search, batch {qs.shape[0]}, top-{args.k}, on a {n} x {args.dim} "
            f"row sample; QPS scaled by {n}/{args.rows} (a flat scan is linear in rows)")
    return scaled, desc, os.cpu_count(), dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        # torchrun pins OMP_NUM_THREADS=1; the CPU arm runs alone on rank 0 and may use every host core
        for v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
            os.environ[v] = str(os.cpu_count())
    steps = max(1, min(args.steps, 10))
    qps, desc, cores, dt = cpu_search_qps(args, steps, args.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3 * args.rows / min(args.cpu_sample_rows, args.rows),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload(args),
            "cpu_baseline": {"value": qps, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": qps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "LanceDB (the reference's search engine, vector_store.py:132-140) is unavailable on this box; "
                    "this arm is the numpy port of its flat scan on the host cores"}
    print(json.dumps(line), flush=True)


def workload(args):
    return {"workload": f"top-{args.k} batch-{args.batch} exhaustive cosine search over {args.rows} x {args.dim} fp32 memories in "
                        f"{args.shards} semantic shards (BASELINE metric; configs[3] at k={args.k})",
            "rows": args.rows, "dim": args.dim, "batch": args.batch, "k": args.k,
            "l2_policy": "inputs (rows) are far larger than the 126 MB L2; no flush needed",
            "sub_block_warmup": "each sub-result (batch1, k100, decay_prune) runs 3 steps + 0.25 s of its own work untimed first: "
                                "after the power-capped tensor-core block the SM clock takes tens of ms to come back, which a "
                                "10 ms timed region would otherwise sit inside",
            "issue": ("two lanes: step i+1's coarse pass overlaps step i's exact pass, all-gather and merge (ShardedSearch lanes)"
                      if getattr(args, "lanes", False) and args.batch >= 8 else "one stream, one step after another"),
            "partitioning": f"rows split evenly over {args.gpus} GPU(s) by semantic shard; per-GPU top-k merged after an all-gather"}


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from lazzaro_b200 import _lib, kernels, synth
    from lazzaro_b200.store import B200Store
    from lazzaro_b200.dist import ShardedSearch

    os.environ.setdefault("NCCL_DEBUG", "WARN")          # keep NCCL's version banner off stdout: one JSON line only
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        # torchrun pins OMP_NUM_THREADS=1; the host side of the host-API legs (staging copies, the tenants' gather) may use
        # this rank's share of the host cores
        torch.set_num_threads(max(1, (os.cpu_count() or 1) // world))
    _lib.load()

    # ---- rows of this rank: whole semantic shards, generated in HBM ---------------------------------
    rows_per_rank = args.rows // world
    lo = rank * rows_per_rank
    n_local = rows_per_rank if rank < world - 1 else args.rows - lo
    shards_local = max(1, args.shards // world)
    rows, _ = synth.clustered_rows_device(n_local, args.dim, shards_local, dev, seed=synth.SEED_DATA + rank)
    store = B200Store(device=str(dev))
    store.adopt_rows(rows, id_prefix="node_")
    sharded = ShardedSearch(store, row_offset=lo, world=world)
    # ---- multi-GPU parity record: before anything is timed, the sharded search must equal a single-GPU exact
    # search bit for bit on the same rows (lazzaro_b200.dist.parity_check); a mismatch ends the run with rc != 0
    dist_parity = None
    if world > 1:
        from lazzaro_b200.dist import parity_check
        cases = parity_check()
        dist_parity = {"status": "ok" if all(c["ok"] for c in cases) else "MISMATCH", "cases": cases,
                       "check": "sharded search (local search + one packed all-gather + strided merge) vs lz_search_exact over "
                                "all rows on one GPU: ids, float64 scores, counts bitwise; device and host API"}
        if dist_parity["status"] != "ok":
            if rank == 0:
                print(json.dumps({"metric": METRIC, "dist_parity": dist_parity, "error": "multi-GPU result differs from single-GPU"}), flush=True)
            dist.destroy_process_group()
            sys.exit(3)
    # the same query batch on every rank; query i is drawn near a row of rank i % world, so the true neighbours of
    # a batch are spread over all GPUs (each rank draws near its own rows, the draws are all-gathered)
    queries = synth.queries_near_rows_device(rows, args.batch, seed=synth.SEED_QUERY + rank)
    if world > 1:
        allq = torch.empty((world,) + tuple(queries.shape), dtype=queries.dtype, device=dev)
        dist.all_gather_into_tensor(allq, queries.contiguous())
        pick = torch.arange(args.batch, device=dev) % world
        queries = allq[pick, torch.arange(args.batch, device=dev)].contiguous()
        del allq
    q_host = queries.cpu().numpy()
    torch.cuda.synchronize()

    def step_device():
        return sharded.search_device(queries, args.k)

    def steps_device_lanes(n_):
        # K searches back to back on two lanes (two streams, two scratch sets): the exact pass + all-gather + merge of
        # search i overlap the coarse pass of search i+1; every search is complete when the loop's join() returns
        for i in range(n_):
            out = sharded.search_device(queries, args.k, lane=i & 1)
        sharded.join()
        return out

    def step_e2e():
        if args.batch == 1 and world == 1:
            return store.search_nodes(q_host[0].tolist(), limit=args.k)       # the reference-shaped call
        return sharded.search_host(q_host, args.k)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def host_pipeline(qh, k, steps, depth=2):
        # K searches through the host API with `depth` kept in flight (staging slots i % (depth + 1), two lanes): every
        # step's inputs come from host memory and every step's result is read back to the host inside the timed region
        pend, out = [], None
        for i in range(steps):
            pend.append(sharded.submit_host(qh, k, slot=i % (depth + 1)))
            if len(pend) > depth:
                out = pend.pop(0).result()
        while pend:
            out = pend.pop(0).result()
        return out

    def e2e_block(steps):
        # K steps through the host API with two searches kept in flight: step i+1's staging + H2D copy overlap step
        # i's search; every step's inputs come from host memory and every step's result is read back to the host
        if args.batch == 1 and world == 1:
            for _ in range(steps):
                out = step_e2e()
            return out
        return host_pipeline(q_host, args.k, steps)

    def timed(fn, steps, block=False):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if block:
            out = fn(steps)
        else:
            for _ in range(steps):
                out = fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), out

    for _ in range(max(3, args.warmup)):
        step_device()
        step_e2e()
    e2e_block(2)
    vis = [v for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v.strip().isdigit()]
    job_gpus = ([int(v) for v in vis][:world] if len(vis) >= world else list(range(world))) if rank == 0 else []
    with ClockSampler(job_gpus, 20 if world == 1 else 50) as clocks:
        clocks.wait_first()
        barrier()
        clocks.mark()
        if args.lanes and store._takes_batched_path(args.batch, args.k, n_local) and args.batch >= store.min_batch:
            steps_device_lanes(4)
            ms_total, out = timed(steps_device_lanes, args.steps, block=True)
        else:
            ms_total, out = timed(step_device, args.steps)
        # the dominant kernel alone (local scan + merge, no collective), for the roofline
        ms_kernel, _ = timed(lambda: store.search_device(queries, 0, args.k), args.steps)
        ms_e2e, out_e2e = timed(e2e_block, args.steps, block=True)
    ms_step = ms_total / args.steps
    value = args.batch * 1e3 / ms_step
    e2e = args.batch * 1e3 / (ms_e2e / args.steps)

    # ---- sanity: the timed result is the right answer (self-consistency on device) ------------------
    ids, scores, counts = out
    torch.cuda.synchronize()
    assert int(counts.min()) == args.k, "search returned fewer than k results"
    assert bool((scores[:, 1:] <= scores[:, :-1]).all()), "results not sorted"
    if not (args.batch == 1 and world == 1):
        assert np.array_equal(out_e2e[0], ids.cpu().numpy()), "host-API result differs from the device-API result"

    hbm_peak, tc_burst, peak_src, tc_peak = peaks()       # tensor bound: the sustained figure (timed inside a long loop)
    traffic = {}
    tp = os.path.join(ROOT, "profiles", "traffic.json")           # dram bytes per launch from ncu --set full captures
    if os.path.exists(tp) and args.rows == 10_000_000 and args.dim == 768 and world == 1:
        traffic = json.load(open(tp))
    def roofline_of(batch, k, ms_per_search, is_batched):
        """Roofline of one local search (no collective).  Tensor-core path: the coarse pass reads the fp16 shadow once and does
        2*B*rows*D flops; whichever of the two takes longer at the measured peaks names the bound (the ridge is near batch 210
        at 768 dims).  Exact path: one pass over the float32 rows per group of four queries."""
        if is_batched:
            fl = 2.0 * batch * n_local * args.dim
            by = float(n_local) * store._dim_p * 2
            cert = float(store.last_certified.float().mean().item())
            if fl / (tc_peak * 1e12) >= by / (hbm_peak * 1e9):
                ach = fl / (ms_per_search * 1e-3) / 1e12
                return {"bound": "tensor", "achieved": ach, "peak": tc_peak, "unit": "TFLOP/s", "frac": ach / tc_peak,
                        "traffic": traffic.get("gemm_topk_kernel_10M_b%d" % batch) if k == 10 else None,
                        "kernel": "gemm_topk_kernel (lz_search_batched)", "peak_source": peak_src, "algorithmic_flops_per_launch": fl,
                        "peak_burst": tc_burst, "frac_of_burst": ach / tc_burst, "certified_fraction": cert, "margin": store.margin,
                        "note": "achieved uses the device time of the whole local batched search (query prep + GEMM/top-k + "
                                "rescoring + fallback launches, no collective) inside the timed loop of back-to-back steps, so peak is "
                                "the measured SUSTAINED bf16 figure (power-capped loop); frac_of_burst is against the burst figure"}
            ach = by / (ms_per_search * 1e-3) / 1e9
            return {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                    "traffic": traffic.get("gemm_topk_kernel_10M_b%d" % batch) if k == 10 else None,
                    "kernel": "gemm_topk_kernel (lz_search_batched), memory-bound regime", "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": by, "certified_fraction": cert, "margin": store.margin,
                    "note": "small batches on a large span go through the fp16 shadow: the coarse pass reads rows*dim*2 bytes, half of "
                            "what the exact float32 scan must read, and the answer is the same (fp64 rescoring + certification); "
                            "achieved uses the device time of the whole local search; peak is the measured COPY bandwidth (read + "
                            "write), which a read-only stream can exceed by a few per cent"}
        by = float(n_local) * args.dim * 4 * ((batch + 3) // 4)
        ach = by / (ms_per_search * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                "traffic": traffic.get("scan_kernel_10M_b1") if (batch == 1 and k == 10) else None,
                "kernel": "scan_kernel (lz_search_exact)", "peak_source": peak_src, "algorithmic_bytes_per_launch": by,
                "note": "achieved uses the device time of the local scan+merge launches (no collective)"}

    batched = store._takes_batched_path(args.batch, args.k, n_local)
    roof = roofline_of(args.batch, args.k, ms_kernel / args.steps, batched)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "fp16 coarse (tcgen05, fp32 accumulate) + f64 exact" if batched else "f64", "data": "synthetic", "config": workload(args),
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(args.batch * store._dim_p * 4),
                "d2h_bytes_per_step": int(args.batch * args.k * 16 + args.batch * 4),
                "api": "B200Store.search_nodes (List[float] -> List[str])" if args.batch == 1 and world == 1
                       else "ShardedSearch.submit_host(...).result() (numpy queries -> numpy ids, scores; one search kept in "
                            "flight so a step's staging and H2D copy overlap the previous step's search)"},
        # batched: query prep, sampling GEMM, seed thresholds, GEMM + top-k, exact pass, fallback scan + merge (empty work list)
        "gpu_launches": int(args.steps * ((7 if batched else 2) + (1 if world > 1 else 0))),
        "roofline": roof,
        "clocks": clocks.summary(),
    }
    if dist_parity is not None:
        line["dist_parity"] = dist_parity
    if batched:
        # the same search under the other certification margin (the store default is the deterministic one)
        margins = {store.margin: {"ms_per_step": ms_kernel / args.steps,
                                  "certified_fraction": float(store.last_certified.float().mean().item())}}
        other = "statistical" if store.margin == "worst_case" else "worst_case"
        keep = store.margin
        store.margin = other
        for _ in range(3):
            store.search_device(queries, 0, args.k)
        ms_o, _ = timed(lambda: store.search_device(queries, 0, args.k), args.steps)
        margins[other] = {"ms_per_step": ms_o / args.steps, "certified_fraction": float(store.last_certified.float().mean().item())}
        store.margin = keep
        line["roofline"]["margins"] = margins
        line["roofline"]["margin"] = keep
    def search_block(batch, k, steps):
        """One (batch, k) search configuration over the same resident rows: device-timed value, e2e through the host API,
        roofline of the dominant kernel."""
        q = queries[:batch].contiguous()
        qh = q_host[:batch]
        is_batched = store._takes_batched_path(batch, k, n_local)
        f_dev = lambda: sharded.search_device(q, k)
        f_loc = lambda: store.search_device(q, 0, k)
        if batch == 1 and world == 1:
            f_e2e_block = lambda n_: [store.search_nodes(qh[0].tolist(), limit=k) for _ in range(n_)][-1]
            api = "B200Store.search_nodes (List[float] -> List[str])"
        else:
            f_e2e_block = lambda n_: host_pipeline(qh, k, n_)
            api = "ShardedSearch.submit_host(...).result(), two searches kept in flight"
        # warm-up: 3 steps, then as many more as fill 0.25 s of this block's own work, so that the SM clock has settled
        # after the previous (tensor-core, power-capped) block before anything here is timed; the count is agreed over the
        # ranks (the sharded search has a collective in it, so every rank must make the same number of calls)
        ms_w, _ = timed(lambda: (f_dev(), f_loc()), 3)
        for _ in range(min(2000, int(250.0 / max(ms_w / 3, 1e-3)))):
            f_dev(); f_loc()
        f_e2e_block(2)
        if is_batched and args.lanes:
            # like the headline: K searches on two lanes, complete when the loop's join() returns
            def f_dev_lanes(n_):
                for i in range(n_):
                    o = sharded.search_device(q, k, lane=i & 1)
                sharded.join()
                return o
            f_dev_lanes(2)
            ms_d, _ = timed(f_dev_lanes, steps, block=True)
        else:
            ms_d, _ = timed(f_dev, steps)
        ms_k, _ = timed(f_loc, steps)
        ms_e, _ = timed(f_e2e_block, steps, block=True)
        out = {"batch": batch, "k": k, "value": batch * 1e3 / (ms_d / steps), "unit": UNIT, "ms_per_step": ms_d / steps, "steps": steps,
               "issue": "two lanes" if (is_batched and args.lanes) else "one stream",
               "e2e": {"value": batch * 1e3 / (ms_e / steps), "unit": UNIT, "h2d_bytes_per_step": int(batch * store._dim_p * 4),
                       "d2h_bytes_per_step": int(batch * k * 16 + batch * 4), "api": api}}
        out["roofline"] = roofline_of(batch, k, ms_k / steps, is_batched)
        out["gpu_launches"] = int(steps * ((7 if is_batched else 2) + (1 if world > 1 else 0)))
        if is_batched and batch == 1:
            # north_star's batch-1 target is stated against the HBM roofline of the float32 rows: the exact scan beside it
            keep_mode = store.batched_search
            store.batched_search = "never"
            try:
                for _ in range(3):
                    f_loc()
                ms_x, _ = timed(f_loc, steps)
            finally:
                store.batched_search = keep_mode
            out["speedup_vs_exact_scan"] = (ms_x / steps) / (ms_k / steps)
            out["exact_scan"] = {"ms_per_step": ms_x / steps, "value_local": 1e3 / (ms_x / steps), "unit": UNIT,
                                 "roofline": roofline_of(1, k, ms_x / steps, False),
                                 "note": "the same query through lz_search_exact (float64 scan of the float32 rows, no shadow): what "
                                         "a store with shadow_for_single_queries = False runs"}
        return out

    n_sub = max(args.steps, 10)
    if not args.no_batch1 and args.batch != 1:
        line["batch1"] = search_block(1, args.k, n_sub)
    if not args.no_k100 and args.k != 100:
        # BASELINE configs[3] as written: top-100, batch 1 and batch 1024, same rows
        line["k100"] = {"workload": f"top-100 over the same {args.rows} x {args.dim} rows (BASELINE configs[3] as written)",
                        f"batch{args.batch}": search_block(args.batch, 100, max(5, args.steps // 2)),
                        "batch1": search_block(1, 100, n_sub)}
    if args.edges > 0:
        # north_star (3): decay + prune sweep over the edge table; edges are partitioned with their shards, so the
        # ranks sweep disjoint slices with no collective (total edges fixed: strong scaling like the search)
        del out, out_e2e
        ne = args.edges // world
        g = torch.Generator(device=dev); g.manual_seed(synth.SEED_EDGES + rank)
        cols = [torch.randint(0, max(1, n_local), (ne,), generator=g, device=dev, dtype=torch.int32),
                torch.randint(0, max(1, n_local), (ne,), generator=g, device=dev, dtype=torch.int32),
                torch.rand(ne, generator=g, device=dev, dtype=torch.float32),
                torch.ones(ne, device=dev, dtype=torch.int32)]
        outs = [torch.empty_like(c) for c in cols]
        kept = torch.zeros(1, dtype=torch.int64, device=dev)
        scratch = torch.empty(int(_lib.load().lz_edge_decay_prune_scratch_bytes(ne)), dtype=torch.uint8, device=dev)
        f_edges = lambda: kernels.edge_decay_prune(*cols, ne, 0.99, 0.5, out=tuple(outs), scratch=scratch, kept=kept)
        ms_w, _ = timed(f_edges, 3)
        for _ in range(min(2000, int(250.0 / max(ms_w / 3, 1e-3)))):     # 0.25 s of its own work: clocks settled (see search_block)
            f_edges()
        n_e = max(args.steps, 10)
        ms_e, _ = timed(f_edges, n_e)
        kf = float(kept.item()) / ne
        ebytes = (16 + 16 * kf) * ne
        ach_e = ebytes / (ms_e / n_e * 1e-3) / 1e9
        line["decay_prune"] = {"metric": "decay+prune edges/s over %d edges (factor 0.99, threshold 0.5)" % args.edges,
                               "value": ne * world / (ms_e / n_e * 1e-3), "unit": "edges/s", "ms_per_step": ms_e / n_e,
                               "kept_fraction": kf, "gpu_launches": int(n_e),
                               "roofline": {"bound": "hbm", "achieved": ach_e, "peak": hbm_peak, "unit": "GB/s",
                                            "frac": ach_e / hbm_peak,
                                            "traffic": traffic.get("prune_stream_kernel_50M") if (ne == 50_000_000) else None,
                                            "kernel": "prune_stream_kernel (lz_edge_decay_prune; one launch + one memset of the tile words)",
                                            "algorithmic_bytes_per_launch": ebytes}}
    if world == 1 and not args.no_consolidation and n_local >= 1_000_000 and args.dim == 768:
        # north_star (2) / BASELINE configs[2]: a 10k-node episodic buffer deduplicated against a 1M-node graph (top-1 +
        # cosine gate), then one k-means style shard reassignment with k = 1024 (assign + centroid update) — on the first
        # 1M resident rows; the 1024 starting centroids are the means of 1024 equal slabs of those rows
        from lazzaro_b200 import consolidate
        n_g, n_new, kc = 1_000_000, 10_000, 1024
        graph = rows[:n_g]
        new = synth.queries_near_rows_device(graph, n_new, noise=0.25, seed=synth.SEED_QUERY + 7)
        f_dedup = lambda: store.search_device(new, 0, 1, row_range=(0, n_g))
        # two centroid layouts, both reported, both with ~10 centroids per true cluster (the first 1M rows hold ~102
        # clusters): 1024 random member rows (spread inside their cluster: some rows see two or more centroids within
        # the rescoring margin), and the means of 1024 equal slabs of the rows (nearly coincident inside a cluster: every
        # row sees ~10).  One centroid per cluster — the easy case, 3.1 ms — is timed by scripts/run_configs.py cfg3
        seg = torch.linspace(0, n_g, kc + 1, device=dev).long()
        cen = consolidate.centroid_means(graph, seg_offsets=seg).float().contiguous()
        cen_sep = graph[torch.randperm(n_g, device=dev, generator=torch.Generator(device=dev).manual_seed(9))[:kc]].contiguous()
        f_assign = lambda: consolidate.assign_to_centroids(graph, cen_sep)
        f_assign_co = lambda: consolidate.assign_to_centroids(graph, cen)
        f_assign_res = lambda: consolidate.assign_store_rows(store, cen_sep, (0, n_g))      # the store's own shadow as the query side
        f_assign_co_res = lambda: consolidate.assign_store_rows(store, cen, (0, n_g))
        f_update = lambda: consolidate.centroid_means(graph, seg_offsets=seg)
        parts = {}
        for name, fn, flops in (("dedup_10k_x_1M_top1", f_dedup, 2.0 * n_new * n_g * args.dim),
                                ("assign_1M_x_1024_member_rows", f_assign, 2.0 * n_g * kc * args.dim),
                                ("assign_1M_x_1024_coincident", f_assign_co, 2.0 * n_g * kc * args.dim),
                                ("assign_1M_x_1024_member_rows_resident", f_assign_res, 2.0 * n_g * kc * args.dim),
                                ("assign_1M_x_1024_coincident_resident", f_assign_co_res, 2.0 * n_g * kc * args.dim),
                                ("centroid_update_1M", f_update, None)):
            for _ in range(2):
                fn()
            ms_c, _ = timed(fn, 5)
            parts[name] = {"ms": ms_c / 5}
            if flops:
                parts[name]["tflops"] = flops / (ms_c / 5 * 1e-3) / 1e12
                parts[name]["frac_of_sustained_bf16_peak"] = parts[name]["tflops"] / tc_peak
            else:
                parts[name]["gbs"] = n_g * args.dim * 4 / (ms_c / 5 * 1e-3) / 1e9
                parts[name]["frac_of_hbm_peak"] = parts[name]["gbs"] / hbm_peak
        line["consolidation"] = {"metric": "consolidation wall time: dedup 10k x 1M + shard reassignment (k=1024) + centroid update, 768-d",
                                 "value": sum(parts[k_]["ms"] for k_ in ("dedup_10k_x_1M_top1", "assign_1M_x_1024_member_rows_resident", "centroid_update_1M")),
                                 "value_with_coincident_centroids": sum(parts[k_]["ms"] for k_ in ("dedup_10k_x_1M_top1", "assign_1M_x_1024_coincident_resident", "centroid_update_1M")),
                                 "value_note": "dedup + reassignment of the graph's rows as they sit in the store (its fp16 shadow is the query side) + centroid update; "
                                               "the parts without `_resident` convert the rows on every call (rows that are not in a store)",
                                 "unit": "ms", "higher_is_better": False,
                                 "parts": parts, "gpu_launches": 5 * (7 + 4 * 5 + 1)}
        del new, cen, cen_sep
    if not args.no_configs:
        from lazzaro_b200.dist import TenantSearch
        # ---- BASELINE configs[4], search half: 100k users x 1k memories over 8 GPUs = 12 500 tenants (12.5M x 768 rows)
        # per GPU, WEAK scaling: every rank holds its own tenants, a query is routed by the host to the rank that owns
        # its user and scans that user's 1000 rows only (the `.where(user_id = ...)` pre-filter, vector_store.py:136-139).
        # No collective on the data path.  One step = a host batch of 4096 queries per rank.
        users_per_rank, per_user, qb = 12_500, 1000, 4096
        t_rows, _ = synth.clustered_rows_device(users_per_rank * per_user, args.dim, 1024, dev, seed=synth.SEED_DATA + 100 + rank)
        tstore = B200Store(device=str(dev))
        tstore.adopt_tenants(t_rows, per_user, first_user=rank * users_per_rank)
        tsearch = TenantSearch(tstore, users_per_rank, rank, world)
        rng = np.random.default_rng(4242)                       # the same global batch on every rank; each routes its share
        g_users = rng.integers(0, users_per_rank * world, qb * world)
        mine = tsearch.route(g_users)
        local_users = g_users[mine] - rank * users_per_rank
        tq = synth.queries_near_rows_device(t_rows, qb * world, seed=synth.SEED_QUERY + 9)[: mine.size].contiguous()
        tq_host_all = np.zeros((qb * world, args.dim), np.float32)
        tq_host_all[mine] = tq.cpu().numpy()
        t_prep = tsearch.prepare(local_users)                       # per-query row spans and tags, resident like the queries
        f_t = lambda: tsearch.search_device(tq, t_prep, args.k)
        def f_te(n_):
            # host batches through TenantSearch.submit_host with two kept in flight: the gather of a batch's routed rows into
            # pinned memory and its H2D copy run while the previous batch is being scanned
            pend, out_ = [], None
            for i in range(n_):
                pend.append(tsearch.submit_host(tq_host_all, g_users, args.k, slot=i % 3))
                if len(pend) > 2:
                    out_ = pend.pop(0).result()
            while pend:
                out_ = pend.pop(0).result()
            return out_
        for _ in range(3):
            f_t()
        f_te(3)
        n_t = max(args.steps // 2, 5)
        ms_t, out_t = timed(f_t, n_t)
        ms_te, out_te = timed(f_te, n_t, block=True)
        assert np.array_equal(out_te[1], out_t[0].cpu().numpy()) and int(out_t[2].min()) == args.k
        own = (out_t[0] // per_user).cpu().numpy() == local_users[:, None]
        assert own.all(), "a per-user search returned another tenant's row"
        served = torch.tensor([float(mine.size)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(served)
        t_bytes = mine.size * per_user * args.dim * 4
        line["cfg5_search"] = {
            "workload": f"per-user top-{args.k} search, {users_per_rank * world} users x {per_user} memories x {args.dim} "
                        f"({users_per_rank} users = {users_per_rank * per_user} rows resident per GPU), {qb} queries per rank and step, "
                        "routed to the owning rank by user (BASELINE configs[4], search half)",
            "metric": "per-user-filtered search QPS", "value": float(served.item()) * 1e3 / (ms_t / n_t), "unit": UNIT,
            "ms_per_step": ms_t / n_t, "steps": n_t, "scaling": "weak", "queries_per_step": int(served.item()),
            "e2e": {"value": float(served.item()) * 1e3 / (ms_te / n_t), "unit": UNIT, "h2d_bytes_per_step": int(mine.size * args.dim * 4),
                    "d2h_bytes_per_step": int(mine.size * (args.k * 8 + 4)), "api": "TenantSearch.submit_host(...).result() (numpy queries + user indices -> row ids), two batches kept in flight"},
            "roofline": {"bound": "hbm", "achieved": t_bytes / (ms_t / n_t * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": t_bytes / (ms_t / n_t * 1e-3) / 1e9 / hbm_peak, "traffic": None, "kernel": "scan_kernel (lz_search_exact, one row span per query)",
                         "algorithmic_bytes_per_launch": t_bytes},
            "gpu_launches": int(n_t * 2)}
        if rank == 0 and not args.no_cpu_baseline:
            from oracle import flat_search
            sample = t_rows[: 2000 * per_user].cpu().numpy()              # 2000 tenants
            qs_c = tq_host_all[mine][:2000]
            t0 = time.perf_counter()
            for i in range(2000):                                          # one filtered flat search per query: the reference's call
                flat_search.fast_topk_f32(sample[i * per_user:(i + 1) * per_user], qs_c[i:i + 1], args.k)
            dt = time.perf_counter() - t0
            line["cfg5_search"]["cpu_baseline"] = {"value": 2000 / dt, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                                   "sample": "2000 per-user searches (numpy sgemv + argpartition over the user's 1000 x 768 rows), one after another"}
            del sample
        del tstore, tsearch, t_rows, tq
        torch.cuda.empty_cache()
    if not args.no_configs and world == 1:
        # ---- BASELINE configs[1]: 1M x 768 in 64 semantic shards, batch 256, top-10 — flat and two-level (nprobe 4), the
        # two-level search once with queries spread over all 64 shards (every shard is touched: HBM-bound like the flat one)
        # and once with the batch concentrated on 8 shards (a session's topics)
        n2, S2, B2 = 1_000_000, 64, 256
        r2, sh2 = synth.clustered_rows_device(n2, args.dim, S2, dev, seed=synth.SEED_DATA + 7)
        st2 = B200Store(device=str(dev))
        st2.adopt_rows(r2, id_prefix="node_")
        off2 = torch.searchsorted(sh2.long().contiguous(), torch.arange(S2 + 1, device=dev)).cpu().numpy()
        st2.set_shards(off2)
        q2 = synth.queries_near_rows_device(r2, B2, seed=synth.SEED_QUERY + 2)
        q2c = synth.queries_near_rows_device(r2[: int(off2[8])], B2, seed=synth.SEED_QUERY + 3)     # near shards 0..7 only
        q2_host, q2c_host = q2.cpu().numpy(), q2c.cpu().numpy()
        n_2 = max(args.steps, 10)
        res2 = {}
        for name, fn, fe in (("flat", lambda: st2.search_device(q2, 0, args.k), lambda: st2.search_batch(q2_host, limit=args.k)),
                             ("two_level_nprobe4", lambda: st2.search_two_level_device(q2, args.k, 4), lambda: st2.search_two_level(q2_host, args.k, 4)),
                             ("flat_concentrated", lambda: st2.search_device(q2c, 0, args.k), None),
                             ("two_level_nprobe4_concentrated", lambda: st2.search_two_level_device(q2c, args.k, 4), None)):
            for _ in range(3):
                fn()
            ms_2, o2 = timed(fn, n_2)
            res2[name] = {"ms_per_step": ms_2 / n_2, "value": B2 * 1e3 / (ms_2 / n_2), "unit": UNIT, "ids": o2[0]}
            if fe is not None:
                fe()
                ms_2e, _ = timed(fe, n_2)
                res2[name]["e2e"] = {"value": B2 * 1e3 / (ms_2e / n_2), "unit": UNIT, "h2d_bytes_per_step": int(B2 * args.dim * 4),
                                     "d2h_bytes_per_step": int(B2 * (args.k * 8 + 4)),
                                     "api": "B200Store.search_batch / search_two_level (numpy queries -> id strings)"}
        for name, base in (("two_level_nprobe4", "flat"), ("two_level_nprobe4_concentrated", "flat_concentrated")):
            a_, b_ = res2[name]["ids"], res2[base]["ids"]
            res2[name]["recall_at_%d" % args.k] = float((a_.unsqueeze(2) == b_.unsqueeze(1)).any(2).float().mean().item())
        fl2 = 2.0 * B2 * n2 * args.dim
        for name, v in res2.items():
            v.pop("ids")
            scanned = n2 if name.startswith("flat") else 4 * n2 // S2
            tf = 2.0 * B2 * scanned * args.dim / (v["ms_per_step"] * 1e-3) / 1e12
            v["roofline"] = {"bound": "tensor", "achieved": tf, "peak": tc_peak, "unit": "TFLOP/s", "frac": tf / tc_peak, "traffic": None,
                             "kernel": "gemm_topk_kernel", "algorithmic_flops_per_launch": 2.0 * B2 * scanned * args.dim,
                             "hbm_bound_alternative": {"achieved": n2 * args.dim * 2 / (v["ms_per_step"] * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                                       "note": "fp16 shadow bytes of all rows / time (batch 256 sits at the ridge: SURVEY §8d)"}}
        line["cfg2"] = {"workload": f"{n2} x {args.dim} in {S2} semantic shards, batch {B2}, top-{args.k} (BASELINE configs[1])", **res2,
                        "note": "nprobe < shards is an approximation of the flat search: reported with recall, never as parity; "
                                "nprobe = shards equals the flat result bit for bit (tests/test_gpu_two_level.py)"}
        if not args.no_cpu_baseline:
            from oracle import flat_search
            rs = r2[:200_000].cpu().numpy()
            t0 = time.perf_counter()
            flat_search.fast_topk_f32(rs, q2_host, args.k)
            dt = time.perf_counter() - t0
            line["cfg2"]["cpu_baseline"] = {"value": B2 / (dt * n2 / 200_000), "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                            "sample": "numpy sgemm + argpartition flat search of the 256 queries over a 200k-row sample, scaled by 5 to 1M rows"}
            del rs
        del st2, r2, q2, q2c
        torch.cuda.empty_cache()
        # ---- BASELINE configs[0] (the reference's own scale): 1 user, 10k x 384, top-10 through the Store call
        r1_np, _ = synth.clustered_rows(10_000, 384, 8, seed=synth.SEED_DATA + 1)
        q1_np = synth.queries_near_rows(r1_np, 64, seed=synth.SEED_QUERY + 1)
        st1 = B200Store(device=str(dev))
        st1.add_rows(r1_np, id_prefix="node_")
        q1_lists = [q.tolist() for q in q1_np]
        for q in q1_lists[:8]:
            st1.search_nodes(q, limit=10)
        n_1 = 200
        ms_1, _ = timed(lambda n_: [st1.search_nodes(q1_lists[i % 64], limit=10) for i in range(n_)][-1], n_1, block=True)
        q1_dev = st1._to_device(q1_np[:1])
        ms_1d, _ = timed(lambda: st1.search_device(q1_dev, 0, 10), n_1)
        line["cfg1"] = {"workload": "1 user, 10k x 384 memories, top-10, one query per call (BASELINE configs[0], the reference's own scale)",
                        "value": 1e3 / (ms_1d / n_1), "unit": UNIT, "ms_per_step": ms_1d / n_1, "steps": n_1,
                        "e2e": {"value": 1e3 / (ms_1 / n_1), "unit": UNIT, "us_per_call": ms_1 / n_1 * 1e3, "h2d_bytes_per_step": 384 * 4,
                                "d2h_bytes_per_step": 10 * 16 + 4, "api": "B200Store.search_nodes (List[float] -> List[str])"},
                        "roofline": {"bound": "hbm", "achieved": 10_000 * 384 * 4 / (ms_1d / n_1 * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                     "frac": 10_000 * 384 * 4 / (ms_1d / n_1 * 1e-3) / 1e9 / hbm_peak, "traffic": None,
                                     "kernel": "scan_kernel + merge_kernel", "algorithmic_bytes_per_launch": 10_000 * 384 * 4,
                                     "note": "15 MB: launch-latency bound, not bandwidth bound"}}
        if not args.no_cpu_baseline:
            from oracle import flat_search, ref_loops
            from oracle.flat_store import FlatStore
            inv1 = (1.0 / np.linalg.norm(r1_np, axis=1)).astype(np.float32)
            names1 = [f"node_{i}" for i in range(10_000)]
            t0 = time.perf_counter()
            for i in range(n_1):                                     # the same call shape as search_nodes: List[float] in, List[str] out
                qq = np.asarray(q1_lists[i % 64], dtype=np.float32).reshape(1, -1)
                got = flat_search.fast_topk_f32(r1_np, qq, 10, "cosine", inv1)
                [names1[r] for r in np.asarray(got[0]).reshape(-1).tolist()]
            dt_f = (time.perf_counter() - t0) / n_1
            t0 = time.perf_counter()
            for i in range(n_1):                                     # the arithmetic alone: float32 array in, row indices out
                flat_search.fast_topk_f32(r1_np, q1_np[i % 64:i % 64 + 1], 10, "cosine", inv1)
            dt_a = (time.perf_counter() - t0) / n_1
            fs = FlatStore(metric="cosine")
            fs.add_nodes([{"id": f"node_{i}", "content": "", "embedding": r1_np[i]} for i in range(10_000)])
            resident = {f"node_{i}": i for i in range(10_000)}
            t0 = time.perf_counter()
            for i in range(20):                                      # C1: search_memories over the injected float64 flat store
                ref_loops.search_memories(fs, resident, q1_lists[i % 64], limit=10)
            dt_r = (time.perf_counter() - t0) / 20
            line["cfg1"]["cpu_baseline"] = {"value": 1.0 / dt_f, "unit": UNIT, "us_per_call": dt_f * 1e6, "cores": os.cpu_count(), "kind": "port",
                                            "sample": "numpy float32 sgemv + argpartition over the 10k x 384 rows, one query per call, List[float] in and id strings out like "
                                                      "search_nodes (the arithmetic alone, array in / indices out: %.1f us)" % (dt_a * 1e6),
                                            "reference_shaped": {"value": 1.0 / dt_r, "unit": UNIT, "us_per_call": dt_r * 1e6, "cores": 1,
                                                                 "sample": "search_memories restated (memory_system.py:1460-1472) over the oracle FlatStore: "
                                                                           "float64 distances over the fp32 rows in numpy, ids -> resident nodes (BASELINE.md C1)"}}
        del st1
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        if "consolidation" in line:
            # the reference's dedup is one flat scan per new memory (memory_system.py:719-742): numpy port on a bounded sample
            rs = np.random.default_rng(5).standard_normal((50_000, args.dim)).astype(np.float32)
            qs_ = np.random.default_rng(6).standard_normal((1_000, args.dim)).astype(np.float32)
            t0 = time.perf_counter()
            from oracle import flat_search
            flat_search.fast_topk_f32(rs, qs_, 1)
            dt = time.perf_counter() - t0
            line["consolidation"]["cpu_baseline"] = {"value": dt * 1e3 * (1_000_000 / 50_000) * (10_000 / 1_000), "unit": "ms",
                                                     "cores": os.cpu_count(), "kind": "port",
                                                     "sample": "numpy sgemm + argmax dedup of 1000 x 50000 rows, scaled linearly to 10k x 1M (dedup part only)"}
        qps, desc, cores, _ = cpu_search_qps(args, 3, 1)
        line["cpu_baseline"] = {"value": qps, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
        if "consolidation" in line:
            # the reference's own linking loop (memory_system.py:838-891 restated, oracle/ref_loops.py): per new
            # memory a Python cosine against every existing node — 20 new x 10k existing, 768-d
            from oracle import ref_loops
            rng_ = np.random.default_rng(8)
            ex = [ref_loops.Node(f"e{i}", v.tolist()) for i, v in enumerate(rng_.standard_normal((10_000, args.dim)))]
            nw = [ref_loops.Node(f"x{i}", v.tolist()) for i, v in enumerate(rng_.standard_normal((20, args.dim)))]
            t0 = time.perf_counter()
            ref_loops.link_to_existing(nw, ex, {})
            dt = time.perf_counter() - t0
            line["consolidation"]["cpu_baseline"]["reference_shaped"] = {
                "value": 20 * 10_000 / dt, "unit": "pairs/s", "cores": 1,
                "sample": "the reference's per-pair Python cosine loop (memory_system.py:838-891 restated) on 20 new x 10k existing 768-d nodes; "
                          "the same linking at 10k x 1M is 1e10 pairs", "seconds_for_10k_x_1M_extrapolated": 1e10 / (20 * 10_000 / dt)}
            del ex, nw
        if "decay_prune" in line:
            from oracle import graph_ops
            es, ed, ew, ec = synth.edge_table(2_000_000, 1_000_000)
            t0 = time.perf_counter()
            for _ in range(3):
                graph_ops.decay_prune_edges(es, ed, ew, ec, 0.99, 0.5)
            line["decay_prune"]["cpu_baseline"] = {"value": 2_000_000 * 3 / (time.perf_counter() - t0), "unit": "edges/s",
                                                   "cores": 1, "kind": "port",
                                                   "sample": "numpy vectorised decay+mask+compress on 2M edges"}
            # the reference's own per-object loop (memory_shard.py:64-84 restated) on 1M Edge objects
            from oracle import ref_loops
            es, ed, ew, _ = synth.edge_table(1_200_000, 1_000_000)
            ed_ = ref_loops.make_edge_dict(es, ed, ew)
            n_obj = len(ed_)
            t0 = time.perf_counter()
            ref_loops.decay_and_prune(ed_, {}, 0.01, 0.5)
            dt = time.perf_counter() - t0
            line["decay_prune"]["cpu_baseline"]["reference_shaped"] = {
                "value": n_obj / dt, "unit": "edges/s", "cores": 1,
                "sample": f"apply_temporal_decay + prune_weak_edges as the reference writes them (per-object Python loop over a dict of {n_obj} Edge objects); "
                          "linear in the edge count"}
            del ed_
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
