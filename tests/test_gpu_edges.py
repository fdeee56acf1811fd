"""GPU parity: decay / prune / salience sweeps (through the C-ABI) against the numpy oracle and
against outputs of the reference's own MemoryShard code (tests/golden). Bit-exact."""
import numpy as np
import pytest
import torch

from oracle import graph_ops

pytestmark = pytest.mark.gpu


def table(src, dst, w, cooc):
    from lazzaro_b200.graph import EdgeTable
    return EdgeTable.from_arrays(src, dst, w, cooc, device="cuda")


def test_decay_prune_matches_reference_shard(golden):
    g = golden
    t = table(g["dp_src"], g["dp_dst"], g["dp_w"], g["dp_cooc"])
    pruned = t.decay_prune(graph_ops.decay_factor(0.01), 0.5)          # memory_shard.py:64-84 fused
    s, d, w, c = t.to_numpy()
    assert pruned == int(g["dp_pruned"])
    np.testing.assert_array_equal(np.stack([s, d], 1), g["dp_kept"])   # same survivors, same order
    np.testing.assert_array_equal(w, g["dp_kept_w"].astype(np.float32))
    assert (c == 1).all()


def test_separate_decay_then_prune_equals_reference(golden):
    g = golden
    t = table(g["dp_src"], g["dp_dst"], g["dp_w"], g["dp_cooc"])
    t.decay(0.99)
    np.testing.assert_array_equal(t.to_numpy()[2], g["dp_w_after"].astype(np.float32))
    # pruning the float32-rounded weights may differ from the float64 reference only inside the
    # documented tie band |w - thr| <= 2^-24 * thr
    pruned = t.decay_prune(1.0, 0.5)
    w64 = g["dp_w_after"]
    band = np.abs(w64 - 0.5) <= 2.0 ** -24 * 0.5
    assert abs(pruned - int(g["dp_pruned"])) <= int(band.sum())


def test_salience_matches_reference(golden):
    from lazzaro_b200 import kernels
    s = torch.tensor(golden["dp_sal"], device="cuda")
    kernels.salience_decay(s, s.numel(), 0.01, 0.2)
    np.testing.assert_array_equal(s.cpu().numpy(), golden["dp_sal_after"].astype(np.float32))


@pytest.mark.parametrize("n", [0, 1, 7, 2047, 2048, 2049, 100_003, 1_000_000])
@pytest.mark.parametrize("thr", [0.5, 0.0, 2.0])
def test_decay_prune_sizes_and_extremes(n, thr):
    from lazzaro_b200 import synth
    src, dst, w, cooc = synth.edge_table(n, 1000, seed=n + 1)
    cooc = np.arange(n, dtype=np.int32)
    t = table(src, dst, w, cooc)
    pruned = t.decay_prune(0.99, thr)
    es, ed, ew, ec, ep = graph_ops.decay_prune_edges(src, dst, w, cooc, 0.99, thr)
    s, d, ww, c = t.to_numpy()
    assert pruned == ep and t.n == n - ep
    np.testing.assert_array_equal(s, es); np.testing.assert_array_equal(d, ed)
    np.testing.assert_array_equal(ww, ew); np.testing.assert_array_equal(c, ec)


@pytest.mark.parametrize("factor,thr", [(0.99, 0.5), (1.0, 0.5), (0.5, 0.3), (1e-3, 1e-4), (3.0, 0.7), (0.99, 0.0),
                                        (0.99, -0.25), (-1.0, 0.5), (0.0, 0.5), (0.99, 1e-40), (1.0 / 3.0, 1.0 / 3.0)])
@pytest.mark.parametrize("in_place", [False, True])
def test_prune_decision_on_the_boundary(factor, thr, in_place):
    """Weights packed into the few float32 values either side of the point where float64(w) * factor crosses the threshold
    (memory_shard.py:79-84 compares in float64): the kept set must be the oracle's, whatever form the kernels decide in."""
    from lazzaro_b200.graph import EdgeTable
    centre = np.float32(thr / factor) if factor != 0 else np.float32(thr)
    band = [centre]
    for _ in range(40):
        band.append(np.nextafter(band[-1], np.float32(np.inf), dtype=np.float32))
    lo = centre
    for _ in range(40):
        lo = np.nextafter(lo, np.float32(-np.inf), dtype=np.float32)
        band.append(lo)
    band = np.array(band + [0.0, -0.0, 1e-45, -1e-45, 1.0, -1.0, 3.4e38, -3.4e38], np.float32)
    rng = np.random.default_rng(5)
    n = 70_001
    w = band[rng.integers(0, band.size, n)]
    src = rng.integers(0, 1000, n).astype(np.int32); dst = rng.integers(0, 1000, n).astype(np.int32)
    cooc = np.arange(n, dtype=np.int32)
    t = EdgeTable(device="cuda", capacity=n, in_place=in_place)
    t.append(src, dst, w, cooc)
    pruned = t.decay_prune(factor, thr)
    es, ed, ew, ec, ep = graph_ops.decay_prune_edges(src, dst, w, cooc, factor, thr)
    s, d, ww, c = t.to_numpy()
    assert pruned == ep
    np.testing.assert_array_equal(c, ec); np.testing.assert_array_equal(s, es)
    np.testing.assert_array_equal(ww.view(np.int32), ew.view(np.int32))


def test_in_place_and_out_of_place_tables_agree():
    from lazzaro_b200 import synth
    from lazzaro_b200.graph import EdgeTable
    n = 777_777
    cols = synth.edge_table(n, 9000, seed=8)
    a = EdgeTable.from_arrays(*cols, device="cuda")                  # ping-pong columns: counts ahead of the data path
    b = EdgeTable(device="cuda", capacity=n, in_place=True)          # in place: counts published with the columns on chip
    b.append(*cols)
    for factor, thr in ((0.99, 0.3), (0.97, 0.3), (1.0, 0.5)):
        assert a.decay_prune(factor, thr) == b.decay_prune(factor, thr)
        for x, y in zip(a.to_numpy(), b.to_numpy()):
            np.testing.assert_array_equal(x, y)


def test_two_sweeps_on_two_streams_at_once():
    """Two tables swept concurrently on two streams, each with its own scratch: the streamed kernel's CTAs take their tiles
    from a ticket, so two grids sharing the SMs cannot wait on each other; both results equal the oracle's."""
    from lazzaro_b200 import kernels, synth, _lib
    lib = _lib.load()
    jobs = []
    for seed, n, thr in ((3, 4_000_037, 0.5), (4, 3_500_011, 0.2)):
        cols = synth.edge_table(n, 7000, seed=seed)
        ins = [torch.as_tensor(a).cuda() for a in cols]
        outs = [torch.empty_like(a) for a in ins]
        scratch = torch.empty(int(lib.lz_edge_decay_prune_scratch_bytes(n)), dtype=torch.uint8, device="cuda")
        kept = torch.zeros(1, dtype=torch.int64, device="cuda")
        jobs.append((cols, ins, outs, scratch, kept, n, thr, torch.cuda.Stream()))
    torch.cuda.synchronize()
    for _ in range(5):                                     # several rounds: the two grids interleave differently each time
        for cols, ins, outs, scratch, kept, n, thr, stream in jobs:
            with torch.cuda.stream(stream):
                kernels.edge_decay_prune(*ins, n, 0.99, thr, out=tuple(outs), scratch=scratch, kept=kept)
    torch.cuda.synchronize()
    for cols, ins, outs, scratch, kept, n, thr, stream in jobs:
        es, ed, ew, ec, ep = graph_ops.decay_prune_edges(*cols, 0.99, thr)
        k = int(kept.item())
        assert k == n - ep
        np.testing.assert_array_equal(outs[0][:k].cpu().numpy(), es)
        np.testing.assert_array_equal(outs[2][:k].cpu().numpy(), ew)
        np.testing.assert_array_equal(outs[3][:k].cpu().numpy(), ec)


def test_in_place_keeps_everything():
    """threshold below every weight, in place: every tile writes exactly onto its own input range, the case where a tile's
    output reaches furthest into its predecessors' columns."""
    from lazzaro_b200 import synth
    from lazzaro_b200.graph import EdgeTable
    n = 2_222_222
    cols = synth.edge_table(n, 9000, seed=12)
    t = EdgeTable(device="cuda", capacity=n, in_place=True)
    t.append(*cols)
    for _ in range(3):
        assert t.decay_prune(1.0, -1.0) == 0
    for x, y in zip(t.to_numpy(), cols):
        np.testing.assert_array_equal(x, np.asarray(y))


def test_prune_many_random_tables():
    """150 random tables (sizes around tile, wave and CTA-count multiples, every kept fraction, in place and out of place, a
    dirty scratch buffer) against the same arithmetic in torch float64 on the device: counts, order and bits."""
    from lazzaro_b200 import kernels, _lib
    lib = _lib.load()
    rng = np.random.default_rng(2024)
    specials = [2048 * 256, 2048 * 256 + 1, 2048 * 296, 2048 * 296 * 3 + 5, 2048 * 512 - 1, 2048 * 257]
    for case in range(150):
        n = int(specials[case % len(specials)] + rng.integers(-3, 4)) if case % 5 == 0 else int(rng.integers(1, 3_000_000))
        factor = float(rng.choice([0.99, 1.0, 0.5, 0.97]))
        thr = float(rng.choice([0.0, 0.1, 0.5, 0.9, 0.999, 2.0]))
        g = torch.Generator(device="cuda"); g.manual_seed(case)
        cols = [torch.randint(0, 1 << 20, (n,), generator=g, device="cuda", dtype=torch.int32),
                torch.randint(0, 1 << 20, (n,), generator=g, device="cuda", dtype=torch.int32),
                torch.rand(n, generator=g, device="cuda"), torch.arange(n, device="cuda", dtype=torch.int32)]
        w64 = cols[2].double() * factor
        keep = ~(w64 < thr)
        want = [cols[0][keep], cols[1][keep], w64.float()[keep], cols[3][keep]]
        in_place = case % 3 == 0
        outs = cols if in_place else [torch.full_like(c, -7) for c in cols]
        scratch = torch.randint(0, 255, (int(lib.lz_edge_decay_prune_scratch_bytes(n)),), device="cuda", dtype=torch.uint8)
        kept, _ = kernels.edge_decay_prune(*cols, n, factor, thr, out=None if in_place else tuple(outs), scratch=scratch)
        m = int(kept.item())
        assert m == int(keep.sum()), (case, n, factor, thr, in_place)
        for o, w in zip(outs, want):
            assert torch.equal(o[:m], w), (case, n, factor, thr, in_place)


def test_out_of_place_and_idempotence():
    from lazzaro_b200 import kernels, synth
    n = 300_000
    src, dst, w, cooc = synth.edge_table(n, 5000, seed=4)
    ins = [torch.as_tensor(a).cuda() for a in (src, dst, w, cooc)]
    outs = [torch.empty_like(a) for a in ins]
    kept, _ = kernels.edge_decay_prune(*ins, n, 0.99, 0.5, out=tuple(outs))
    k = int(kept.item())
    es, ed, ew, ec, ep = graph_ops.decay_prune_edges(src, dst, w, cooc, 0.99, 0.5)
    assert k == n - ep
    np.testing.assert_array_equal(outs[2][:k].cpu().numpy(), ew)
    np.testing.assert_array_equal(ins[2].cpu().numpy(), w)             # inputs untouched
    # pruning the survivors again with factor 1 removes nothing (idempotence)
    kept2, _ = kernels.edge_decay_prune(*[o[:k] for o in outs], k, 1.0, 0.5)
    assert int(kept2.item()) == k
    np.testing.assert_array_equal(outs[0][:k].cpu().numpy(), es)


def test_full_size_50m_edges_properties():
    """cfg5 size: 50M edges — sizes the oracle still handles in seconds with numpy."""
    from lazzaro_b200 import synth
    n = 50_000_000
    src, dst, w, cooc = synth.edge_table(n, 100_000_000 // 1000)
    t = table(src, dst, w, cooc)
    pruned = t.decay_prune(0.99, 0.5)
    es, ed, ew, ec, ep = graph_ops.decay_prune_edges(src, dst, w, cooc, 0.99, 0.5)
    assert pruned == ep
    s, d, ww, c = t.to_numpy()
    assert np.array_equal(s, es) and np.array_equal(d, ed) and np.array_equal(ww, ew)
    assert (ww >= np.float32(0.5) - 1e-7).all()


def test_memory_shard_mirror_runs_reference_scenario():
    """tests/test_profile_update.py:32-78 through our MemoryShard mirror: edges 0.8 survive a decay
    and a prune at 0.5; a fresh chain edge of 0.5 decays to 0.495 and is pruned."""
    from lazzaro_b200.graph import BufferGraph, Edge, MemoryShard, Node
    shard = MemoryShard("default", device="cuda")
    for i in range(4):
        shard.add_node(Node(id=f"seed_{i}", content=f"m{i}", salience=0.9))
    for i in range(3):
        shard.add_edge(Edge(source=f"seed_{i}", target=f"seed_{i+1}", weight=0.8))
    shard.add_edge(Edge(source="seed_0", target="seed_3", weight=0.5))
    shard.add_edge(Edge(source="seed_0", target="seed_1", weight=0.8))        # reinforce: 0.9, cooc 2
    bg = BufferGraph({"default": shard})
    bg.apply_temporal_decay(0.01)
    assert bg.prune_weak_edges(0.5) == 1
    edges = shard.edges
    assert set(edges) == {("seed_0", "seed_1"), ("seed_1", "seed_2"), ("seed_2", "seed_3")}
    assert edges[("seed_0", "seed_1")].co_occurrence == 2
    assert edges[("seed_0", "seed_1")].weight == pytest.approx(np.float32(np.float32(0.9) * 0.99), rel=1e-7)
    assert shard.nodes["seed_0"].salience == pytest.approx(0.893, abs=1e-6)   # 0.2 + 0.7*0.99
    assert shard.get_neighbors("seed_1") == ["seed_0", "seed_2"]
    assert shard.size() == (4, 3)


def _components_oracle(src, dst, n):
    parent = list(range(n))

    def find(v):
        while parent[v] != v:
            parent[v] = parent[parent[v]]
            v = parent[v]
        return v
    for u, v in zip(src.tolist(), dst.tolist()):
        ru, rv = find(u), find(v)
        if ru != rv:
            parent[max(ru, rv)] = min(ru, rv)
    return np.array([find(v) for v in range(n)], dtype=np.int32)


@pytest.mark.parametrize("n_nodes,n_edges", [(1, 0), (10, 3), (1000, 700), (50_000, 40_000), (200_000, 1_000_000)])
def test_connected_components_match_union_find(n_nodes, n_edges):
    from lazzaro_b200 import kernels
    rng = np.random.default_rng(n_nodes + n_edges)
    src = rng.integers(0, n_nodes, n_edges, dtype=np.int32)
    dst = rng.integers(0, n_nodes, n_edges, dtype=np.int32)
    if n_edges == 0:
        s = d = torch.zeros(1, dtype=torch.int32, device="cuda")
    else:
        s, d = torch.as_tensor(src).cuda(), torch.as_tensor(dst).cuda()
    got = kernels.connected_components(s, d, n_edges, n_nodes).cpu().numpy()
    np.testing.assert_array_equal(got, _components_oracle(src, dst, n_nodes))       # label = min index: deterministic


def test_shard_components_reference_scenario():
    # tests/test_profile_update.py:32-78: a chain of four memories is one component of size >= 3
    from lazzaro_b200.graph import Edge, MemoryShard, Node
    shard = MemoryShard("default", device="cuda")
    for i in range(6):
        shard.add_node(Node(id=f"seed_{i}", content=f"m{i}"))
    for i in range(3):
        shard.add_edge(Edge(source=f"seed_{i}", target=f"seed_{i+1}", weight=0.8))
    comps = shard.connected_components()
    assert sorted(sorted(c) for c in comps) == [["seed_0", "seed_1", "seed_2", "seed_3"], ["seed_4"], ["seed_5"]]
