"""GPU: the allocation-free decay + prune entry point (lz_edge_decay_prune_noalloc).

Some users of the library cannot allocate device memory at the moment the sweep runs — a service
that handed all of HBM to the row store at start-up (a late cudaMalloc may fail, and it would
synchronise the device under the searches in flight), or a bare C-ABI binding that has no device
allocator of its own.  For them the sweep's per-tile words live in an array inside the library.
This test drives that entry point through the C-ABI and requires bit-identical output to
lz_edge_decay_prune (our edge tests compare exactly, so this one does too), on two calls in a row:
the second call finds the first call's words still in the shared array and must not be disturbed
by them.  The file name makes pytest collect it last, so a `-x` run has already run everything else.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cols(n, seed):
    from lazzaro_b200 import synth
    src, dst, w, _ = synth.edge_table(n, 4096, seed=seed)
    cooc = np.arange(n, dtype=np.int32)
    return [torch.as_tensor(a).cuda() for a in (src, dst, w, cooc)]


def _with_scratch(cols, n, factor, thr):
    from lazzaro_b200 import kernels
    outs = [torch.empty_like(c) for c in cols]
    kept, _ = kernels.edge_decay_prune(*cols, n, factor, thr, out=tuple(outs))
    torch.cuda.synchronize()
    k = int(kept.item())
    return k, [o[:k].cpu().numpy() for o in outs]


def _no_alloc(cols, n, factor, thr):
    from lazzaro_b200 import _lib
    lib = _lib.load()
    outs = [torch.empty_like(c) for c in cols]          # the caller still owns inputs and outputs
    kept = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = lib.lz_edge_decay_prune_noalloc(p(cols[0]), p(cols[1]), p(cols[2]), p(cols[3]), n, factor, thr,
                                         p(outs[0]), p(outs[1]), p(outs[2]), p(outs[3]), p(kept),
                                         C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, "lz_edge_decay_prune_noalloc")
    torch.cuda.synchronize()
    k = int(kept.item())
    return k, [o[:k].cpu().numpy() for o in outs]


def test_no_alloc_matches_scratch_version_twice_in_a_row():
    # first a large table (many tiles published), then a smaller, different one: stale words from the
    # first call would wreck the second if they were not cleared
    for n, seed, factor, thr in [(1_300_000, 101, 0.99, 0.5), (70_001, 202, 0.97, 0.3)]:
        cols = _cols(n, seed)
        k_ref, ref = _with_scratch(cols, n, factor, thr)
        k_new, new = _no_alloc(cols, n, factor, thr)
        assert k_new == k_ref and 0 < k_ref < n
        for a, b in zip(new, ref):
            np.testing.assert_array_equal(a, b)


def test_no_alloc_rejects_oversize_and_handles_empty():
    from lazzaro_b200 import _lib
    lib = _lib.load()
    kept = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    dummy = torch.zeros(16, dtype=torch.int32, device="cuda")
    p = C.c_void_p(dummy.data_ptr())
    too_many = 67_108_864 + 1
    rc = lib.lz_edge_decay_prune_noalloc(p, p, p, p, too_many, 1.0, 0.5, p, p, p, p, C.c_void_p(kept.data_ptr()), None)
    assert rc == _lib.ERR_ARG
    msg = lib.lz_last_error().decode()
    assert "67108864" in msg and "lz_edge_decay_prune" in msg
    rc = lib.lz_edge_decay_prune_noalloc(None, None, None, None, 0, 1.0, 0.5, None, None, None, None,
                                         C.c_void_p(kept.data_ptr()), None)
    torch.cuda.synchronize()
    assert rc == 0 and int(kept.item()) == 0


def test_no_alloc_in_place_and_against_the_oracle():
    """The mode this entry point's single-pass kernel exists for: outputs aliasing the inputs (compaction in place, as a
    binding that owns one copy of the edge columns would call it) — checked directly against the numpy oracle
    (oracle/graph_ops.py decay_prune_edges, memory_shard.py:64-84), not only against the scratch-taking entry point."""
    from lazzaro_b200 import _lib
    from oracle import graph_ops
    lib = _lib.load()
    for n, seed, factor, thr in [(900_001, 303, 0.99, 0.5), (5_000, 404, 1.0, 0.25)]:
        cols = _cols(n, seed)
        host = [c.cpu().numpy() for c in cols]
        ws, wd, ww, wc, pruned = graph_ops.decay_prune_edges(*host, factor, thr)
        k_out, out = _no_alloc(cols, n, factor, thr)                        # out of place first
        work = [c.clone() for c in cols]
        kept = torch.full((1,), -7, dtype=torch.int64, device="cuda")
        p = lambda t: C.c_void_p(t.data_ptr())
        rc = lib.lz_edge_decay_prune_noalloc(p(work[0]), p(work[1]), p(work[2]), p(work[3]), n, factor, thr,
                                             p(work[0]), p(work[1]), p(work[2]), p(work[3]), p(kept),
                                             C.c_void_p(torch.cuda.current_stream().cuda_stream))
        _lib.check(rc, "lz_edge_decay_prune_noalloc (in place)")
        torch.cuda.synchronize()
        k = int(kept.item())
        assert k == k_out == n - pruned == ws.shape[0]
        for got, alias, want in zip(out, work, (ws, wd, ww, wc)):
            np.testing.assert_array_equal(got, want)
            np.testing.assert_array_equal(alias[:k].cpu().numpy(), want)
