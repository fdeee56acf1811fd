"""CPU-only: the C-ABI library builds, loads and exports every symbol the header declares; the
product package has no CPU fallback and never imports the oracle."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


@pytest.fixture(scope="module")
def lib():
    from lazzaro_b200 import build, _lib
    build.build()
    return _lib.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "lazzaro_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lz_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from lazzaro_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/lazzaro_b200.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature in lazzaro_b200/_lib.py"
    assert set(_lib.SIGNATURES) == set(names)


def test_host_only_calls(lib):
    assert lib.lz_version() >= 100
    assert lib.lz_last_error() == b""
    assert lib.lz_search_exact_scratch_bytes(4, 10, 296) >= 4 * 296 * (10 * 12 + 4)
    assert lib.lz_search_exact_scratch_bytes(0, 10, 1) == 0
    assert lib.lz_edge_decay_prune_scratch_bytes(50_000_000) >= (50_000_000 // 2048 + 1) * 8


def test_sass_is_sm100_only():
    out = subprocess.run(["cuobjdump", "-lelf", os.path.join(ROOT, "lazzaro_b200", "liblazzaro_b200.so")],
                         capture_output=True, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out and "sm_80" not in out


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "lazzaro_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "ref_harness" not in src, f


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from lazzaro_b200 import _lib
    from lazzaro_b200.store import B200Store
    from lazzaro_b200.graph import EdgeTable
    with pytest.raises(_lib.LazzaroB200Error):
        B200Store(dim=8)
    with pytest.raises(_lib.LazzaroB200Error):
        EdgeTable()
    from lazzaro_b200 import kernels
    with pytest.raises(_lib.LazzaroB200Error):
        kernels.row_norms(torch.zeros(4, 8))


def test_bulk_id_formatting_matches_per_row_lookup():
    """B200Store._ids_of (the batch searches' id strings) against _id_of row by row: one adopted block, explicit
    id overrides, several blocks, short counts, rows outside every block."""
    import numpy as np
    from lazzaro_b200.store import B200Store
    st = B200Store.__new__(B200Store)                       # host-side bookkeeping only: no device needed
    ids = np.array([[3, 999, -1, 1000], [0, 1, 2, 7]], np.int64)
    for row_id, blocks in (({}, [(0, 1000, "node_", 5)]), ({1: "x", 999: "y"}, [(0, 1000, "node_", 5)]),
                           ({}, [(0, 10, "a", 0), (10, 2000, "b", 3)]), ({}, [])):
        st._row_id, st._blocks = dict(row_id), list(blocks)
        for cnt in (np.array([4, 4]), np.array([2, 0]), np.array([0, 3])):
            want = [[st._id_of(int(r)) for r in ids[b, :cnt[b]]] for b in range(2)]
            assert st._ids_of(ids, cnt) == want
