"""CPU oracle for the Lazzaro memory hot path — TEST INFRASTRUCTURE ONLY.

This package restates, in numpy, the arithmetic of the reference's hot path
(`/root/reference/src/lazzaro/core/{vector_store,memory_system,memory_shard}.py`) so that the CUDA
path in `lazzaro_b200/` can be checked against it.  Nothing in the product package
`lazzaro_b200/` imports it.  Only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` /
`--impl reference` legs of `bench.py` may import it, and there only as the checker / the CPU
number reported beside the GPU one — never as the thing shipped.

Parity status
-------------
* Python side of the path (cosine, dedup gate, linking, centroid mean, decay, prune): PINNED — the
  functions here are checked against outputs of the reference's own code, run in the build
  container under an injected store (`oracle/ref_harness.py`, fixtures in `tests/golden/`,
  generator `tests/golden/make_golden.py`), and against the expectations of the reference's own
  tests (`tests/test_consolidation_semantic.py:63-71`, `tests/test_profile_update.py:32-78`,
  `tests/test_vector_store.py:22-68`).
* LanceDB boundary (`vector_store.py:132-140`): the flat kNN arithmetic lives in the third-party
  `lancedb` package (unpinned in `pyproject.toml:24`, not vendored, not installable here), so
  metric, tie order and accumulation precision are **parity unpinned** beyond the three trivial
  pins the reference tests hold.  We pin them ourselves: exhaustive search, float32 rows, float64
  accumulation, order by (distance asc, insertion index asc), filter before top-k.
"""
