"""ctypes binding of liblazzaro_b200.so (the C-ABI declared in include/lazzaro_b200.h).

The product path has no CPU fallback: if the shared library is missing, or a call returns a
non-zero status, this raises — loudly — rather than computing anything on the host.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblazzaro_b200.so")

OK, ERR_ARG, ERR_CUDA, ERR_SCRATCH, ERR_DEVICE = 0, -1, -2, -3, -4
COSINE, L2 = 0, 1
MAX_K, MAX_GROUP, BATCHED_MAX_K = 1024, 4, 128
METRICS = {"cosine": COSINE, "l2": L2}

_p, _i32, _i64, _f64, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_size_t

# name -> (restype, argtypes); one entry per symbol declared in include/lazzaro_b200.h
SIGNATURES = {
    "lz_version": (_i32, []),
    "lz_last_error": (C.c_char_p, []),
    "lz_device_info": (_i32, [C.POINTER(_i32)] * 3),
    "lz_row_norms": (_i32, [_p, _i64, _i32, _i64, _p, _p]),
    "lz_search_exact_scratch_bytes": (_sz, [_i32, _i32, _i32]),
    "lz_search_exact_default_chunks": (_i32, [_i64, _i32]),
    "lz_search_exact": (_i32, [_p, _i64, _i32, _p, _p, _i64, _i64, _p, _p, _p, _p, _i64, _i32, _i32,
                               _i32, _i32, _i32, _i64, _p, _sz, _p, _p, _p, _p]),
    "lz_rows_to_shadow": (_i32, [_p, _i64, _i32, _i64, _p, _i32, _p, _p, _p]),
    "lz_search_batched_scratch_bytes": (_sz, [_i64, _i32, _i32, _i32]),
    "lz_search_batched": (_i32, [_p, _p, _i64, _i32, _p, _p, _i32, _i64, _i64, _i64, _p, _i64, _i32, _i32, _f64,
                                 _i32, _f64, _f64, _i64, _p, _sz, _p, _p, _p, _p, _p]),
    "lz_search_one_scratch_bytes": (_sz, [_i32]),
    "lz_search_one": (_i32, [_p, _i64, _i32, _p, _p, _i32, _i64, _i64, _p, _i32, _i32, _i64, _p, _sz, _p, _i32, _p]),
    "lz_search_exact_ranges": (_i32, [_p, _i64, _i32, _p, _p, _p, _p, _i32, _i64, _p, _i64, _i32, _i32, _i32, _i32, _i64, _p, _sz,
                                      _p, _p, _p, _p]),
    "lz_search_two_level_scratch_bytes": (_sz, [_i32, _i32, _i32, _i32, _i32]),
    "lz_search_two_level": (_i32, [_p, _p, _i64, _i32, _p, _i64, _p, _i32, _p, _p, _i64, _p, _f64, _p, _i64, _i32, _i32, _i32,
                                   _f64, _i32, _f64, _f64, _i64, _p, _sz, _p, _p, _p, _p, _p, _p]),
    "lz_assign_nearest_scratch_bytes": (_sz, [_i32, _i64, _i32, _i32]),
    "lz_assign_nearest": (_i32, [_p, _p, _i64, _i32, _p, _i32, _f64, _p, _i64, _i64, _p, _p, _f64, _p, _sz, _p, _p, _p, _p, _p]),
    "lz_merge_topk": (_i32, [_p, _p, _i32, _i32, _i32, _i32, _p, _p, _p, _p]),
    "lz_merge_topk_strided": (_i32, [_p, _p, _i64, _i32, _i32, _i32, _i32, _p, _p, _p, _p]),
    "lz_allgather_merge": (_i32, [_p, _i32, _p, _p, _i32, _i32, _i32, _p, _p, _p, _p]),
    "lz_score_pairs": (_i32, [_p, _i64, _i32, _p, _p, _i64, _p, _p, _i64, _i32, _p, _p]),
    "lz_group_by_label_scratch_bytes": (_sz, [_i64, _i32]),
    "lz_group_by_label": (_i32, [_p, _i64, _i32, _p, _sz, _p, _p, _p]),
    "lz_segment_mean": (_i32, [_p, _i64, _i32, _p, _p, _i32, _p, _p]),
    "lz_edge_decay_prune_scratch_bytes": (_sz, [_i64]),
    "lz_edge_decay_prune": (_i32, [_p, _p, _p, _p, _i64, _f64, _f64, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "lz_edge_decay_prune_noalloc": (_i32, [_p, _p, _p, _p, _i64, _f64, _f64, _p, _p, _p, _p, _p, _p]),
    "lz_edge_decay": (_i32, [_p, _i64, _f64, _p]),
    "lz_connected_components": (_i32, [_p, _p, _i64, _i32, _p, _p]),
    "lz_reach_components_scratch_bytes": (_sz, [_i64, _i32]),
    "lz_reach_components": (_i32, [_p, _p, _p, _i64, _i32, _p, _sz, _p, _p]),
    "lz_edge_endpoints_to_global": (_i32, [_p, _p, _p, _i64, _p, _p, _i32, _f64, _p, _p, _p, _p]),
    "lz_salience_decay": (_i32, [_p, _i64, _f64, _f64, _p]),
    "lz_evict_lowest_scratch_bytes": (_sz, [_i64, _i32]),
    "lz_evict_lowest": (_i32, [_p, _i32, _p, _p, _p, _i64, _f64, _i32, _p, _sz, _p, _p, _p, _p, _p]),
    "lz_csr_build_scratch_bytes": (_sz, [_i64, _i32]),
    "lz_csr_build": (_i32, [_p, _p, _i64, _i32, _p, _sz, _p, _p, _p]),
    "lz_neighbors": (_i32, [_p, _p, _p, _p, _p, _i32, _p, _i32, _f64, _i32, _p, _p, _p, _p]),
    "lz_edge_find": (_i32, [_p, _p, _p, _i32, _p, _p, _i32, _p, _p]),
    "lz_edge_reinforce": (_i32, [_p, _p, _i64, _p, _p, _i64, _p]),
}


class LazzaroB200Error(RuntimeError):
    pass


_lib = None


def load():
    """Loads the library (once). Raises if it has not been built — there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LazzaroB200Error(
            f"{LIB_PATH} not found: build it with `python -m lazzaro_b200.build` "
            "(or __graft_entry__.build()). lazzaro_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header and library out of sync
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().lz_last_error().decode("utf-8", "replace")
        raise LazzaroB200Error(f"{what or 'lazzaro_b200 call'} failed (status {rc}): {msg}")


def ptr(t):
    """Device pointer of a torch tensor (or None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise LazzaroB200Error("no CUDA device: lazzaro_b200 runs its hot path on a B200 only (no CPU fallback)")
