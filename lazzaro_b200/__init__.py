"""lazzaro_b200 — B200-native implementation of the Lazzaro memory hot path.

Host code is Python and mirrors the reference's own interfaces for the path (the `Store` protocol,
the legacy vector-store calls, the MemoryShard decay/prune entry points); all arithmetic runs in
hand-written sm_100a CUDA kernels behind the C-ABI of include/lazzaro_b200.h.  There is no CPU
fallback: importing works anywhere, but every compute call needs the built library and a B200.
"""
__version__ = "0.1.0"

__all__ = ["B200Store", "EdgeTable", "MemoryShard", "BufferGraph", "Node", "Edge"]


def __getattr__(name):
    if name in ("B200Store", "LanceDBVectorStore"):
        from . import store
        return getattr(store, name)
    if name in ("EdgeTable", "MemoryShard", "BufferGraph", "Node", "Edge"):
        from . import graph
        return getattr(graph, name)
    raise AttributeError(name)
