"""Run the reference's own Python on the hot path, in the BUILD CONTAINER only.
TEST INFRASTRUCTURE — see oracle/__init__.py.

`/root/reference` does not exist on the GPU box, so nothing that runs there may import this
module; it is used by `tests/golden/make_golden.py` (to write fixtures) and by the `not gpu`
tests that are skipped when `/root/reference` is absent.

`import lazzaro` fails at core/vector_store.py:2 (`import lancedb`, third-party, not installed), so
a stub `lancedb` module whose `connect()` raises is registered first; the reference is then driven
through `MemorySystem(store=<injected Store>)` (memory_system.py:83,104-107) and never touches the
stub.  Everything else executed is the reference's unmodified code.
"""
from __future__ import annotations

import json
import os
import sys
import types

REFERENCE_SRC = "/root/reference/src"


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "lazzaro"))


def import_reference():
    """Returns the reference's `lazzaro` package (imported from /root/reference/src)."""
    if not available():
        raise RuntimeError("/root/reference is not present on this machine")
    if "lancedb" not in sys.modules:
        stub = types.ModuleType("lancedb")

        def connect(*a, **k):
            raise RuntimeError("lancedb is not installed; inject a Store via MemorySystem(store=...)")

        stub.connect = connect
        sys.modules["lancedb"] = stub
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import lazzaro  # noqa: F401
    return lazzaro


class ScriptedLLM:
    """LLMProvider (interfaces.py:3-31) that replays a queue of JSON fact lists."""

    def __init__(self):
        self.queue = []

    def push_memories(self, memories):
        self.queue.append(json.dumps({"memories": memories}))

    def completion(self, messages, response_format=None):
        return self.queue.pop(0) if self.queue else "{}"

    def completion_stream(self, messages, response_format=None):
        yield self.completion(messages, response_format)


class TableEmbedder:
    """EmbeddingProvider (interfaces.py:33-53) backed by a text -> vector table."""

    def __init__(self, table=None, dim=8):
        self.table = dict(table or {})
        self.dim = dim

    def embed(self, text):
        return list(self.table.get(text, [0.0] * self.dim))

    def batch_embed(self, texts):
        return [self.embed(t) for t in texts]


def make_memory_system(store, llm=None, embedder=None, **kw):
    """A reference `MemorySystem` wired for deterministic offline runs."""
    lz = import_reference()
    from lazzaro.core.memory_system import MemorySystem
    args = dict(llm_provider=llm or ScriptedLLM(), embedding_provider=embedder or TableEmbedder(),
                store=store, enable_async=False, load_from_disk=False, enable_caching=False)
    args.update(kw)
    return MemorySystem(**args)
