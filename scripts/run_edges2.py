"""Decay + prune: one form (LZ_PRUNE_FORM: 1 streamed, 2 chained, 3/unset two-pass) checked against a torch restatement on
several table sizes (ragged, tiny, in place), then timed.   python scripts/run_edges2.py [n_edges]"""
import os, sys
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
from lazzaro_b200 import kernels as K, _lib
_lib.LIB_PATH = os.environ.get('LZ_LIB', _lib.LIB_PATH)       # an older build for same-box A/B timing
lib = _lib.load()
dev = "cuda"

def make(n, seed=1):
    g = torch.Generator(device=dev); g.manual_seed(seed)
    return [torch.randint(0, 1 << 20, (n,), generator=g, device=dev, dtype=torch.int32) for _ in range(2)] + \
           [torch.rand(n, generator=g, device=dev), torch.randint(1, 9, (n,), generator=g, device=dev, dtype=torch.int32)]

def expect(cols, f, t):
    w64 = cols[2].double() * f
    keep = ~(w64 < t)
    return [cols[0][keep], cols[1][keep], w64.float()[keep], cols[3][keep]]

def check(n, f=0.99, t=0.5, inplace=False):
    cols = make(n, seed=n % 97 + 1)
    want = expect(cols, f, t)
    kept = torch.zeros(1, dtype=torch.int64, device=dev)
    scratch = torch.empty(int(lib.lz_edge_decay_prune_scratch_bytes(n)), dtype=torch.uint8, device=dev)
    outs = cols if inplace else [torch.full_like(c, -1) for c in cols]
    K.edge_decay_prune(*cols, n, f, t, out=None if inplace else tuple(outs), scratch=scratch, kept=kept)
    m = int(kept.item())
    ok = m == want[0].numel() and all(torch.equal(o[:m], w) for o, w in zip(outs, want))
    print(f"n={n} f={f} t={t} inplace={inplace}: kept {m} {'ok' if ok else 'MISMATCH'}", flush=True)
    return ok

good = True
for n in () if os.environ.get('LZ_SKIP_CHECK') else (1, 7, 2047, 2048, 2049, 5000, 1 << 16, (1 << 20) + 13, 3_000_001):
    good &= check(n)
    good &= check(n, inplace=True)
good &= bool(os.environ.get('LZ_SKIP_CHECK')) or check(1_000_000, t=0.0) and check(1_000_000, t=2.0) and check(1_000_000, t=0.9, inplace=True)
if not good:
    sys.exit(1)

def timed(fn, it=20):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

for n in ([int(a) for a in sys.argv[1:]] or [50_000_000, 6_250_000]):
    cols = make(n)
    outs = [torch.empty_like(c) for c in cols]
    kept = torch.zeros(1, dtype=torch.int64, device=dev)
    scratch = torch.empty(int(lib.lz_edge_decay_prune_scratch_bytes(n)), dtype=torch.uint8, device=dev)
    f = lambda: K.edge_decay_prune(*cols, n, 0.99, 0.5, out=tuple(outs), scratch=scratch, kept=kept)
    for _ in range(200 if n < 10_000_000 else 50): f()           # clocks settled
    ms = min(timed(f) for _ in range(3))
    kf = float(kept.item()) / n
    gb = (16 + 16 * kf) * n / 1e9
    print(f"form={os.environ.get('LZ_PRUNE_FORM','0')} thr={os.environ.get('LZ_PRUNE_THREADS','-')} st={os.environ.get('LZ_PRUNE_STAGES','-')} "
          f"ctas={os.environ.get('LZ_PRUNE_CTAS','-')} ahead={os.environ.get('LZ_PRUNE_AHEAD','-')} n={n}: {ms*1e3:.1f} us  {gb/ms*1e3:.0f} GB/s  kept {kf:.4f}", flush=True)
