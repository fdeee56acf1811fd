"""GPU, needs >= 2 devices: the sharded search over NCCL equals the single-GPU search (scripts/check_dist.py under
torchrun).  Skipped on a one-GPU box; the bookkeeping is covered on CPU by tests/test_dist_gloo.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_sharded_search_matches_single_gpu():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU visible")
    world = 2 if n < 4 else 4
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                          "--master-addr", "127.0.0.1", "--master-port", "29577", os.path.join(ROOT, "scripts", "check_dist.py")],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert res.stdout.count("OK") == 4 and "MISMATCH" not in res.stdout


def test_c_abi_exchange_matches_single_gpu():
    """lz_allgather_merge (ncclAllGather + merge behind the C-ABI, own communicator): scripts/check_cabi_exchange.py."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU visible")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29579", os.path.join(ROOT, "scripts", "check_cabi_exchange.py")],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert res.stdout.count("OK") == 3 and "MISMATCH" not in res.stdout
