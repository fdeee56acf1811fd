import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_python_path.npz"))


@pytest.fixture(scope="session")
def golden_next():
    """Reference outputs for the SURVEY §8f rows (eviction, neighbours); tests/golden/make_golden.py."""
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_next_rows.npz"))
