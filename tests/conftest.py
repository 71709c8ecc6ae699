"""pytest configuration: `gpu` marker, in-tree native build once per session, shared helpers."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def native_libs():
    """Build (if stale) libbandcu.so, libband_c.so, libband_testing.so and the oracle."""
    from band_b200 import build as B
    return B.build_all()


@pytest.fixture(scope="session")
def hippo():
    z = np.load(os.path.join(GOLDEN, "hippo.npz"))
    return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def O():
    from oracle import oracle
    oracle.lib()
    return oracle
