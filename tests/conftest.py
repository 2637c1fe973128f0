"""pytest configuration: the `gpu` marker splits the suite.

  -m "not gpu"  runs in the build container (no GPU): oracle vs the reference-generated golden
                fixtures, host logic, C-ABI symbol export, gloo world_size-2 sharding.
  -m gpu        runs on a B200: parity of the CUDA path (through the C ABI) against the oracle.
Only this directory (plus __graft_entry__.smoke and bench.py's CPU legs) may import `oracle`.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name + ".npz"))
    return load


@pytest.fixture(scope="session")
def engine():
    from unast_b200.engine import get_engine
    return get_engine()


def snr_db(ref, x):
    ref = np.asarray(ref, np.float64)
    x = np.asarray(x, np.float64)
    assert ref.shape == x.shape, (ref.shape, x.shape)
    err = ref - x
    return float(10 * np.log10((ref ** 2).sum() / max((err ** 2).sum(), 1e-300)))
