import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    from oracle.oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference (oracle/_ref); skipped where it was not built / cannot run."""
    from oracle.oracle import Ref, ref_variant
    if ref_variant() is None:
        pytest.skip("oracle/_ref not built for this host CPU")
    return Ref()


@pytest.fixture(scope="session")
def golden():
    return np.load(ROOT / "tests" / "golden" / "ref_vectors.npz")


@pytest.fixture(scope="session")
def b200():
    """The product's C ABI through ctypes. Fails loudly (no fallback) if the extension is missing."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from __graft_entry__ import load_pkg
    return load_pkg()


def nmse(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.sum((a - b) ** 2) / max(np.sum(b ** 2), 1e-300))
