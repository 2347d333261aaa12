import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")
    config.addinivalue_line("markers", "slow: tens of seconds of reference (CPU oracle) time; still part of `-m gpu`")


@pytest.fixture(scope="session")
def golden():
    """Vectors produced by the unmodified reference (tests/golden/make_golden.py)."""
    path = os.path.join(ROOT, "tests", "golden", "slq_golden.npz")
    return dict(np.load(path))


@pytest.fixture(scope="session")
def oracle():
    """CPU oracle (checker only).  Builds the C restatement on demand (gcc, ~1 s)."""
    from oracle import build_oracle, slq_oracle

    build_oracle.build_c_oracle(verbose=False)
    build_oracle.build_reference_extension(verbose=False)
    return slq_oracle


def laplacian_2d(g: int, dtype=np.float64):
    import scipy.sparse as sp

    T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g), dtype=dtype)
    I = sp.identity(g, dtype=dtype)  # noqa: E741
    return (sp.kron(I, T) + sp.kron(T, I)).tocsr().astype(dtype)


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
