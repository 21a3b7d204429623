"""pytest configuration: registers the `gpu` marker and exposes golden fixtures.

`-m "not gpu"`: oracle vs golden vectors, host logic, C-ABI symbol checks (no GPU needed).
`-m gpu`: parity of the CUDA path against the oracle and the golden vectors on a real B200.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


class Golden:
    """Dictionary view over tests/golden/golden_small.npz with 'case/field' keys."""

    def __init__(self, path):
        self._z = np.load(path, allow_pickle=False)
        self.keys = list(self._z.keys())

    def __call__(self, case, field):
        x = self._z[f"{case}/{field}"]
        return x[()] if x.ndim == 0 else x

    def has(self, case, field):
        return f"{case}/{field}" in self._z

    def cases(self, prefix):
        out = []
        for k in self.keys:
            c = k.rsplit("/", 1)[0]
            if c.startswith(prefix) and c not in out:
                out.append(c)
        return out


@pytest.fixture(scope="session")
def golden():
    return Golden(os.path.join(GOLDEN_DIR, "golden_small.npz"))


@pytest.fixture(scope="session")
def golden_c1():
    return np.load(os.path.join(GOLDEN_DIR, "golden_c1.npz"), allow_pickle=False)


@pytest.fixture(scope="session")
def gs():
    """Golden vectors of the callers after the path (ot.solve, unbalanced, divergence): golden_solvers.npz."""
    return Golden(os.path.join(GOLDEN_DIR, "golden_solvers.npz"))


@pytest.fixture(scope="session")
def gstab():
    """Golden vectors of sinkhorn_stabilized / sinkhorn_epsilon_scaling: golden_stabilized.npz."""
    return Golden(os.path.join(GOLDEN_DIR, "golden_stabilized.npz"))


@pytest.fixture(scope="session")
def gstabm():
    """Golden vectors of sinkhorn_stabilized with 2-D b: golden_stabilized_multi.npz."""
    return Golden(os.path.join(GOLDEN_DIR, "golden_stabilized_multi.npz"))


@pytest.fixture(scope="session")
def ggw():
    """Golden vectors of entropic_gromov_wasserstein: golden_gromov.npz."""
    return Golden(os.path.join(GOLDEN_DIR, "golden_gromov.npz"))


@pytest.fixture(scope="session")
def ggwb():
    """Golden vectors of ot.batch.solve_gromov_batch: golden_gromov_batch.npz."""
    return Golden(os.path.join(GOLDEN_DIR, "golden_gromov_batch.npz"))
