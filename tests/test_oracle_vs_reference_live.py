"""CPU: the oracle against the UNMODIFIED reference, live, on seeded random inputs beyond the committed fixtures.

Runs where `oracle/_ref` holds the reference built by oracle/build_ref.py (the build container, and any box the directory
travelled to); skipped otherwise.  Nothing here touches /root/reference.  fp64, fixed iteration counts and converging
runs; tolerance 1e-12 on values the two sides compute by the same arithmetic.
"""
import os
import sys
import warnings

import numpy as np
import pytest

from oracle import sinkhorn_oracle as orc
from oracle import solvers_oracle as sorc

_REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")


@pytest.fixture(scope="module")
def ref():
    if not os.path.isdir(os.path.join(_REF, "ot")):
        pytest.skip("oracle/_ref is not built here")
    sys.path.insert(0, _REF)
    try:
        import ot  # the reference
    except Exception as e:  # pragma: no cover - a broken build is a skip, not a failure of the oracle
        pytest.skip(f"oracle/_ref does not import: {e}")
    finally:
        sys.path.remove(_REF)
    return ot


def problem(seed, n, m, h=0):
    rng = np.random.RandomState(seed)
    xs, xt = rng.randn(n, 3), rng.randn(m, 3) + 0.5
    M = orc.sqeuclidean(xs, xt)
    M /= M.max()
    a = rng.rand(n) + 0.1
    b = rng.rand(m, h) + 0.1 if h else rng.rand(m) + 0.1
    return a / a.sum(), b / b.sum(0), M, xs, xt


def close(x, y, tol=1e-12):
    x, y = np.asarray(x, dtype=np.float64), np.asarray(y, dtype=np.float64)
    assert x.shape == y.shape
    np.testing.assert_allclose(x, y, rtol=tol, atol=tol * max(1.0, float(np.abs(y).max()) if y.size else 1.0))


SEEDS = [11, 12, 13]


@pytest.mark.parametrize("seed", SEEDS)
@pytest.mark.parametrize("h", [0, 3])
def test_knopp_and_log(ref, seed, h):
    a, b, M, _, _ = problem(seed, 37 + seed, 29 + 2 * seed, h)
    for kw in (dict(numItermax=55, stopThr=0.0), dict(numItermax=2000, stopThr=1e-9)):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            r0, l0 = ref.bregman.sinkhorn_knopp(a, b, M, 0.1, log=True, **kw)
            r1, l1 = orc.sinkhorn_knopp(a, b, M, 0.1, log=True, **kw)
        assert l0["niter"] == l1["niter"]
        close(r1, r0)
        close(l1["u"], l0["u"])
        close(l1["err"], l0["err"], 1e-9)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            r0, l0 = ref.bregman.sinkhorn_log(a, b, M, 0.1, log=True, **kw)
            r1, l1 = orc.sinkhorn_log(a, b, M, 0.1, log=True, **kw)
        close(r1, r0, 1e-11)
        close(l1["log_v"], l0["log_v"], 1e-11)


@pytest.mark.parametrize("seed", SEEDS)
@pytest.mark.parametrize("h", [0, 2])
@pytest.mark.parametrize("reg,tau", [(0.05, 1e3), (0.01, 10.0)])
def test_stabilized(ref, seed, h, reg, tau):
    a, b, M, _, _ = problem(seed, 30 + seed, 41 - seed, h)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r0, l0 = ref.bregman.sinkhorn_stabilized(a, b, M, reg, numItermax=120, tau=tau, stopThr=0.0, log=True)
        r1, l1 = sorc.sinkhorn_stabilized(a, b, M, reg, numItermax=120, tau=tau, stopThr=0.0, log=True)
    assert l0["n_iter"] == l1["n_iter"]
    close(r1, r0, 1e-11)
    close(l1["alpha"], l0["alpha"], 1e-11)
    close(l1["err"], l0["err"], 1e-8)


@pytest.mark.parametrize("seed", SEEDS[:2])
def test_epsilon_scaling(ref, seed):
    a, b, M, _, _ = problem(seed, 25, 31)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r0, l0 = ref.bregman.sinkhorn_epsilon_scaling(a, b, M, 0.02, numItermax=15, numInnerItermax=40, stopThr=0.0,
                                                      log=True)
        r1, l1 = sorc.sinkhorn_epsilon_scaling(a, b, M, 0.02, numItermax=15, numInnerItermax=40, stopThr=0.0, log=True)
    close(r1, r0, 1e-10)
    close(l1["alpha"], l0["alpha"], 1e-10)


@pytest.mark.parametrize("seed", SEEDS)
@pytest.mark.parametrize("h", [0, 3])
@pytest.mark.parametrize("reg_m", [1.0, (0.5, 2.0), (float("inf"), 1.0)])
def test_unbalanced(ref, seed, h, reg_m):
    a, b, M, _, _ = problem(seed, 33, 27 + seed, h)
    b = 1.3 * b
    for kw in (dict(numItermax=40, stopThr=0.0), dict(numItermax=3000, stopThr=1e-9)):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            r0, l0 = ref.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, 0.1, reg_m, log=True, **kw)
            r1, l1 = sorc.sinkhorn_knopp_unbalanced(a, b, M, 0.1, reg_m, log=True, **kw)
        assert len(l0["err"]) == len(l1["err"])
        close(r1, r0)
        close(l1["logu"], l0["logu"])
        close(l1["err"], l0["err"], 1e-9)
        if not h:
            close(l1["total_cost"], l0["total_cost"], 1e-11)


@pytest.mark.parametrize("seed", SEEDS[:2])
def test_empirical_lazy(ref, seed):
    _, _, _, xs, xt = problem(seed, 40, 33)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f0, g0, l0 = ref.bregman.empirical_sinkhorn(xs, xt, 0.5, numIterMax=60, stopThr=-1.0, isLazy=True, batchSize=16,
                                                    log=True)
        f1, g1, l1 = orc.empirical_sinkhorn_lazy(xs, xt, 0.5, numIterMax=60, stopThr=-1.0, batchSize=16, log=True)
    assert l0["niter"] == l1["niter"]
    close(f1, f0, 1e-11)
    close(g1, g0, 1e-11)


@pytest.mark.parametrize("seed", SEEDS[:2])
def test_batch_projections(ref, seed):
    rng = np.random.RandomState(seed)
    M = rng.rand(5, 12, 9)
    a = rng.rand(5, 12) + 0.1
    b = rng.rand(5, 9) + 0.1
    a /= a.sum(1, keepdims=True)
    b /= b.sum(1, keepdims=True)
    K = np.exp(-M / 0.1)
    r0 = ref.batch.bregman_projection_batch(K, a, b, max_iter=50, tol=0.0)
    r1 = orc.bregman_projection_batch(K, a, b, max_iter=50, tol=0.0)
    close(r1["T"], r0["T"])
    r0 = ref.batch.bregman_log_projection_batch(-M / 0.1, a, b, max_iter=50, tol=0.0)
    r1 = orc.bregman_log_projection_batch(-M / 0.1, a, b, max_iter=50, tol=0.0)
    close(r1["T"], r0["T"], 1e-11)
