"""CPU: the host logic of the joint 2-D b solves (pot_b200._dense.multi_hist_steps / multi_hist_select: probe step, chunked
enqueue, which buffer holds the result) driven by a numpy model of the device state machine
(oracle/multi_hist_protocol.py), against the oracle's sinkhorn_knopp / sinkhorn_knopp_unbalanced with a 2-D b."""
import warnings

import numpy as np
import pytest

from oracle import multi_hist_protocol as proto
from oracle import sinkhorn_oracle as orc
from oracle import solvers_oracle as sorc
from pot_b200 import _lib
from pot_b200._dense import multi_hist_select, multi_hist_steps


def problem(n, m, h, seed=0, mass=1.0):
    rng = np.random.RandomState(seed + n + m + h)
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = orc.sqeuclidean(xs, xt)
    M /= M.max()
    a = rng.rand(n) + 0.1
    b = rng.rand(m, h) + 0.1
    return a / a.sum(), mass * b / b.sum(0), M


def test_steps_and_select():
    assert multi_hist_steps(True, 45, 10) == 45 and multi_hist_steps(True, 41, 10) == 42   # probe step
    assert multi_hist_steps(True, 1, 10) == 2 and multi_hist_steps(False, 41, 10) == 41
    assert multi_hist_select(0, 0, 40) == (0, 39, _lib.ST_MAXITER)
    assert multi_hist_select(1, 20, 40) == (1, 20, _lib.ST_CONVERGED)
    assert multi_hist_select(2, 7, 40) == (1, 7, _lib.ST_NUMERR)
    assert multi_hist_select(2, 41, 41) == (1, 40, _lib.ST_MAXITER)  # a guard raised by the probe step is not the solve's
    assert multi_hist_select(1, 40, 41) == (1, 40, _lib.ST_CONVERGED)


@pytest.mark.parametrize("n,m,h,reg,its,thr,poll", [
    (60, 50, 3, 0.1, 45, 0.0, 50), (60, 50, 3, 0.1, 41, 0.0, 7), (60, 50, 3, 0.1, 1, 0.0, 50),
    (130, 77, 5, 0.05, 3000, 1e-9, 50), (33, 100, 9, 0.1, 1000, 1e-7, 3), (40, 52, 1, 0.1, 21, 1e-8, 50),
])
def test_knopp_protocol_matches_oracle(n, m, h, reg, its, thr, poll):
    a, b, M = problem(n, m, h)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        _, lr = orc.sinkhorn_knopp(a, b, M, reg, numItermax=its, stopThr=thr, log=True)
    K = np.exp(M / -reg)
    u0, v0 = np.full((n, h), 1.0 / n), np.full((m, h), 1.0 / m)
    u, v, errs, niter, status = proto.solve(proto.KNOPP, K, a, b, u0, v0, its, thr, 10, (1.0, 1.0), poll,
                                            multi_hist_steps, multi_hist_select)
    assert niter == lr["niter"]
    assert len(errs) == len(lr["err"])
    np.testing.assert_allclose(errs, lr["err"], rtol=1e-6, atol=1e-15)
    np.testing.assert_allclose(u, lr["u"], rtol=1e-11)
    np.testing.assert_allclose(v, lr["v"], rtol=1e-11)
    assert status == (_lib.ST_CONVERGED if niter < its - 1 or (errs and errs[-1] < thr) else _lib.ST_MAXITER)


def test_knopp_protocol_guard():
    a, b, M = problem(40, 30, 3, seed=5)
    M[:, :4] = 1e6
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        _, lr = orc.sinkhorn_knopp(a, b, M, 1.0, numItermax=100, stopThr=1e-9, log=True)
        K = np.exp(M / -1.0)
        u0, v0 = np.full((40, 3), 1.0 / 40), np.full((30, 3), 1.0 / 30)
        u, v, errs, niter, status = proto.solve(proto.KNOPP, K, a, b, u0, v0, 100, 1e-9, 10, (1.0, 1.0), 50,
                                                multi_hist_steps, multi_hist_select)
    assert status == _lib.ST_NUMERR and niter == lr["niter"] == 0 and len(errs) == len(lr["err"]) == 0
    np.testing.assert_array_equal(u, lr["u"])


@pytest.mark.parametrize("n,m,h,reg,reg_m,its,thr,poll", [
    (60, 50, 3, 0.1, (1.0, 1.0), 40, 0.0, 50), (130, 77, 5, 0.05, (0.5, 2.0), 3000, 1e-10, 50),
    (33, 100, 9, 0.1, (float("inf"), 1.0), 2000, 1e-9, 7), (40, 52, 1, 0.1, (1.0, 1.0), 500, 1e-9, 1),
])
def test_unbalanced_protocol_matches_oracle(n, m, h, reg, reg_m, its, thr, poll):
    a, b, M = problem(n, m, h, mass=1.3)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        _, lr = sorc.sinkhorn_knopp_unbalanced(a, b, M, reg, reg_m, numItermax=its, stopThr=thr, log=True)
    fi = tuple(rm / (rm + reg) if rm != float("inf") else 1.0 for rm in reg_m)
    K = np.exp(-M / reg)
    u, v, errs, niter, status = proto.solve(proto.UNBALANCED, K, a, b, np.ones((n, h)), np.ones((m, h)), its, thr, 10, fi,
                                            poll, multi_hist_steps, multi_hist_select)
    assert len(errs) == len(lr["err"])
    np.testing.assert_allclose(errs, lr["err"], rtol=1e-6, atol=1e-15)
    np.testing.assert_allclose(np.log(u + 1e-300), lr["logu"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(np.log(v + 1e-300), lr["logv"], rtol=1e-10, atol=1e-12)
    assert status == (_lib.ST_CONVERGED if errs[-1] < thr else _lib.ST_MAXITER)


def test_unbalanced_protocol_guard():
    a, b, M = problem(40, 30, 3, seed=7, mass=0.7)
    M[:4] *= 1e7
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        _, lr = sorc.sinkhorn_knopp_unbalanced(a, b, M, 0.01, 1.0, numItermax=30, stopThr=0.0, log=True)
        K = np.exp(-M / 0.01)
        u, v, errs, niter, status = proto.solve(proto.UNBALANCED, K, a, b, np.ones((40, 3)), np.ones((30, 3)), 30, 0.0,
                                                10, (1.0 / 1.01, 1.0 / 1.01), 50, multi_hist_steps, multi_hist_select)
    assert status == _lib.ST_NUMERR and len(errs) == len(lr["err"]) == 0
    # back at the initial scalings (the reference's (dim_a, 1) column of ones broadcasts over the histograms)
    np.testing.assert_array_equal(np.log(u + 1e-300), np.broadcast_to(lr["logu"], u.shape))
