"""Pins the CPU oracle (oracle/sinkhorn_oracle.py) to outputs of the unmodified reference.

Golden vectors were produced by tests/golden/make_golden.py importing PythonOT/POT from
/root/reference (numpy backend).  The oracle follows the reference's operation order,
so agreement is at round-off level (far tighter than the north_star tolerances).
"""
import warnings

import numpy as np
import pytest

from oracle import sinkhorn_oracle as orc

DENSE = [f"dense_{t}_{s}" for t in ("f64", "f32") for s in ("50x60", "33x47", "128x96")]
EMP = [f"emp_{t}_{s}" for t in ("f64", "f32") for s in ("40x50x3", "64x32x8", "37x41x2")]


def tol(case):
    return dict(rtol=1e-12, atol=1e-300) if "f64" in case else dict(rtol=2e-5, atol=1e-30)


@pytest.mark.parametrize("case", DENSE)
def test_knopp_fixed_conv_warm(golden, case):
    a, b, M = golden(case, "a"), golden(case, "b"), golden(case, "M")
    reg, its = float(golden(case, "reg")), int(golden(case, "its"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = orc.sinkhorn_knopp(a, b, M, reg, numItermax=its, stopThr=0, log=True)
    assert P.dtype == M.dtype
    np.testing.assert_allclose(P, golden(case + "/knopp_fixed", "plan"), **tol(case))
    np.testing.assert_allclose(lg["u"], golden(case + "/knopp_fixed", "u"), **tol(case))
    np.testing.assert_allclose(lg["v"], golden(case + "/knopp_fixed", "v"), **tol(case))
    np.testing.assert_allclose(lg["err"], golden(case + "/knopp_fixed", "err"), rtol=1e-4, atol=1e-12)
    assert lg["niter"] == int(golden(case + "/knopp_fixed", "niter")) == its - 1
    thr = float(golden(case + "/knopp_conv", "thr"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = orc.sinkhorn_knopp(a, b, M, reg, numItermax=2000, stopThr=thr, log=True)
    assert lg["niter"] == int(golden(case + "/knopp_conv", "niter"))
    np.testing.assert_allclose(P, golden(case + "/knopp_conv", "plan"), **tol(case))
    ws = (golden(case + "/knopp_warm", "ws_u"), golden(case + "/knopp_warm", "ws_v"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = orc.sinkhorn_knopp(a, b, M, reg, numItermax=20, stopThr=0, log=True, warmstart=ws)
    np.testing.assert_allclose(P, golden(case + "/knopp_warm", "plan"), **tol(case))


@pytest.mark.parametrize("case", DENSE)
def test_log_fixed_conv_warm(golden, case):
    a, b, M = golden(case, "a"), golden(case, "b"), golden(case, "M")
    reg, its = float(golden(case, "reg")), int(golden(case, "its"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = orc.sinkhorn_log(a, b, M, reg, numItermax=its, stopThr=0, log=True)
    assert P.dtype == M.dtype
    np.testing.assert_allclose(P, golden(case + "/log_fixed", "plan"), **tol(case))
    np.testing.assert_allclose(lg["log_u"], golden(case + "/log_fixed", "log_u"), **tol(case))
    np.testing.assert_allclose(lg["log_v"], golden(case + "/log_fixed", "log_v"), **tol(case))
    assert lg["niter"] == its - 1
    thr = float(golden(case + "/log_conv", "thr"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = orc.sinkhorn_log(a, b, M, reg, numItermax=2000, stopThr=thr, log=True)
    assert lg["niter"] == int(golden(case + "/log_conv", "niter"))
    np.testing.assert_allclose(P, golden(case + "/log_conv", "plan"), **tol(case))
    ws = (golden(case + "/knopp_warm", "ws_u"), golden(case + "/knopp_warm", "ws_v"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = orc.sinkhorn_log(a, b, M, reg, numItermax=20, stopThr=0, log=True, warmstart=ws)
    np.testing.assert_allclose(P, golden(case + "/log_warm", "plan"), **tol(case))


@pytest.mark.parametrize("case", DENSE)
def test_sinkhorn2_loss(golden, case):
    a, b, M = golden(case, "a"), golden(case, "b"), golden(case, "M")
    reg, its = float(golden(case, "reg")), int(golden(case, "its"))
    t = dict(rtol=1e-12) if "f64" in case else dict(rtol=1e-5)
    np.testing.assert_allclose(orc.sinkhorn2(a, b, M, reg, numItermax=its, stopThr=0),
                               golden(case + "/loss", "knopp"), **t)
    np.testing.assert_allclose(orc.sinkhorn2(a, b, M, reg, method="sinkhorn_log", numItermax=its, stopThr=0),
                               golden(case + "/loss", "log"), **t)


def test_multi_histogram(golden):
    a, b, M = golden("multib", "a"), golden("multib", "b"), golden("multib", "M")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        L, lg = orc.sinkhorn_knopp(a, b, M, 0.1, numItermax=100, stopThr=0, log=True)
        L2, lg2 = orc.sinkhorn_log(a, b, M, 0.1, numItermax=100, stopThr=0, log=True)
    np.testing.assert_allclose(L, golden("multib", "knopp_loss"), rtol=1e-12)
    np.testing.assert_allclose(lg["u"], golden("multib", "knopp_u"), rtol=1e-12)
    np.testing.assert_allclose(L2, golden("multib", "log_loss"), rtol=1e-12)
    np.testing.assert_allclose(lg2["log_v"], golden("multib", "log_v"), rtol=1e-12)


def test_numerical_guard(golden):
    a, b, M = golden("guard", "a"), golden("guard", "b"), golden("guard", "M")
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        P, lg = orc.sinkhorn_knopp(a, b, M, 1e-3, numItermax=50, stopThr=0, log=True)
    assert int(golden("guard", "warned")) == int(any("numerical errors" in str(x.message) for x in w))
    assert lg["niter"] == int(golden("guard", "niter"))
    np.testing.assert_allclose(P, golden("guard", "plan"), rtol=1e-12, atol=0)
    np.testing.assert_allclose(lg["u"], golden("guard", "u"), rtol=1e-12)


def test_empty_weights(golden):
    M = golden("empty", "M")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P = orc.sinkhorn([], [], M, 0.2, numItermax=30, stopThr=0)
    np.testing.assert_allclose(P, golden("empty", "plan"), rtol=1e-12)


def test_unknown_method():
    with pytest.raises(ValueError, match="Unknown method"):
        orc.sinkhorn([0.5, 0.5], [0.5, 0.5], np.eye(2), 1.0, method="nope")


@pytest.mark.parametrize("case", EMP)
def test_empirical_lazy(golden, case):
    xs, xt, a, b = (golden(case, k) for k in ("xs", "xt", "a", "b"))
    reg, its = float(golden(case, "reg")), int(golden(case, "its"))
    bs = tuple(int(x) for x in golden(case, "bs"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g, lg = orc.empirical_sinkhorn_lazy(xs, xt, reg, a, b, numIterMax=its, stopThr=-1, batchSize=bs, log=True)
    t = dict(rtol=1e-11, atol=1e-13) if "f64" in case else dict(rtol=1e-4, atol=1e-5)
    assert f.dtype == xs.dtype
    np.testing.assert_allclose(f, golden(case, "f"), **t)
    np.testing.assert_allclose(g, golden(case, "g"), **t)
    assert lg["niter"] == int(golden(case, "niter")) == its - 1
    np.testing.assert_allclose(lg["err"], golden(case, "err"), rtol=1e-3, atol=1e-6)
    np.testing.assert_allclose(orc.lazy_plan_block(xs, xt, f, g, reg), golden(case, "plan"),
                               rtol=1e-9 if "f64" in case else 1e-3, atol=1e-12)
    np.testing.assert_allclose(orc.lazy_plan_block(xs, xt, f, g, reg, slice(3, 11), slice(5, 20)),
                               golden(case, "plan_block"), rtol=1e-9 if "f64" in case else 1e-3, atol=1e-12)
    thr = float(golden(case + "/conv", "thr"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g, lg = orc.empirical_sinkhorn_lazy(xs, xt, reg, a, b, numIterMax=3000, stopThr=thr, batchSize=bs, log=True)
    assert lg["niter"] == int(golden(case + "/conv", "niter"))
    np.testing.assert_allclose(f, golden(case + "/conv", "f"), **t)


def test_empirical_default_weights(golden):
    xs, xt = golden("emp_default", "xs"), golden("emp_default", "xt")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g = orc.empirical_sinkhorn_lazy(xs, xt, 0.7, numIterMax=50, stopThr=-1, batchSize=7)
        P = orc.sinkhorn([], [], orc.sqeuclidean(xs, xt), 0.7, numItermax=50, stopThr=0)
    np.testing.assert_allclose(f, golden("emp_default", "f"), rtol=1e-11)
    np.testing.assert_allclose(g, golden("emp_default", "g"), rtol=1e-11)
    np.testing.assert_allclose(P, golden("emp_default", "plan"), rtol=1e-11)


@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("solver", ["log_sinkhorn", "sinkhorn"])
@pytest.mark.parametrize("reg_type", ["entropy", "kl"])
def test_batch(golden, tag, solver, reg_type):
    base = f"batch_{tag}"
    M, a, b = golden(base, "M"), golden(base, "a"), golden(base, "b")
    X, Y = golden(base, "X"), golden(base, "Y")
    Mb = orc.dist_sqeuclidean_batch(X, Y)
    np.testing.assert_allclose((Mb / Mb.max()).astype(M.dtype), M, rtol=1e-6, atol=1e-7)
    t = dict(rtol=1e-11, atol=1e-14) if tag == "f64" else dict(rtol=2e-4, atol=1e-7)
    r = orc.solve_batch(M, 0.1, a, b, max_iter=40, tol=0, solver=solver, reg_type=reg_type)
    c = f"{base}/{solver}_{reg_type}_fixed"
    assert r["n_iter"] == int(golden(c, "n_iter")) == 39
    np.testing.assert_allclose(r["plan"], golden(c, "plan"), **t)
    np.testing.assert_allclose(r["value"], golden(c, "value"), **t)
    np.testing.assert_allclose(r["value_linear"], golden(c, "value_linear"), **t)
    np.testing.assert_allclose(r["potentials"][0], golden(c, "pa"), **t)
    np.testing.assert_allclose(r["potentials"][1], golden(c, "pb"), **t)
    r = orc.solve_batch(M, 0.1, a, b, max_iter=1000, tol=1e-5, solver=solver, reg_type=reg_type)
    c = f"{base}/{solver}_{reg_type}_conv"
    assert r["n_iter"] == int(golden(c, "n_iter"))
    np.testing.assert_allclose(r["plan"], golden(c, "plan"), **t)


def test_batch_uniform(golden):
    M = golden("batch_f64", "M")
    r = orc.solve_batch(M, 0.1, max_iter=25, tol=0)
    np.testing.assert_allclose(r["plan"], golden("batch_f64/uniform", "plan"), rtol=1e-11)


def test_doctest_known_answers(golden):
    a, b, M = [0.5, 0.5], [0.5, 0.5], [[0.0, 1.0], [1.0, 0.0]]
    want = np.array([[0.36552929, 0.13447071], [0.13447071, 0.36552929]])  # _sinkhorn.py:127-129
    for fn in (orc.sinkhorn, orc.sinkhorn_knopp, orc.sinkhorn_log):
        np.testing.assert_allclose(fn(a, b, np.asarray(M), 1), want, atol=1e-8)
    np.testing.assert_allclose(orc.sinkhorn(a, b, np.asarray(M), 1), golden("doctest", "sinkhorn_2x2"), rtol=1e-12)
    np.testing.assert_allclose(orc.sinkhorn2(a, b, np.asarray(M), 1), 0.26894142136999516, rtol=1e-12)  # :346-347
    xs = np.arange(2, dtype=np.float64).reshape(2, 1)
    P = orc.sinkhorn([], [], orc.sqeuclidean(xs, xs.copy()), 0.1)
    np.testing.assert_allclose(P, [[4.99977301e-01, 2.26989344e-05], [2.26989344e-05, 4.99977301e-01]], rtol=1e-6)
    np.testing.assert_allclose(P, golden("doctest", "emp_2x2"), rtol=1e-12)  # _empirical.py:165-167


def test_config1_full(golden_c1):
    """BASELINE config 1 exactly (n=m=1000, fp64, reg=0.05, 1000 fixed iterations)."""
    rng = np.random.RandomState(0)
    x = rng.randn(1000, 2)
    y = rng.randn(1000, 2) + 1
    M = orc.sqeuclidean(x, y)
    np.testing.assert_allclose(M.sum(), golden_c1["M_checksum"], rtol=1e-13)
    a = np.ones(1000) / 1000
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = orc.sinkhorn_knopp(a, a.copy(), M, 0.05, numItermax=1000, stopThr=0, log=True)
    assert lg["niter"] == 999
    np.testing.assert_allclose(lg["u"], golden_c1["u"], rtol=1e-11)
    np.testing.assert_allclose(lg["v"], golden_c1["v"], rtol=1e-11)
    np.testing.assert_allclose(P[0], golden_c1["plan_row0"], rtol=1e-11, atol=1e-300)
    np.testing.assert_allclose(np.sum(P * M), golden_c1["loss"], rtol=1e-11)
    np.testing.assert_allclose(lg["err"], golden_c1["err"], rtol=1e-6, atol=1e-18)
