"""The reference's own tests for the hot path, run on the CUDA path (SURVEY section 8c lists them: test/test_bregman.py
:21-39, :130-160, :228-279, :282-314, :482-556, :1116-1158, :1197-1277, :1280-1312, :1450-1548).

Same inputs, same calls, same tolerances as the reference's tests; `ot` is pot_b200, the "numpy backend" side of the
reference's comparisons is the CPU oracle.  Left out on purpose: the methods outside the accelerated path (greenkhorn,
screenkhorn, barycenters) and the exact-LP duals the warm-start tests start from (any warm start must lead to the same
plan: a short log-domain solve provides one).
"""
import warnings
from itertools import product

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot
from oracle import sinkhorn_oracle as orc

DEV = "cuda:0"


def dev(*xs):
    out = [torch.as_tensor(np.ascontiguousarray(x), device=DEV) for x in xs]
    return out[0] if len(out) == 1 else out


def to_np(x):
    return x.detach().cpu().numpy()


def unif(n):
    return np.ones(n) / n


def gauss(n, m, s):  # ot.datasets.make_1D_gauss
    x = np.arange(n, dtype=np.float64)
    h = np.exp(-((x - m) ** 2) / (2 * s**2))
    return h / h.sum()


def dist0(n):  # ot.utils.dist0
    x = np.arange(n, dtype=np.float64).reshape((n, 1))
    return orc.sqeuclidean(x, x)


@pytest.mark.parametrize("verbose, warn", list(product([True, False], [True, False])))
def test_sinkhorn(verbose, warn):  # test/test_bregman.py:21-39
    n = 100
    rng = np.random.RandomState(0)
    x = rng.randn(n, 2)
    u = unif(n)
    M = orc.sqeuclidean(x, x)
    ud, Md = dev(u, M)
    G = to_np(ot.sinkhorn(ud, ud, Md, 1, stopThr=1e-10, verbose=verbose, warn=warn))
    np.testing.assert_allclose(u, G.sum(1), atol=1e-05)
    np.testing.assert_allclose(u, G.sum(0), atol=1e-05)
    with pytest.warns(UserWarning):
        ot.sinkhorn(ud, ud, Md, 1, stopThr=0, numItermax=1)


@pytest.mark.parametrize("method", ["sinkhorn", "sinkhorn_stabilized", "sinkhorn_epsilon_scaling", "sinkhorn_log"])
def test_convergence_warning(method):  # :52-74 (sinkhorn / sinkhorn2 part)
    n = 100
    a1, a2, M = dev(gauss(n, 30, 10), gauss(n, 40, 10), dist0(n))
    with pytest.warns(UserWarning):
        ot.sinkhorn(a1, a2, M, 1.0, method=method, stopThr=0, numItermax=1)
    if method in ["sinkhorn", "sinkhorn_stabilized", "sinkhorn_log"]:
        with pytest.warns(UserWarning):
            ot.sinkhorn2(a1, a2, M, 1, method=method, stopThr=0, numItermax=1, warn=True)
        with warnings.catch_warnings():
            warnings.simplefilter("error")
            ot.sinkhorn2(a1, a2, M, 1, method=method, stopThr=0, numItermax=1, warn=False)


def test_not_implemented_method():  # :77-87
    n = 100
    rng = np.random.RandomState(42)
    A_flat = rng.rand(2, 10, 10).reshape(n, 2)
    a1, a2, M = dev(A_flat[:, 0], A_flat[:, 1], dist0(n))
    with pytest.raises(ValueError):
        ot.sinkhorn(a1, a2, M, 0.01, method="new_method")
    with pytest.raises(ValueError):
        ot.sinkhorn2(a1, a2, M, 0.01, method="new_method")


def test_sinkhorn_stabilization():  # :118-126
    n = 100
    a1, a2, M = dev(gauss(n, 30, 10), gauss(n, 40, 10), dist0(n))
    reg = 1e-5
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        loss1 = float(ot.sinkhorn2(a1, a2, M, reg, method="sinkhorn_log"))
        loss2 = float(ot.sinkhorn2(a1, a2, M, reg, tau=1, method="sinkhorn_stabilized"))
    np.testing.assert_allclose(loss1, loss2, atol=1e-06)


def test_sinkhorn_empty():  # :228-279 (the four methods of the path)
    n = 100
    rng = np.random.RandomState(0)
    x = rng.randn(n, 2)
    u = unif(n)
    Md = dev(orc.sqeuclidean(x, x))
    for method in ("sinkhorn_log", "sinkhorn", "sinkhorn_stabilized", "sinkhorn_epsilon_scaling"):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            G, log = ot.sinkhorn([], [], Md, 1, stopThr=1e-10, method=method, verbose=True, log=True)
        G = to_np(G)
        np.testing.assert_allclose(u, G.sum(1), atol=1e-05, err_msg=method)
        np.testing.assert_allclose(u, G.sum(0), atol=1e-05, err_msg=method)


def test_sinkhorn_variants():  # :282-314
    n = 100
    rng = np.random.RandomState(0)
    x = rng.randn(n, 2)
    u = unif(n)
    M = orc.sqeuclidean(x, x)
    ub, Mb = dev(u, M)
    G = orc.sinkhorn(u, u, M, 1, method="sinkhorn", stopThr=1e-10)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        Gl = to_np(ot.sinkhorn(ub, ub, Mb, 1, method="sinkhorn_log", stopThr=1e-10))
        G0 = to_np(ot.sinkhorn(ub, ub, Mb, 1, method="sinkhorn", stopThr=1e-10))
        Gs = to_np(ot.sinkhorn(ub, ub, Mb, 1, method="sinkhorn_stabilized", stopThr=1e-10))
        Ges = to_np(ot.sinkhorn(ub, ub, Mb, 1, method="sinkhorn_epsilon_scaling", stopThr=1e-10))
    np.testing.assert_allclose(G, G0, atol=1e-05)
    np.testing.assert_allclose(G, Gl, atol=1e-05)
    np.testing.assert_allclose(G0, Gs, atol=1e-05)
    np.testing.assert_allclose(G0, Ges, atol=1e-05)


@pytest.mark.parametrize("method", ["sinkhorn", "sinkhorn_stabilized", "sinkhorn_epsilon_scaling", "sinkhorn_log"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_sinkhorn_variants_dtype_device(method, dtype):  # :351-368, and :371-390 for the methods sinkhorn2 takes
    n = 100
    rng = np.random.RandomState(0)
    x = rng.randn(n, 2)
    ub = torch.as_tensor(unif(n), device=DEV, dtype=dtype)
    Mb = torch.as_tensor(orc.sqeuclidean(x, x), device=DEV, dtype=dtype)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        Gb = ot.sinkhorn(ub, ub, Mb, 1, method=method, stopThr=1e-10)
    assert Gb.dtype == Mb.dtype and Gb.device == Mb.device
    if method == "sinkhorn_epsilon_scaling":  # not a sinkhorn2 method in the reference either (_sinkhorn.py:433)
        with pytest.raises(ValueError):
            ot.sinkhorn2(ub, ub, Mb, 1, method=method, stopThr=1e-10)
        return
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        lb = ot.sinkhorn2(ub, ub, Mb, 1, method=method, stopThr=1e-10)
    assert lb.dtype == Mb.dtype and lb.device == Mb.device


def test_sinkhorn_variants_log():  # :482-515
    n = 50
    rng = np.random.RandomState(0)
    x = rng.randn(n, 2)
    u, M = dev(unif(n), orc.sqeuclidean(x, x))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G0, log0 = ot.sinkhorn(u, u, M, 1, method="sinkhorn", stopThr=1e-10, log=True)
        Gl, logl = ot.sinkhorn(u, u, M, 1, method="sinkhorn_log", stopThr=1e-10, log=True)
        Gs, logs = ot.sinkhorn(u, u, M, 1, method="sinkhorn_stabilized", stopThr=1e-10, log=True)
        Ges, loges = ot.sinkhorn(u, u, M, 1, method="sinkhorn_epsilon_scaling", stopThr=1e-10, log=True)
    np.testing.assert_allclose(to_np(G0), to_np(Gs), atol=1e-05)
    np.testing.assert_allclose(to_np(G0), to_np(Gl), atol=1e-05)
    np.testing.assert_allclose(to_np(G0), to_np(Ges), atol=1e-05)


@pytest.mark.parametrize("verbose, warn", list(product([True, False], [True, False])))
def test_sinkhorn_variants_log_multib(verbose, warn):  # :518-556
    n = 50
    rng = np.random.RandomState(0)
    x = rng.randn(n, 2)
    b = rng.rand(n, 3)
    b = b / np.sum(b, 0, keepdims=True)
    u, b, M = dev(unif(n), b, orc.sqeuclidean(x, x))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G0, log0 = ot.sinkhorn(u, b, M, 1, method="sinkhorn", stopThr=1e-10, log=True)
        Gl, logl = ot.sinkhorn(u, b, M, 1, method="sinkhorn_log", stopThr=1e-10, log=True, verbose=verbose, warn=warn)
        Gs, logs = ot.sinkhorn(u, b, M, 1, method="sinkhorn_stabilized", stopThr=1e-10, log=True, verbose=verbose,
                               warn=warn)
    np.testing.assert_allclose(to_np(G0), to_np(Gs), atol=1e-05)
    np.testing.assert_allclose(to_np(G0), to_np(Gl), atol=1e-05)


def test_empirical_sinkhorn():  # :1116-1158
    n = 10
    a, b = unif(n), unif(n)
    X_s = np.reshape(1.0 * np.arange(n), (n, 1))
    X_t = np.reshape(1.0 * np.arange(0, n), (n, 1))
    M = orc.sqeuclidean(X_s, X_t)
    M_m = np.sqrt(M)
    ab, bb, X_sb, X_tb, Mb, M_mb = dev(a, b, X_s, X_t, M, M_m)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G_sqe = to_np(ot.bregman.empirical_sinkhorn(X_sb, X_tb, 1))
        sinkhorn_sqe = to_np(ot.sinkhorn(ab, bb, Mb, 1))
        G_log, log_es = ot.bregman.empirical_sinkhorn(X_sb, X_tb, 0.1, log=True)
        sinkhorn_log, log_s = ot.sinkhorn(ab, bb, Mb, 0.1, log=True)
        G_log, sinkhorn_log = to_np(G_log), to_np(sinkhorn_log)
        G_m = to_np(ot.bregman.empirical_sinkhorn(X_sb, X_tb, 1, metric="euclidean"))
        sinkhorn_m = to_np(ot.sinkhorn(ab, bb, M_mb, 1))
        loss_emp_sinkhorn = float(ot.bregman.empirical_sinkhorn2(X_sb, X_tb, 1))
        loss_sinkhorn = float(ot.sinkhorn2(ab, bb, Mb, 1))
    np.testing.assert_allclose(sinkhorn_sqe.sum(1), G_sqe.sum(1), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_sqe.sum(0), G_sqe.sum(0), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_log.sum(1), G_log.sum(1), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_log.sum(0), G_log.sum(0), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_m.sum(1), G_m.sum(1), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_m.sum(0), G_m.sum(0), atol=1e-05)
    np.testing.assert_allclose(loss_emp_sinkhorn, loss_sinkhorn, atol=1e-05)


def test_lazy_empirical_sinkhorn():  # :1197-1277
    n = 10
    a, b = unif(n), unif(n)
    numIterMax = 1000
    X_s = np.reshape(np.arange(n, dtype=np.float64), (n, 1))
    X_t = np.reshape(np.arange(0, n, dtype=np.float64), (n, 1))
    M = orc.sqeuclidean(X_s, X_t)
    M_m = np.sqrt(M)
    ab, bb, X_sb, X_tb, Mb, M_mb = dev(a, b, X_s, X_t, M, M_m)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g = ot.bregman.empirical_sinkhorn(X_sb, X_tb, 1, numIterMax=numIterMax, isLazy=True, batchSize=(1, 3),
                                             verbose=True)
        f, g = to_np(f), to_np(g)
        G_sqe = np.exp(f[:, None] + g[None, :] - M / 1)
        sinkhorn_sqe = to_np(ot.sinkhorn(ab, bb, Mb, 1))
        f, g, log_es = ot.bregman.empirical_sinkhorn(X_sb, X_tb, 1, numIterMax=numIterMax, isLazy=True, batchSize=5,
                                                     log=True)
        f, g = to_np(f), to_np(g)
        G_log = np.exp(f[:, None] + g[None, :] - M / 1)
        sinkhorn_log, log_s = ot.sinkhorn(ab, bb, Mb, 1, log=True)
        sinkhorn_log = to_np(sinkhorn_log)
        f, g = ot.bregman.empirical_sinkhorn(X_sb, X_tb, 1, metric="euclidean", numIterMax=numIterMax, isLazy=True,
                                             batchSize=1)
        f, g = to_np(f), to_np(g)
        G_m = np.exp(f[:, None] + g[None, :] - M_m / 1)
        sinkhorn_m = to_np(ot.sinkhorn(ab, bb, M_mb, 1))
        loss_emp_sinkhorn, log = ot.bregman.empirical_sinkhorn2(X_sb, X_tb, 1, numIterMax=numIterMax, isLazy=True,
                                                                batchSize=5, log=True)
        G_lazy = to_np(log["lazy_plan"][:])
        loss_sinkhorn = float(ot.sinkhorn2(ab, bb, Mb, 1))
        loss_emp_sinkhorn = float(ot.bregman.empirical_sinkhorn2(X_sb, X_tb, 1, numIterMax=numIterMax, isLazy=True,
                                                                 batchSize=1, log=False))
    np.testing.assert_allclose(G_sqe[1, 1], G_lazy[1, 1], atol=1e-03)
    np.testing.assert_allclose(sinkhorn_sqe.sum(1), G_sqe.sum(1), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_sqe.sum(0), G_sqe.sum(0), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_log.sum(1), G_log.sum(1), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_log.sum(0), G_log.sum(0), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_m.sum(1), G_m.sum(1), atol=1e-05)
    np.testing.assert_allclose(sinkhorn_m.sum(0), G_m.sum(0), atol=1e-05)
    np.testing.assert_allclose(loss_emp_sinkhorn, loss_sinkhorn, atol=1e-05)
    np.testing.assert_allclose(G_log, G_lazy, atol=1e-05)


def test_empirical_sinkhorn_divergence():  # :1280-1312
    n = 10
    a = np.linspace(1, n, n)
    a /= a.sum()
    b = unif(n)
    X_s = np.reshape(np.arange(n, dtype=np.float64), (n, 1))
    X_t = np.reshape(np.arange(0, n * 2, 2, dtype=np.float64), (n, 1))
    M, M_s, M_t = orc.sqeuclidean(X_s, X_t), orc.sqeuclidean(X_s, X_s), orc.sqeuclidean(X_t, X_t)
    ab, bb, X_sb, X_tb, Mb, M_sb, M_tb = dev(a, b, X_s, X_t, M, M_s, M_t)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        emp_sinkhorn_div = float(ot.bregman.empirical_sinkhorn_divergence(X_sb, X_tb, 1, a=ab, b=bb))
        sinkhorn_div = float(ot.sinkhorn2(ab, bb, Mb, 1) - 1 / 2 * ot.sinkhorn2(ab, ab, M_sb, 1)
                             - 1 / 2 * ot.sinkhorn2(bb, bb, M_tb, 1))
        emp_sinkhorn_div_np = float(orc.sinkhorn2(a, b, M, 1) - 1 / 2 * orc.sinkhorn2(a, a, M_s, 1)
                                    - 1 / 2 * orc.sinkhorn2(b, b, M_t, 1))
        ot.bregman.empirical_sinkhorn_divergence(X_sb, X_tb, 1, a=ab, b=bb, log=True)
    np.testing.assert_allclose(emp_sinkhorn_div, emp_sinkhorn_div_np, atol=1e-05)
    np.testing.assert_allclose(emp_sinkhorn_div, sinkhorn_div, atol=1e-05)


def _warmstart(a, b, M, reg):
    """A warm start that is not the uniform one: the log-scalings after a few log-domain iterations."""
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        _, log = ot.bregman.sinkhorn_log(a, b, M, reg, numItermax=5, stopThr=0, log=True)
    return log["log_u"], log["log_v"]


def test_sinkhorn_warmstart():  # :1450-1486
    m, n = 10, 20
    Xs, Xt = np.arange(m) * 1.0, np.arange(n) * 1.0
    a, b, M = dev(unif(m), unif(n), orc.sqeuclidean(Xs.reshape(-1, 1), Xt.reshape(-1, 1)))
    reg = 1
    warmstart = _warmstart(a, b, M, reg)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        pi_unif, _ = ot.bregman.sinkhorn(a, b, M, reg, method="sinkhorn", log=True, warmstart=None)
        res = {meth: to_np(ot.bregman.sinkhorn(a, b, M, reg, method=meth, log=True, warmstart=warmstart)[0])
               for meth in ("sinkhorn", "sinkhorn_log", "sinkhorn_stabilized", "sinkhorn_epsilon_scaling")}
    for meth, pi in res.items():
        np.testing.assert_allclose(to_np(pi_unif), pi, atol=1e-05, err_msg=meth)


def test_empirical_sinkhorn_warmstart():  # :1489-1518
    m, n = 10, 20
    Xs, Xt = np.arange(m).reshape(-1, 1) * 1.0, np.arange(n).reshape(-1, 1) * 1.0
    M = orc.sqeuclidean(Xs, Xt)
    Xsb, Xtb, Mb = dev(Xs, Xt, M)
    reg = 1
    warmstart = _warmstart(*dev(unif(m), unif(n)), Mb, reg)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g, _ = ot.bregman.empirical_sinkhorn(X_s=Xsb, X_t=Xtb, reg=reg, isLazy=True, log=True, warmstart=None)
        pi_unif = np.exp(to_np(f)[:, None] + to_np(g)[None, :] - M / reg)
        f, g, _ = ot.bregman.empirical_sinkhorn(X_s=Xsb, X_t=Xtb, reg=reg, isLazy=True, log=True, warmstart=warmstart)
        pi_ws_lazy = np.exp(to_np(f)[:, None] + to_np(g)[None, :] - M / reg)
        pi_ws_not_lazy, _ = ot.bregman.empirical_sinkhorn(X_s=Xsb, X_t=Xtb, reg=reg, isLazy=False, log=True,
                                                          warmstart=warmstart)
    np.testing.assert_allclose(pi_unif, pi_ws_lazy, atol=1e-05)
    np.testing.assert_allclose(pi_unif, to_np(pi_ws_not_lazy), atol=1e-05)


def test_empirical_sinkhorn_divergence_warmstart():  # :1521-1548
    m, n = 10, 20
    Xs, Xt = np.arange(m).reshape(-1, 1) * 1.0, np.arange(n).reshape(-1, 1) * 1.0
    Xsb, Xtb, Mb = dev(Xs, Xt, orc.sqeuclidean(Xs, Xt))
    reg = 1
    warmstart = _warmstart(*dev(unif(m), unif(n)), Mb, reg)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        sd_unif, _ = ot.bregman.empirical_sinkhorn_divergence(X_s=Xsb, X_t=Xtb, reg=reg, isLazy=True, log=True,
                                                              warmstart=None)
        sd_ws_lazy, _ = ot.bregman.empirical_sinkhorn_divergence(X_s=Xsb, X_t=Xtb, reg=reg, isLazy=True, log=True,
                                                                 warmstart=warmstart)
        sd_ws_not_lazy, _ = ot.bregman.empirical_sinkhorn_divergence(X_s=Xsb, X_t=Xtb, reg=reg, isLazy=False, log=True,
                                                                     warmstart=warmstart)
    np.testing.assert_allclose(float(sd_unif), float(sd_ws_lazy), atol=1e-05)
    np.testing.assert_allclose(float(sd_unif), float(sd_ws_not_lazy), atol=1e-05)
