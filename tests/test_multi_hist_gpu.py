"""GPU parity of the solvers that keep SEVERAL target histograms (2-D b) on one kernel matrix, against the CPU oracle:

    ot.bregman.sinkhorn_knopp             ot/bregman/_sinkhorn.py:626-663   joint guard, joint error every 10 iterations
    ot.unbalanced.sinkhorn_knopp_unbalanced   ot/unbalanced/_sinkhorn.py:749-787   joint guard, joint error every iteration

(sinkhorn_stabilized with a 2-D b is in test_stabilized_gpu.py.)  Everything goes through b2ot_multi_scaling_step.
Tolerances: 1e-10 relative in fp64, 1e-4 in fp32 (north_star); iteration counts and stop decisions must be equal in fp64.
"""
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot
    from pot_b200 import _dense, _lib
from oracle import sinkhorn_oracle as orc
from oracle import solvers_oracle as sorc
from test_solvers_gpu import assert_vec, to_dev

DEV = "cuda:0"


def problem(n, m, h, dtype, seed=0, mass=1.0):
    rng = np.random.RandomState(seed + n + m + h)
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = orc.sqeuclidean(xs, xt)
    M = (M / M.max()).astype(dtype)
    a = rng.rand(n) + 0.1
    b = rng.rand(m, h) + 0.1
    return (a / a.sum()).astype(dtype), (mass * b / b.sum(0)).astype(dtype), M


def errs_of(log):
    return np.array([float(e) for e in log["err"]], dtype=np.float64)


@pytest.mark.parametrize("dtype,n,m,h,reg,kw", [
    (np.float64, 60, 50, 3, 0.1, dict(numItermax=45, stopThr=0.0)),          # fixed count, last iterate not a checked one
    (np.float64, 60, 50, 3, 0.1, dict(numItermax=41, stopThr=0.0)),          # last iterate IS a checked one (probe step)
    (np.float64, 130, 77, 5, 0.05, dict(numItermax=3000, stopThr=1e-9)),     # joint stop: all histograms below the threshold
    (np.float64, 33, 300, 9, 0.1, dict(numItermax=1000, stopThr=1e-7)),      # two chunks of histograms
    (np.float64, 40, 52, 1, 0.1, dict(numItermax=200, stopThr=1e-8)),
    (np.float32, 60, 50, 4, 0.1, dict(numItermax=45, stopThr=0.0)),
    (np.float32, 200, 333, 2, 0.1, dict(numItermax=500, stopThr=1e-4)),
])
def test_knopp_joint_histograms_vs_oracle(dtype, n, m, h, reg, kw):
    a, b, M = problem(n, m, h, dtype)
    f64 = dtype == np.float64
    tol = 1e-10 if f64 else 1e-4
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        rr, lr = orc.sinkhorn_knopp(a, b, M, reg, log=True, **kw)
        rg, lg = ot.bregman.sinkhorn_knopp(*to_dev(a, b, M), reg, log=True, **kw)
    if f64 or kw["stopThr"] == 0.0:
        assert lg["niter"] == lr["niter"]
        assert len(lg["err"]) == len(lr["err"])
        np.testing.assert_allclose(errs_of(lg), np.asarray(lr["err"], dtype=np.float64), rtol=1e-6 if f64 else 5e-2,
                                   atol=1e-13 if f64 else 2e-6)
    assert tuple(rg.shape) == (h,)
    assert_vec(rg, rr, tol)
    assert_vec(torch.log(lg["u"]), np.log(lr["u"]), tol)
    assert_vec(torch.log(lg["v"]), np.log(lr["v"]), tol)


def test_knopp_joint_stop_is_not_per_histogram():
    """One well-conditioned and one slow histogram: the reference keeps iterating BOTH until the joint error passes, so the
    fast one ends with more iterations than it would take alone."""
    a, b, M = problem(50, 60, 2, np.float64, seed=3)
    b[:, 1] = np.linspace(0.01, 1.0, 60) ** 4
    b[:, 1] /= b[:, 1].sum()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        _, l_joint = orc.sinkhorn_knopp(a, b, M, 0.02, numItermax=5000, stopThr=1e-9, log=True)
        _, l_first = orc.sinkhorn_knopp(a, b[:, 0].copy(), M, 0.02, numItermax=5000, stopThr=1e-9, log=True)
        rg, lg = ot.sinkhorn(*to_dev(a, b, M), 0.02, numItermax=5000, stopThr=1e-9, log=True)
    assert l_joint["niter"] > l_first["niter"]
    assert lg["niter"] == l_joint["niter"]
    assert_vec(torch.log(lg["u"]), np.log(l_joint["u"]), 1e-10)


def test_knopp_joint_guard_and_warnings():
    a, b, M = problem(40, 30, 3, np.float64, seed=5)
    Mbad = M.copy()
    Mbad[:, :4] = 1e6  # four columns of K underflow to 0: K^T u == 0 there
    with warnings.catch_warnings(record=True) as wr:
        warnings.simplefilter("always")
        rr, lr = orc.sinkhorn_knopp(a, b, Mbad, 1.0, numItermax=100, stopThr=1e-9, log=True)
    with warnings.catch_warnings(record=True) as wg:
        warnings.simplefilter("always")
        rg, lg = ot.bregman.sinkhorn_knopp(*to_dev(a, b, Mbad), 1.0, numItermax=100, stopThr=1e-9, log=True)
    ref = [str(w.message) for w in wr if "numerical errors" in str(w.message)]
    got = [str(w.message) for w in wg if "numerical errors" in str(w.message)]
    assert ref and got == ref
    assert lg["niter"] == lr["niter"] and len(lg["err"]) == len(lr["err"])
    assert_vec(lg["u"], lr["u"], 1e-10)  # the iterate before the guard fired
    with pytest.warns(UserWarning, match="did not converge"):
        ot.bregman.sinkhorn_knopp(*to_dev(a, b, M), 0.01, numItermax=5, stopThr=1e-12)


@pytest.mark.parametrize("dtype,n,m,h,reg,reg_m,kw", [
    (np.float64, 60, 50, 3, 0.1, 1.0, dict(numItermax=40, stopThr=0.0)),
    (np.float64, 130, 77, 5, 0.05, (0.5, 2.0), dict(numItermax=3000, stopThr=1e-10)),
    (np.float64, 33, 300, 9, 0.1, (float("inf"), 1.0), dict(numItermax=2000, stopThr=1e-9)),
    (np.float64, 40, 52, 1, 0.1, 1.0, dict(numItermax=500, stopThr=1e-9)),
    (np.float32, 60, 50, 4, 0.1, 1.0, dict(numItermax=40, stopThr=0.0)),
    (np.float32, 200, 333, 2, 0.1, (2.0, 0.5), dict(numItermax=1000, stopThr=1e-5)),
])
def test_unbalanced_joint_histograms_vs_oracle(dtype, n, m, h, reg, reg_m, kw):
    a, b, M = problem(n, m, h, dtype, mass=1.3)
    f64 = dtype == np.float64
    tol = 1e-10 if f64 else 1e-4
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        rr, lr = sorc.sinkhorn_knopp_unbalanced(a, b, M, reg, reg_m, log=True, **kw)
        rg, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(*to_dev(a, b, M), reg, reg_m, log=True, **kw)
    if f64 or kw["stopThr"] == 0.0:
        assert len(lg["err"]) == len(lr["err"])
        np.testing.assert_allclose(errs_of(lg), np.asarray(lr["err"], dtype=np.float64), rtol=1e-6 if f64 else 5e-2,
                                   atol=1e-13 if f64 else 2e-6)
    assert tuple(rg.shape) == (h,)
    assert_vec(rg, rr, tol)
    assert_vec(lg["logu"], lr["logu"], tol)
    assert_vec(lg["logv"], lr["logv"], tol)


def test_unbalanced_joint_guard():
    a, b, M = problem(40, 30, 3, np.float64, seed=7, mass=0.7)
    Mbad = M.copy()
    Mbad[:4] *= 1e7
    with warnings.catch_warnings(record=True) as wr:
        warnings.simplefilter("always")
        rr, lr = sorc.sinkhorn_knopp_unbalanced(a, b, Mbad, 0.01, 1.0, numItermax=30, stopThr=0.0, log=True)
    with warnings.catch_warnings(record=True) as wg:
        warnings.simplefilter("always")
        rg, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(*to_dev(a, b, Mbad), 0.01, 1.0, numItermax=30, stopThr=0.0,
                                                         log=True)
    ref = [str(w.message) for w in wr if "Numerical errors" in str(w.message)]
    got = [str(w.message) for w in wg if "Numerical errors" in str(w.message)]
    assert ref and got == ref
    assert len(lg["err"]) == len(lr["err"])
    assert_vec(lg["logu"], lr["logu"], 1e-10)


def test_multi_step_unbalanced_order_cabi():
    """b2ot_multi_scaling_step, UNBALANCED mode against fp64 torch: u-then-v with exponents, error, stop at the new iterate."""
    lib = _lib.load()
    gen = torch.Generator().manual_seed(9)
    n, m, h = 90, 141, 3
    K = (torch.rand(n, m, generator=gen, dtype=torch.float64) + 0.05).to(DEV)
    a = (torch.rand(n, generator=gen, dtype=torch.float64) + 0.1).to(DEV)
    B = (torch.rand(m, h, generator=gen, dtype=torch.float64) + 0.1).to(DEV)
    up = (torch.rand(n, h, generator=gen, dtype=torch.float64) + 0.1).to(DEV)
    vp = (torch.rand(m, h, generator=gen, dtype=torch.float64) + 0.1).to(DEV)
    un, vn = torch.empty_like(up), torch.empty_like(vp)
    stats = torch.zeros(16, dtype=torch.float64, device=DEV)
    eh = torch.zeros(4, dtype=torch.float64, device=DEV)
    scratch = torch.empty(lib.b2ot_multi_scratch_bytes(n, m, h), dtype=torch.uint8, device=DEV)

    def step(it, thr):
        return lib.b2ot_multi_scaling_step(_lib.F64, _lib.MULTI_UNBALANCED, K.data_ptr(), n, m, K.stride(0), h,
                                           a.data_ptr(), B.data_ptr(), up.data_ptr(), vp.data_ptr(), un.data_ptr(),
                                           vn.data_ptr(), it, 1, thr, 0.75, 0.6, stats.data_ptr(), eh.data_ptr(), 4,
                                           scratch.data_ptr(), _dense.stream_ptr(torch.device(DEV)))

    assert step(0, 0.0) == 0
    uref = (a[:, None] / (K @ vp)) ** 0.75
    vref = (B / (K.T @ uref)) ** 0.6
    torch.testing.assert_close(un, uref, rtol=1e-12, atol=0)
    torch.testing.assert_close(vn, vref, rtol=1e-12, atol=0)
    err = 0.5 * ((uref - up).abs().max() / max(uref.abs().max().item(), up.abs().max().item(), 1.0)
                 + (vref - vp).abs().max() / max(vref.abs().max().item(), vp.abs().max().item(), 1.0))
    st = stats.tolist()
    assert st[12] == 0.0 and st[14] == 1.0
    np.testing.assert_allclose(st[15], float(err), rtol=1e-12)
    np.testing.assert_allclose(eh[0].item(), float(err), rtol=1e-12)
    # a generous threshold: converged at this step, and the next step is a no-op
    assert step(1, 10.0) == 0
    st = stats.tolist()
    assert st[12] == 1.0 and st[13] == 1.0 and st[14] == 2.0
    un.fill_(-1.0)
    assert step(2, 10.0) == 0
    assert float(un.min()) == -1.0 and stats.tolist()[14] == 2.0


def test_knopp_joint_warmstart_and_sinkhorn2():
    """warmstart = log-scalings of shape (dim, n_hists) (:614); ot.sinkhorn2 with a 2-D b returns the same losses."""
    a, b, M = problem(45, 38, 3, np.float64, seed=13)
    rng = np.random.RandomState(4)
    ws = (0.1 * rng.randn(45, 3), 0.1 * rng.randn(38, 3))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        rr, lr = orc.sinkhorn_knopp(a, b, M, 0.1, numItermax=25, stopThr=0.0, log=True, warmstart=ws)
        ad, bd, Md = to_dev(a, b, M)
        rg, lg = ot.sinkhorn(ad, bd, Md, 0.1, numItermax=25, stopThr=0.0, log=True, warmstart=tuple(to_dev(*ws)))
        r2 = ot.sinkhorn2(ad, bd, Md, 0.1, numItermax=25, stopThr=0.0, warmstart=tuple(to_dev(*ws)))
    assert lg["niter"] == lr["niter"] == 24
    assert_vec(rg, rr, 1e-10)
    assert_vec(r2, rr, 1e-10)
    assert_vec(torch.log(lg["v"]), np.log(lr["v"]), 1e-10)


# ---- the reference's own multi-histogram tests (test/test_bregman.py), on the CUDA path ----
def _multi_b_problem(n):
    rng = np.random.RandomState(0)
    x = rng.randn(n, 2)
    u = np.full(n, 1.0 / n)
    b = rng.rand(n, 3)
    b = b / np.sum(b, 0, keepdims=True)
    return u, b, orc.sqeuclidean(x, x)


@pytest.mark.parametrize("method", ["sinkhorn", "sinkhorn_stabilized", "sinkhorn_log"])
def test_reference_sinkhorn_multi_b(method):
    """test/test_bregman.py:128-160: the losses of a 2-D b equal the sinkhorn2 values of its columns."""
    u, b, M = _multi_b_problem(10)
    ud, bd, Md = to_dev(u, b, M)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        loss0, log = ot.sinkhorn(ud, bd, Md, 0.1, method=method, stopThr=1e-10, log=True)
        loss = [float(ot.sinkhorn2(ud, bd[:, k].contiguous(), Md, 0.1, method=method, stopThr=1e-10)) for k in range(3)]
    np.testing.assert_allclose(loss0.cpu().numpy(), loss, atol=1e-4)


def test_reference_sinkhorn_variants_multi_b():
    """test/test_bregman.py:422-479: the three methods agree on a 2-D b (sinkhorn and sinkhorn2), and with numpy."""
    u, b, M = _multi_b_problem(50)
    ud, bd, Md = to_dev(u, b, M)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G = orc.sinkhorn(u, b, M, 1, method="sinkhorn", stopThr=1e-10)
        res = {m: ot.sinkhorn(ud, bd, Md, 1, method=m, stopThr=1e-10).cpu().numpy()
               for m in ("sinkhorn", "sinkhorn_log", "sinkhorn_stabilized")}
        res2 = {m: ot.sinkhorn2(ud, bd, Md, 1, method=m, stopThr=1e-10).cpu().numpy()
                for m in ("sinkhorn", "sinkhorn_log", "sinkhorn_stabilized")}
    for r in (res, res2):
        np.testing.assert_allclose(G, r["sinkhorn"], atol=1e-5)
        np.testing.assert_allclose(G, r["sinkhorn_log"], atol=1e-5)
        np.testing.assert_allclose(r["sinkhorn"], r["sinkhorn_stabilized"], atol=1e-5)


def test_reference_stabilized_vs_sinkhorn_multidim():
    """test/test_bregman.py:1346-1374: 1-D Gaussians, two target histograms, stabilized == plain."""
    n = 100
    x = np.arange(n, dtype=np.float64)

    def gauss(m, s):
        h = np.exp(-((x - m) ** 2) / (2 * s**2))
        return h / h.sum()

    a, b = gauss(20, 5), np.vstack((gauss(60, 8), gauss(30, 4))).T
    M = (x[:, None] - x[None, :]) ** 2
    M /= np.median(M)
    ad, bd, Md = to_dev(a, np.ascontiguousarray(b), M)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G_np, _ = orc.sinkhorn_knopp(a, b, M, 0.1, log=True)
        G, log = ot.bregman.sinkhorn(ad, bd, Md, reg=0.1, method="sinkhorn_stabilized", log=True)
        G2, log2 = ot.bregman.sinkhorn(ad, bd, Md, 0.1, method="sinkhorn", log=True)
    np.testing.assert_allclose(G_np, G2.cpu().numpy())
    np.testing.assert_allclose(G.cpu().numpy(), G2.cpu().numpy())
