"""CPU: the solvers oracle (oracle/solvers_oracle.py) against the reference's golden vectors
(tests/golden/golden_solvers.npz, produced by tests/golden/make_golden_solvers.py from the unmodified reference)."""
import warnings

import numpy as np
import pytest

from oracle import solvers_oracle as orc


def rtol(case):
    return 1e-10 if "f64" in case else 2e-4


def close(x, ref, rt, scale=None):
    ref = np.asarray(ref)
    s = np.abs(ref).max() if scale is None else scale
    np.testing.assert_allclose(np.asarray(x), ref, rtol=rt, atol=rt * max(s, 1e-300))


SOLVE_CASES = [f"solve_{t}_{s}" for t in ("f64", "f32") for s in ("40x52", "128x96")]


@pytest.mark.parametrize("case", SOLVE_CASES)
@pytest.mark.parametrize("rt", ["KL", "entropy"])
def test_solve_balanced(gs, case, rt):
    M, a, b, reg, its = (gs(case, k) for k in ("M", "a", "b", "reg", "its"))
    r = orc.solve(M, a, b, reg=float(reg), reg_type=rt, max_iter=int(its), tol=0)
    t = rtol(case)
    close(r["plan"], gs(f"{case}/{rt}", "plan"), t)
    close(r["potentials"][0], gs(f"{case}/{rt}", "f"), t)
    close(r["potentials"][1], gs(f"{case}/{rt}", "g"), t)
    close(r["value"], gs(f"{case}/{rt}", "value"), t)
    close(r["value_linear"], gs(f"{case}/{rt}", "value_linear"), t)
    # envelope gradients
    close(r["grad_M"], gs(f"{case}/{rt}_envelope", "gM"), t)
    close(r["grad_a"], gs(f"{case}/{rt}_envelope", "ga"), t, scale=np.abs(gs(f"{case}/{rt}_envelope", "ga")).max())
    close(r["grad_b"], gs(f"{case}/{rt}_envelope", "gb"), t, scale=np.abs(gs(f"{case}/{rt}_envelope", "gb")).max())
    # last step
    r = orc.solve(M, a, b, reg=float(reg), reg_type=rt, max_iter=int(its), tol=0, grad="last_step")
    close(r["plan"], gs(f"{case}/{rt}_last_step", "plan"), t)
    close(r["value"], gs(f"{case}/{rt}_last_step", "value"), t)
    close(r["potentials"][0], gs(f"{case}/{rt}_last_step", "f"), t)


@pytest.mark.parametrize("case", SOLVE_CASES)
def test_solve_default_and_unbalanced(gs, case):
    M, a, b, reg, its = (gs(case, k) for k in ("M", "a", "b", "reg", "its"))
    t = rtol(case)
    r = orc.solve(M, reg=float(reg), tol=1e-9 if "f64" in case else 1e-6, max_iter=3000)
    close(r["plan"], gs(f"{case}/default", "plan"), t)
    close(r["value"], gs(f"{case}/default", "value"), t)
    for utag, um in (("unb1", 1.0), ("unb_pair", (0.5, 2.0))):
        r = orc.solve_unbalanced(M, a, b, reg=float(reg), unbalanced=um, max_iter=int(its), tol=0)
        close(r["plan"], gs(f"{case}/{utag}", "plan"), t)
        close(r["value"], gs(f"{case}/{utag}", "value"), t)
        close(r["value_linear"], gs(f"{case}/{utag}", "value_linear"), t)
        close(r["potentials"][0], gs(f"{case}/{utag}", "f"), t, scale=1.0)


UNB_CASES = [f"unb_{t}_{s}" for t in ("f64", "f32") for s in ("33x47", "128x160")]
UNB_VARIANTS = {"scalar": dict(reg_m=1.0), "pair": dict(reg_m=(0.5, 2.0)), "semi": dict(reg_m=(float("inf"), 1.0)),
                "entropy": dict(reg_m=1.0, reg_type="entropy")}


@pytest.mark.parametrize("case", UNB_CASES)
@pytest.mark.parametrize("variant", list(UNB_VARIANTS))
def test_unbalanced_fixed(gs, case, variant):
    M, a, b, reg, its = (gs(case, k) for k in ("M", "a", "b", "reg", "its"))
    t = rtol(case)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = orc.sinkhorn_knopp_unbalanced(a, b, M, float(reg), numItermax=int(its), stopThr=0, log=True,
                                              **UNB_VARIANTS[variant])
    c = f"{case}/{variant}"
    close(P, gs(c, "plan"), t)
    close(lg["logu"], gs(c, "logu"), t, scale=1.0)
    close(lg["logv"], gs(c, "logv"), t, scale=1.0)
    close(lg["cost"], gs(c, "cost"), t)
    close(lg["total_cost"], gs(c, "total_cost"), t, scale=abs(float(gs(c, "cost"))))
    assert len(lg["err"]) == len(gs(c, "err"))
    close(lg["err"][:5], gs(c, "err")[:5], 1e-6 if "f64" in case else 1e-2)


@pytest.mark.parametrize("case", UNB_CASES)
def test_unbalanced_conv_c_warm_multib(gs, case):
    M, a, b, reg, its = (gs(case, k) for k in ("M", "a", "b", "reg", "its"))
    t = rtol(case)
    P, lg = orc.sinkhorn_knopp_unbalanced(a, b, M, float(reg), 1.0, numItermax=5000,
                                          stopThr=1e-6 if "f32" in case else 1e-9, log=True)
    if "f64" in case:  # fp32 stop iteration can move by rounding
        assert len(lg["err"]) == int(gs(f"{case}/conv", "nerr"))
    close(P, gs(f"{case}/conv", "plan"), 1e-4 if "f32" in case else 1e-8)
    c = f"{case}/c_warm"
    P, lg = orc.sinkhorn_knopp_unbalanced(a, b, M, float(reg), 1.0, c=gs(c, "c"), warmstart=(gs(c, "ws_u"), gs(c, "ws_v")),
                                          numItermax=int(its), stopThr=0, log=True)
    close(P, gs(c, "plan"), t)
    close(lg["total_cost"], gs(c, "total_cost"), t, scale=abs(float(gs(c, "cost"))))
    c = f"{case}/multib"
    L, lg = orc.sinkhorn_knopp_unbalanced(a, gs(c, "b2"), M, float(reg), 1.0, numItermax=int(its), stopThr=0, log=True)
    close(L, gs(c, "loss"), t)
    close(lg["logv"], gs(c, "logv"), t, scale=1.0)


def test_unbalanced_guard(gs):
    M, a, b, reg = (gs("unb_guard", k) for k in ("M", "a", "b", "reg"))
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        P, lg = orc.sinkhorn_knopp_unbalanced(a, b, M, float(reg), 1.0, numItermax=50, stopThr=0, log=True)
    assert int(any("Numerical errors" in str(x.message) for x in w)) == int(gs("unb_guard", "warned"))
    assert len(lg["err"]) == int(gs("unb_guard", "nerr"))
    close(lg["logu"], gs("unb_guard", "logu"), 1e-10, scale=1.0)
    close(P, gs("unb_guard", "plan"), 1e-10)


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_empirical_loss_divergence_solve_sample(gs, tag):
    base = f"emp2_{tag}"
    xs, xt, a, b, reg, its = (gs(base, k) for k in ("xs", "xt", "a", "b", "reg", "its"))
    t = 1e-10 if tag == "f64" else 2e-4
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        L, lg = orc.empirical_sinkhorn2_lazy(xs, xt, float(reg), a, b, numIterMax=int(its), stopThr=-1, batchSize=32,
                                             log=True)
        close(L, gs(base + "/loss", "loss"), t)
        close(lg["u"], gs(base + "/loss", "f"), t)
        assert lg["niter"] == int(gs(base + "/loss", "niter"))
        D, (l_ab, l_a, l_b) = orc.empirical_sinkhorn_divergence_lazy(xs, xt, float(reg), a, b, numIterMax=int(its),
                                                                     stopThr=-1, batchSize=32)
        close(l_ab, gs(base + "/div", "l_ab"), t)
        close(l_a, gs(base + "/div", "l_a"), t)
        close(l_b, gs(base + "/div", "l_b"), t)
        close(D, gs(base + "/div", "div"), t, scale=abs(float(gs(base + "/div", "l_ab"))))
        r = orc.solve_sample_lazy(xs, xt, a, b, reg=float(reg), max_iter=int(its), tol=-1, batch_size=32)
        close(r["value_linear"], gs(base + "/solve_sample", "value_linear"), t)
        close(r["potentials"][1], gs(base + "/solve_sample", "g"), t)
