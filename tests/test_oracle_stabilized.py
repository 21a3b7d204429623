"""CPU: oracle sinkhorn_stabilized / sinkhorn_epsilon_scaling against the reference's golden vectors
(tests/golden/golden_stabilized.npz, produced by tests/golden/make_golden_stabilized.py from the unmodified reference)."""
import warnings

import numpy as np
import pytest

from oracle import solvers_oracle as orc

STAB = ["mild_f64", "absorb_f64", "absorb_tau10_f64", "period7_f64", "conv_f64", "conv_absorb_f64", "wide_f64",
        "mild_f32", "absorb_f32"]
KW = ("numItermax", "stopThr", "tau", "print_period", "numInnerItermax", "epsilon0")


def kwargs(g, case):
    kw = {}
    for k in KW:
        if g.has(case, f"kw_{k}"):
            x = g(case, f"kw_{k}")
            kw[k] = int(x) if k in ("numItermax", "print_period", "numInnerItermax") else float(x)
    return kw


def rtol(case):
    return 1e-10 if "f64" in case else 2e-4


def close(x, ref, rt):
    ref = np.asarray(ref)
    np.testing.assert_allclose(np.asarray(x), ref, rtol=rt, atol=rt * max(np.abs(ref).max(), 1e-300))


@pytest.mark.parametrize("name", STAB)
def test_stabilized(gstab, name):
    case = f"stab_{name}"
    M, a, b, reg = (gstab(case, k) for k in ("M", "a", "b", "reg"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, log = orc.sinkhorn_stabilized(a, b, M, float(reg), log=True, **kwargs(gstab, case))
    t = rtol(case)
    assert log["n_iter"] == int(gstab(case, "n_iter"))
    if "absorb" in name:
        assert log["n_absorb"] >= 1
    close(P, gstab(case, "plan"), t)
    for k in ("logu", "logv", "alpha", "beta"):
        close(log[k], gstab(case, k), t)
    ref = gstab(case, "err")
    assert len(log["err"]) == len(ref)
    np.testing.assert_allclose(log["err"], ref, rtol=1e-6 if "f64" in case else 5e-2, atol=1e-13 if "f64" in case else 1e-6)


def test_stabilized_warmstart(gstab):
    case = "stab_absorb_f64"
    M, a, b, reg = (gstab(case, k) for k in ("M", "a", "b", "reg"))
    ws = (gstab(case, "alpha"), gstab(case, "beta"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, log = orc.sinkhorn_stabilized(a, b, M, float(reg), numItermax=40, stopThr=0.0, warmstart=ws, log=True)
    close(P, gstab(case + "_warm", "plan"), 1e-10)
    close(log["alpha"], gstab(case + "_warm", "alpha"), 1e-10)
    close(log["beta"], gstab(case + "_warm", "beta"), 1e-10)
    assert log["n_iter"] == int(gstab(case + "_warm", "n_iter"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P = orc.sinkhorn_stabilized(a, b, M, float(reg), numItermax=200, stopThr=0.0)
    close(np.sum(P * M), gstab(case + "_loss", "value"), 1e-10)


@pytest.mark.parametrize("name", ["f64", "short_f64"])
def test_epsilon_scaling(gstab, name):
    case = f"eps_{name}"
    M, a, b, reg = (gstab(case, k) for k in ("M", "a", "b", "reg"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, log = orc.sinkhorn_epsilon_scaling(a, b, M, float(reg), log=True, **kwargs(gstab, case))
    assert log["niter"] == int(gstab(case, "niter"))
    close(P, gstab(case, "plan"), 1e-9)
    close(log["alpha"], gstab(case, "alpha"), 1e-9)
    close(log["beta"], gstab(case, "beta"), 1e-9)
    ref = gstab(case, "err")
    assert len(log["err"]) == len(ref)
    np.testing.assert_allclose(log["err"], ref, rtol=1e-5, atol=1e-20)


# ---- several target histograms (2-D b): tests/golden/golden_stabilized_multi.npz ----
STABM = ["mild_f64", "absorb_f64", "absorb_tau10_f64", "period7_f64", "conv_f64", "conv_absorb_f64", "wide_f64",
         "one_hist_f64", "mild_f32", "absorb_f32"]


@pytest.mark.parametrize("name", STABM)
def test_stabilized_multi(gstabm, name):
    case = f"stabm_{name}"
    M, a, b, reg = (gstabm(case, k) for k in ("M", "a", "b", "reg"))
    assert b.ndim == 2
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        res, log = orc.sinkhorn_stabilized(a, b, M, float(reg), log=True, **kwargs(gstabm, case))
    t = rtol(case)
    assert log["n_iter"] == int(gstabm(case, "n_iter"))
    if "absorb" in name:
        assert log["n_absorb"] >= 1
    assert res.shape == (b.shape[1],)
    close(res, gstabm(case, "losses"), t)
    for k in ("logu", "logv", "alpha", "beta"):
        assert log[k].shape == gstabm(case, k).shape
        close(log[k], gstabm(case, k), t)
    ref = gstabm(case, "err")
    assert len(log["err"]) == len(ref)
    np.testing.assert_allclose(log["err"], ref, rtol=1e-6 if "f64" in case else 5e-2, atol=1e-13 if "f64" in case else 1e-6)


def test_stabilized_multi_warmstart(gstabm):
    case = "stabm_absorb_f64"
    M, a, b, reg = (gstabm(case, k) for k in ("M", "a", "b", "reg"))
    ws = (gstabm(case + "_warm", "ws_alpha"), gstabm(case + "_warm", "ws_beta"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        res, log = orc.sinkhorn_stabilized(a, b, M, float(reg), numItermax=40, stopThr=0.0, warmstart=ws, log=True)
    close(res, gstabm(case + "_warm", "losses"), 1e-10)
    close(log["alpha"], gstabm(case + "_warm", "alpha"), 1e-10)
    close(log["beta"], gstabm(case + "_warm", "beta"), 1e-10)
    assert log["n_iter"] == int(gstabm(case + "_warm", "n_iter"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        res = orc.sinkhorn_stabilized(a, b, M, float(reg), numItermax=200, stopThr=0.0)
    close(res, gstabm(case + "_loss", "value"), 1e-10)
