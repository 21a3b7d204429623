"""CPU: oracle entropic_gromov_wasserstein against the reference's golden vectors
(tests/golden/golden_gromov.npz, produced by tests/golden/make_golden_gromov.py from the unmodified reference)."""
import warnings

import numpy as np
import pytest

from oracle import solvers_oracle as orc

CASES = ["pgd_sq_f64", "pgd_kl_f64", "ppa_sq_f64", "pgd_asym_f64", "pgd_warm_f64", "ppa_warm_f64", "pgd_sq_f32"]
INNER = dict(numItermax=100, stopThr=0.0)


def gw_kwargs(g, case):
    kw = {}
    for k in ("loss_fun", "solver"):
        kw[k] = str(g(case, f"kw_{k}"))
    kw["epsilon"] = float(g(case, "kw_epsilon"))
    kw["max_iter"] = int(g(case, "kw_max_iter"))
    if g.has(case, "kw_warmstart"):
        kw["warmstart"] = bool(g(case, "kw_warmstart"))
    return kw


@pytest.mark.parametrize("name", CASES)
def test_entropic_gw(ggw, name):
    case = f"egw_{name}"
    C1, C2, p, q = (ggw(case, k) for k in ("C1", "C2", "p", "q"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        T, log = orc.entropic_gromov_wasserstein(C1, C2, p, q, tol=0.0, log=True, **gw_kwargs(ggw, case), **INNER)
    rt = 1e-10 if "f64" in name else 5e-4
    ref = ggw(case, "T")
    np.testing.assert_allclose(T, ref, rtol=rt, atol=rt * ref.max())
    np.testing.assert_allclose(log["gw_dist"], ggw(case, "gw_dist"), rtol=rt * 10, atol=1e-14)
    np.testing.assert_allclose(log["gw_dist"], ggw(case, "gw2"), rtol=rt * 10, atol=1e-14)
    eref = ggw(case, "err")
    assert len(log["err"]) == len(eref)
    np.testing.assert_allclose(log["err"], eref, rtol=1e-6 if "f64" in name else 1e-2, atol=1e-12)


def test_entropic_gw_default_weights_converging(ggw):
    case = "egw_default_conv_f64"
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        T, log = orc.entropic_gromov_wasserstein(ggw(case, "C1"), ggw(case, "C2"), epsilon=0.1, tol=1e-7, max_iter=200,
                                                 log=True, **INNER)
    assert len(log["err"]) == len(ggw(case, "err"))
    np.testing.assert_allclose(T, ggw(case, "T"), rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(log["gw_dist"], ggw(case, "gw_dist"), rtol=1e-9)
