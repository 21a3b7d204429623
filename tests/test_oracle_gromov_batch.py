"""CPU: oracle solve_gromov_batch against the reference's golden vectors
(tests/golden/golden_gromov_batch.npz, produced by tests/golden/make_golden_gromov_batch.py from the unmodified
reference)."""
import numpy as np
import pytest

from oracle import solvers_oracle as orc

CASES = ["gw_f64", "gw_rect_f64", "fgw_f64", "gw_asym_f64", "gw_feat_f64", "gw_kl_f64", "gw_f32"]
FIX = dict(tol=0.0, tol_inner=0.0)


def gwb_kwargs(g, case):
    kw = {"reg": float(g(case, "kw_reg")), "max_iter": int(g(case, "kw_max_iter")),
          "max_iter_inner": int(g(case, "kw_max_iter_inner"))}
    if g.has(case, "kw_loss"):
        kw["loss"] = str(g(case, "kw_loss"))
        kw["logits"] = bool(g(case, "kw_logits"))
    alpha = float(g(case, "alpha"))
    if alpha >= 0:
        kw["alpha"] = alpha
    return kw


@pytest.mark.parametrize("name", CASES)
def test_solve_gromov_batch(ggwb, name):
    case = f"gwb_{name}"
    kw = gwb_kwargs(ggwb, case)
    if "alpha" in kw:
        kw["M"] = ggwb(case, "M")
    r = orc.solve_gromov_batch(ggwb(case, "C1"), ggwb(case, "C2"), a=ggwb(case, "a"), b=ggwb(case, "b"), **kw, **FIX)
    rt = 1e-10 if "f64" in name else 5e-4
    ref = ggwb(case, "plan")
    assert r["n_iter"] == int(ggwb(case, "n_iter"))
    np.testing.assert_allclose(r["plan"], ref, rtol=rt, atol=rt * ref.max())
    for k in ("value", "value_linear", "value_quad"):
        np.testing.assert_allclose(r[k], ggwb(case, k), rtol=rt * 10, atol=rt)


def test_default_weights_converging(ggwb):
    case = "gwb_default_conv_f64"
    r = orc.solve_gromov_batch(ggwb(case, "C1"), ggwb(case, "C2"), reg=0.1, max_iter=50, tol=1e-6, max_iter_inner=50,
                               tol_inner=1e-7)
    assert r["n_iter"] == int(ggwb(case, "n_iter"))
    np.testing.assert_allclose(r["plan"], ggwb(case, "plan"), rtol=1e-8, atol=1e-12)
    np.testing.assert_allclose(r["value"], ggwb(case, "value"), rtol=1e-8)
