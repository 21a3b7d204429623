"""GPU parity tests of ot.gromov.entropic_gromov_wasserstein(2) against the golden vectors
of the reference (tests/golden/golden_gromov.npz) and the CPU oracle.  The inner entropic solves run in libb2ot.so.

Fixed work (outer tol = 0, inner numItermax fixed with stopThr = 0); 1e-10 relative in fp64, 1e-4 in fp32 (north_star).
"""
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot
from oracle import solvers_oracle as sorc
from test_oracle_gromov import CASES, INNER, gw_kwargs
from test_solvers_gpu import assert_plan, assert_scalar, to_dev

DEV = "cuda:0"


@pytest.mark.parametrize("name", CASES)
def test_entropic_gw_golden(ggw, name):
    case = f"egw_{name}"
    C1, C2, p, q = to_dev(*(ggw(case, k) for k in ("C1", "C2", "p", "q")))
    kw = gw_kwargs(ggw, case)
    tol = 1e-10 if "f64" in name else 1e-4
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        T, log = ot.gromov.entropic_gromov_wasserstein(C1, C2, p, q, tol=0.0, log=True, **kw, **INNER)
        val = ot.gromov.entropic_gromov_wasserstein2(C1, C2, p, q, tol=0.0, **kw, **INNER)
    assert T.dtype == C1.dtype and T.device == C1.device
    assert_plan(T, ggw(case, "T"), tol)
    ref = float(ggw(case, "gw_dist"))
    assert_scalar(log["gw_dist"], ref, tol * 10, scale=max(abs(ref), 1e-3))
    assert_scalar(val, ref, tol * 10, scale=max(abs(ref), 1e-3))
    eref = ggw(case, "err")
    errs = [float(e) for e in log["err"]]
    assert len(errs) == len(eref)
    np.testing.assert_allclose(errs, eref, rtol=1e-5 if "f64" in name else 2e-2, atol=1e-11 if "f64" in name else 1e-6)
    # marginal violation of the coupling next to the reference's
    Tr = ggw(case, "T")
    assert abs(float((T.sum(1) - p).norm()) - np.linalg.norm(Tr.sum(1) - ggw(case, "p"))) < (1e-10 if "f64" in name else 1e-5)


def test_entropic_gw_default_weights_converging(ggw):
    case = "egw_default_conv_f64"
    C1, C2 = to_dev(ggw(case, "C1"), ggw(case, "C2"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        T, log = ot.gromov.entropic_gromov_wasserstein(C1, C2, epsilon=0.1, tol=1e-7, max_iter=200, log=True, **INNER)
        val, log2 = ot.gromov.entropic_gromov_wasserstein2(C1, C2, epsilon=0.1, tol=1e-7, max_iter=200, log=True,
                                                           **INNER)
    assert len(log["err"]) == len(ggw(case, "err"))
    assert_plan(T, ggw(case, "T"), 1e-8)
    assert_scalar(log["gw_dist"], ggw(case, "gw_dist"), 1e-8)
    assert_scalar(val, ggw(case, "gw_dist"), 1e-8)
    assert torch.equal(log2["T"], T)


@pytest.mark.parametrize("solver", ["PGD", "PPA"])
def test_entropic_gw_vs_oracle_larger(solver):
    """n = 400, m = 350: the inner solves take the fused cluster sweep."""
    rng = np.random.RandomState(3)
    xs, xt = rng.randn(400, 2), rng.randn(350, 3)
    C1 = ((xs[:, None] - xs[None]) ** 2).sum(-1)
    C2 = ((xt[:, None] - xt[None]) ** 2).sum(-1)
    C1, C2 = C1 / C1.max(), C2 / C2.max()
    kw = dict(epsilon=0.05 if solver == "PGD" else 0.1, solver=solver, max_iter=6, tol=0.0, numItermax=60, stopThr=0.0)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        Tr, lr = sorc.entropic_gromov_wasserstein(C1, C2, log=True, **kw)
        T, lg = ot.gromov.entropic_gromov_wasserstein(*to_dev(C1, C2), log=True, **kw)
    assert_plan(T, Tr, 1e-10)
    assert_scalar(lg["gw_dist"], lr["gw_dist"], 1e-9)
    assert abs(float(T.sum()) - 1) < 1e-6


def test_errors():
    C = torch.rand(6, 6, device=DEV, dtype=torch.float64)
    with pytest.raises(ValueError, match="Unknown solver"):
        ot.gromov.entropic_gromov_wasserstein(C, C, solver="BAPG")
    with pytest.raises(ValueError, match="Unknown `loss_fun"):
        ot.gromov.entropic_gromov_wasserstein(C, C, loss_fun="l1")
    with pytest.raises(TypeError):
        ot.gromov.entropic_gromov_wasserstein(C.cpu().numpy(), C.cpu().numpy())
    with pytest.raises(ValueError, match="CUDA"):
        ot.gromov.entropic_gromov_wasserstein(C.cpu(), C.cpu())
