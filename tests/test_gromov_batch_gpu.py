"""GPU parity tests of ot.batch.solve_gromov_batch against the golden vectors of the
reference (tests/golden/golden_gromov_batch.npz) and the CPU oracle.  Every inner projection runs in libb2ot.so.

Fixed work (tol = tol_inner = 0); 1e-10 relative in fp64, 1e-4 in fp32 (north_star).
"""
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot
from oracle import solvers_oracle as sorc
from test_oracle_gromov_batch import CASES, FIX, gwb_kwargs
from test_solvers_gpu import npf, to_dev

DEV = "cuda:0"


def close(x, ref, tol):
    x, ref = npf(x), np.asarray(ref, dtype=np.float64)
    np.testing.assert_allclose(x, ref, rtol=tol, atol=tol * max(np.abs(ref).max(), 1e-300))


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("grad", ["detach", "envelope"])
def test_solve_gromov_batch_golden(ggwb, name, grad):
    case = f"gwb_{name}"
    kw = gwb_kwargs(ggwb, case)
    C1, C2, a, b = to_dev(*(ggwb(case, k) for k in ("C1", "C2", "a", "b")))
    if "alpha" in kw:
        kw["M"] = to_dev(ggwb(case, "M"))[0]
    res = ot.batch.solve_gromov_batch(C1, C2, a=a, b=b, grad=grad, **kw, **FIX)
    tol = 1e-10 if "f64" in name else 1e-4
    assert res.log["n_iter"] == int(ggwb(case, "n_iter"))
    assert res.plan.dtype == C1.dtype and res.plan.device == C1.device
    close(res.plan, ggwb(case, "plan"), tol)
    close(res.value, ggwb(case, "value"), tol * 10)
    close(res.value_linear, ggwb(case, "value_linear"), tol * 10)
    close(res.value_quad, ggwb(case, "value_quad"), tol * 10)
    # row-marginal violation of the returned plans next to the reference's (the last half-step fixes the columns)
    viol = float((res.plan.sum(2) - a).abs().max())
    ref_viol = float(np.abs(ggwb(case, "plan").sum(2) - ggwb(case, "a")).max())
    assert abs(viol - ref_viol) < (1e-10 if "f64" in name else 1e-5)


def test_default_weights_converging(ggwb):
    case = "gwb_default_conv_f64"
    C1, C2 = to_dev(ggwb(case, "C1"), ggwb(case, "C2"))
    res = ot.batch.solve_gromov_batch(C1, C2, reg=0.1, grad="detach", max_iter=50, tol=1e-6, max_iter_inner=50,
                                      tol_inner=1e-7)
    assert res.log["n_iter"] == int(ggwb(case, "n_iter"))
    close(res.plan, ggwb(case, "plan"), 1e-7)
    close(res.value, ggwb(case, "value"), 1e-7)


def test_envelope_gradients_match_torch_autograd_of_the_value():
    """grad='envelope': value is differentiable wrt (C1, C2, M) with the plan held fixed (:472-483)."""
    rng = np.random.RandomState(0)
    B, n = 2, 12
    C1 = torch.as_tensor(rng.rand(B, n, n), device=DEV).requires_grad_(True)
    C2 = torch.as_tensor(rng.rand(B, n, n), device=DEV).requires_grad_(True)
    M = torch.as_tensor(rng.rand(B, n, n), device=DEV).requires_grad_(True)
    res = ot.batch.solve_gromov_batch(C1, C2, reg=0.1, M=M, alpha=0.5, symmetric=False, max_iter=5, max_iter_inner=30,
                                      **FIX)
    res.value.sum().backward()
    T = res.plan.detach()
    assert not res.plan.requires_grad
    torch.testing.assert_close(M.grad, 0.5 * T, rtol=1e-12, atol=1e-14)
    assert C1.grad is not None and C2.grad is not None and torch.isfinite(C1.grad).all()
    # the same value rebuilt from the public helpers
    a = torch.full((B, n), 1.0 / n, device=DEV, dtype=torch.float64)
    q = ot.batch.loss_quadratic_samples_batch(a, a, C1.detach(), C2.detach(), T, symmetric=False, recompute_const=True)
    torch.testing.assert_close(res.value_quad.detach(), 0.5 * q, rtol=1e-12, atol=1e-14)


def test_vs_oracle_larger_batch():
    """B = 64 problems of 64 x 48: the cluster-of-CTAs on-chip solver."""
    rng = np.random.RandomState(9)
    B, n, m = 64, 64, 48
    xs, xt = rng.randn(B, n, 2), rng.randn(B, m, 2)
    C1 = ((xs[:, :, None] - xs[:, None]) ** 2).sum(-1)
    C2 = ((xt[:, :, None] - xt[:, None]) ** 2).sum(-1)
    C1, C2 = C1 / C1.max(), C2 / C2.max()
    kw = dict(reg=0.05, max_iter=4, max_iter_inner=30, **FIX)
    r = sorc.solve_gromov_batch(C1, C2, **kw)
    res = ot.batch.solve_gromov_batch(*to_dev(C1, C2), grad="detach", **kw)
    close(res.plan, r["plan"], 1e-10)
    close(res.value, r["value"], 1e-9)


def test_errors():
    C = torch.rand(2, 6, 6, device=DEV, dtype=torch.float64)
    M = torch.rand(2, 6, 6, device=DEV, dtype=torch.float64)
    with pytest.raises(ValueError, match="alpha must also be provided"):
        ot.batch.solve_gromov_batch(C, C, M=M)
    with pytest.raises(ValueError, match="M must also be provided"):
        ot.batch.solve_gromov_batch(C, C, alpha=0.5)
    with pytest.raises(ValueError, match="Alpha must be in"):
        ot.batch.solve_gromov_batch(C, C, M=M, alpha=1.5)
    with pytest.raises(ValueError, match="does not match"):
        ot.batch.solve_gromov_batch(C, C, M=M[:, :5], alpha=0.5)
    with pytest.raises(ValueError, match="Unknown loss"):
        ot.batch.solve_gromov_batch(C, C, loss="l1")
    with pytest.raises(TypeError):
        ot.batch.solve_gromov_batch(C.cpu().numpy(), C.cpu().numpy())
    with pytest.warns(UserWarning, match="forward-only"):
        ot.batch.solve_gromov_batch(C, C, grad="autodiff", max_iter=2, max_iter_inner=5)
