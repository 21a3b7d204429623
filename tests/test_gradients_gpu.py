"""Gradients of the loss-returning entry points against goldens from the reference's torch
backend (tests/golden/golden_gradients.npz, golden_solvers.npz): sinkhorn2, empirical_sinkhorn2 (dense and lazy),
empirical_sinkhorn_divergence, solve_sample, the differentiable dist; plus the port of the reference's
test_sinkhorn2_gradients (test/test_bregman.py:203-225) and the warnings of the plan-returning functions.

The kernels are forward-only: the loss carries the envelope gradient of ot/solvers.py:485-494 (pot_b200/_autograd.py).
Tolerances: 1e-8 relative in fp64 (a gradient chains two kernels and a matmul), 2e-4 in fp32."""
import os
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot

DEV = "cuda:0"
Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_gradients.npz"))
CASES = sorted({k.split("/")[0] for k in Z.keys()})


def dev(x, grad=False):
    t = torch.as_tensor(np.ascontiguousarray(x), device=DEV)
    return t.requires_grad_(True) if grad else t


def close(x, ref, tol, what):
    x, ref = x.detach().cpu().numpy().astype(np.float64), np.asarray(ref, dtype=np.float64)
    err = np.abs(x - ref).max() / max(np.abs(ref).max(), 1e-300)
    assert err < tol, f"{what}: err {err:.3e} >= {tol}"


def tol_of(case):
    return 1e-8 if "_f64_" in case else 2e-4


def test_sinkhorn2_gradients():
    """Port of test/test_bregman.py:203-225."""
    n_samples, n_features = 100, 2
    rng = np.random.RandomState(0)
    x, y = rng.randn(n_samples, n_features), rng.randn(n_samples, n_features)
    a = np.full(n_samples, 1.0 / n_samples)
    M = ot.dist(dev(x), dev(y))
    a1, b1, M1 = dev(a, True), dev(a, True), M.clone().requires_grad_(True)
    with pytest.warns(UserWarning, match="envelope gradient"):
        val = ot.sinkhorn2(a1, b1, M1, 1)
    val.backward()
    assert a1.shape == a1.grad.shape and b1.shape == b1.grad.shape and M1.shape == M1.grad.shape
    assert torch.isfinite(M1.grad).all()


@pytest.mark.parametrize("case", [f"solve_{t}_{s}" for t in ("f64", "f32") for s in ("40x52", "128x96")])
@pytest.mark.parametrize("method", ["sinkhorn_log", "sinkhorn"])
def test_sinkhorn2_envelope_matches_reference_solve(gs, case, method):
    """The gradient attached to sinkhorn2 is the one the reference's ot.solve(grad="envelope") prescribes for the same
    inputs and iteration count (golden_solvers.npz); the scaling-domain method reaches the same potentials."""
    M, a, b = (dev(gs(case, k), True) for k in ("M", "a", "b"))
    reg, its = float(gs(case, "reg")), int(gs(case, "its"))
    tol = 1e-8 if "f64" in case else 2e-4
    if method == "sinkhorn":
        tol = max(tol, 1e-6)  # v-then-u in the scaling domain vs the log domain: same iterates up to round-off
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        val = ot.sinkhorn2(a, b, M, reg, method=method, numItermax=its, stopThr=0)
    assert val.requires_grad
    (2.0 * val).backward()
    c = f"{case}/KL_envelope"
    close(M.grad, 2.0 * gs(c, "gM"), tol, "dM")
    close(a.grad, 2.0 * gs(c, "ga"), tol, "da")
    close(b.grad, 2.0 * gs(c, "gb"), tol, "db")
    # no grad requested: no graph, same value
    with torch.no_grad():
        v0 = ot.sinkhorn2(a, b, M, reg, method=method, numItermax=its, stopThr=0)
    assert not v0.requires_grad and abs(float(v0) - float(val)) <= 1e-6 * abs(float(val))


def test_sinkhorn2_multi_histogram_gradients(gs):
    case = "solve_f64_40x52"
    M0, a0, b0 = gs(case, "M"), gs(case, "a"), gs(case, "b")
    reg, its = float(gs(case, "reg")), int(gs(case, "its"))
    b2 = np.stack([b0, b0[::-1].copy()], 1)
    M, a, b = dev(M0, True), dev(a0, True), dev(b2, True)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        losses = ot.sinkhorn2(a, b, M, reg, method="sinkhorn_log", numItermax=its, stopThr=0)
    assert losses.shape == (2,)
    losses[0].backward()
    c = f"{case}/KL_envelope"
    close(M.grad, gs(c, "gM"), 1e-8, "dM (histogram 0)")
    close(b.grad[:, 0], gs(c, "gb"), 1e-8, "db (histogram 0)")
    assert float(b.grad[:, 1].abs().max()) == 0.0


def test_sinkhorn2_joint_histograms_gradients_scaling_domain(gs):
    """2-D b with method="sinkhorn": the histograms iterate jointly on one kernel (b2ot_multi_scaling_step); each loss
    carries the envelope gradient of its own histogram."""
    case = "solve_f64_40x52"
    M0, a0, b0 = gs(case, "M"), gs(case, "a"), gs(case, "b")
    reg, its = float(gs(case, "reg")), int(gs(case, "its"))
    b2 = np.stack([b0[::-1].copy(), b0], 1)
    M, a, b = dev(M0, True), dev(a0, True), dev(b2, True)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        losses = ot.sinkhorn2(a, b, M, reg, method="sinkhorn", numItermax=its, stopThr=0)
    assert losses.shape == (2,) and losses.requires_grad
    losses[1].backward()
    c = f"{case}/KL_envelope"
    close(M.grad, gs(c, "gM"), 1e-6, "dM (histogram 1)")
    close(b.grad[:, 1], gs(c, "gb"), 1e-6, "db (histogram 1)")
    close(a.grad, gs(c, "ga"), 1e-6, "da (histogram 1)")
    assert float(b.grad[:, 0].abs().max()) == 0.0


@pytest.mark.parametrize("case", CASES)
def test_empirical_sinkhorn2_dense_gradients(case):
    """X, Y -> dist kernel (differentiable) -> sinkhorn2 envelope: against the reference's
    ot.solve(ot.dist(X, Y), grad="envelope") chain on torch."""
    same = case.endswith("_same")
    metric = "euclidean" if "_euclidean" in case else "sqeuclidean"
    X = dev(Z[f"{case}/X"], True)
    Y = X if same else dev(Z[f"{case}/Y"], True)
    a, b = dev(Z[f"{case}/a"], True), dev(Z[f"{case}/b"], True)
    reg, its = float(Z[f"{case}/reg"]), int(Z[f"{case}/its"])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        val = ot.empirical_sinkhorn2(X, Y, reg, a, b, metric=metric, numIterMax=its, stopThr=0, isLazy=False,
                                     method="sinkhorn_log")
    val.backward()
    tol = tol_of(case)
    close(X.grad, Z[f"{case}/dense/gX"], tol, "dX")
    if not same:
        close(Y.grad, Z[f"{case}/dense/gY"], tol, "dY")
    close(a.grad, Z[f"{case}/dense/ga"], tol, "da")
    close(b.grad, Z[f"{case}/dense/gb"], tol, "db")


@pytest.mark.parametrize("case", CASES)
def test_empirical_sinkhorn2_lazy_gradients(case):
    """Lazy path: the plan is rebuilt block by block in backward; against the envelope gradient formed from the
    reference's own lazy potentials."""
    same = case.endswith("_same")
    metric = "euclidean" if "_euclidean" in case else "sqeuclidean"
    X = dev(Z[f"{case}/X"], True)
    Y = X if same else dev(Z[f"{case}/Y"], True)
    a, b = dev(Z[f"{case}/a"], True), dev(Z[f"{case}/b"], True)
    reg, its = float(Z[f"{case}/reg"]), int(Z[f"{case}/its"])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        val, lg = ot.empirical_sinkhorn2(X, Y, reg, a, b, metric=metric, numIterMax=its, stopThr=-1, isLazy=True,
                                         batchSize=16, log=True)
    tol = tol_of(case)
    close(lg["u"], Z[f"{case}/lazy/f"], tol, "f")
    assert abs(float(val) - float(Z[f"{case}/lazy/loss"])) <= tol * abs(float(Z[f"{case}/lazy/loss"]))
    val.backward()
    close(X.grad, Z[f"{case}/lazy/gX"], tol, "dX")
    if not same:
        close(Y.grad, Z[f"{case}/lazy/gY"], tol, "dY")
    close(a.grad, Z[f"{case}/lazy/ga"], tol, "da")
    close(b.grad, Z[f"{case}/lazy/gb"], tol, "db")


def test_divergence_and_solve_sample_are_differentiable():
    case = "pc_f64_sqeuclidean"
    X, Y = dev(Z[f"{case}/X"], True), dev(Z[f"{case}/Y"], True)
    reg = float(Z[f"{case}/reg"])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for lazy in (False, True):
            X.grad = Y.grad = None
            div = ot.empirical_sinkhorn_divergence(X, Y, reg, numIterMax=50, stopThr=-1, isLazy=lazy)
            div.backward()
            assert X.grad.shape == X.shape and torch.isfinite(X.grad).all() and float(X.grad.abs().max()) > 0
            assert Y.grad.shape == Y.shape and torch.isfinite(Y.grad).all()
        X.grad = Y.grad = None
        r = ot.solve_sample(X, Y, reg=reg, max_iter=50, tol=0, grad="envelope")
        r.value.backward()
        assert float(X.grad.abs().max()) > 0
        X.grad = None
        r = ot.solve_sample(X, Y, reg=reg, max_iter=50, tol=0, lazy=True)
        r.value_linear.backward()
        assert float(X.grad.abs().max()) > 0


def test_dist_gradient_matches_torch():
    rng = np.random.RandomState(1)
    for metric in ("sqeuclidean", "euclidean"):
        X, Y = dev(rng.randn(33, 5), True), dev(rng.randn(21, 5), True)
        W = dev(rng.randn(33, 21))
        (ot.dist(X, Y, metric=metric) * W).sum().backward()
        Xt, Yt = X.detach().clone().requires_grad_(True), Y.detach().clone().requires_grad_(True)
        C = torch.cdist(Xt, Yt) if metric == "euclidean" else torch.cdist(Xt, Yt) ** 2
        (C * W).sum().backward()
        close(X.grad, Xt.grad.cpu().numpy(), 1e-9, f"dX {metric}")
        close(Y.grad, Yt.grad.cpu().numpy(), 1e-9, f"dY {metric}")


def test_plan_returning_functions_warn_instead_of_silently_detaching():
    rng = np.random.RandomState(0)
    M = dev(rng.rand(20, 30), True)
    a, b = dev(np.full(20, 1 / 20)), dev(np.full(30, 1 / 30))
    for fn in (ot.sinkhorn, ot.sinkhorn_knopp, ot.sinkhorn_log):
        with pytest.warns(UserWarning, match="forward-only"):
            G = fn(a, b, M, 0.5, numItermax=20, stopThr=0, warn=False)
        assert not G.requires_grad
    X = dev(rng.randn(20, 3), True)
    with pytest.warns(UserWarning, match="forward-only"):
        ot.empirical_sinkhorn(X, X.detach() + 1, 0.5, numIterMax=10, stopThr=-1, isLazy=True, warn=False)
    with warnings.catch_warnings():  # nothing requires grad: no warning
        warnings.simplefilter("error")
        ot.sinkhorn(a, b, M.detach(), 0.5, numItermax=20, stopThr=0, warn=False)
