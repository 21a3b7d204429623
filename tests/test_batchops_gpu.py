"""GPU parity of the batched helpers of ot.batch (dist_batch, loss_linear_batch, loss_linear_samples_batch,
entropy_batch, solve_sample_batch) against the reference's outputs (golden_batchops.npz) and the numpy oracle.
Ports test/batch/test_solve_batch.py::test_metrics / test_gradients_torch of the reference.

Tolerances: 1e-10 relative in fp64, 1e-4 in fp32 (north_star)."""
import os
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot
    from pot_b200.batch import (dist_batch, entropy_batch, loss_linear_batch, loss_linear_samples_batch,
                                solve_batch, solve_sample_batch)
from oracle import sinkhorn_oracle as orc

DEV = "cuda:0"
Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_batchops.npz"))
CASES = sorted({k.rsplit("/", 1)[0] for k in Z.keys()})


def dev(x):
    return torch.as_tensor(np.ascontiguousarray(x), device=DEV)


def npf(x):
    return x.detach().cpu().numpy().astype(np.float64)


def close(x, ref, tol, what):
    x, ref = npf(x), np.asarray(ref, dtype=np.float64)
    err = np.abs(x - ref).max() / max(np.abs(ref).max(), 1e-300)
    assert err < tol, f"{what}: err {err:.3e} >= {tol}"


@pytest.mark.parametrize("case", CASES)
def test_batchops_golden(case):
    X, Y, p, T = Z[f"{case}/X"], Z[f"{case}/Y"], int(Z[f"{case}/p"]), Z[f"{case}/T"]
    metric = case.split("_")[1]
    tol = 1e-10 if case.startswith("f64") else 1e-4
    M = dist_batch(dev(X), dev(Y), metric=metric, p=p)
    assert M.dtype == dev(X).dtype and M.shape == Z[f"{case}/M"].shape
    close(M, Z[f"{case}/M"], tol, "dist_batch")
    close(loss_linear_batch(dev(Z[f"{case}/M"]), dev(T)), Z[f"{case}/loss"], tol, "loss_linear_batch")
    close(entropy_batch(dev(T)), Z[f"{case}/ent"], tol, "entropy_batch")
    if f"{case}/loss_samples" in Z:
        close(loss_linear_samples_batch(dev(X), dev(Y), dev(T), metric=metric), Z[f"{case}/loss_samples"], tol,
              "loss_linear_samples_batch")
    res = solve_sample_batch(dev(X), dev(Y), 0.1, metric=metric, p=p, max_iter=20, tol=0)
    close(res.plan, Z[f"{case}/ssb_plan"], 10 * tol, "solve_sample_batch plan")
    close(res.value, Z[f"{case}/ssb_value"], 10 * tol, "solve_sample_batch value")
    close(res.value_linear, Z[f"{case}/ssb_value_linear"], 10 * tol, "solve_sample_batch value_linear")


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(1, 1, 1, 1), (5, 64, 64, 2), (3, 129, 67, 17), (2, 300, 513, 40), (64, 256, 256, 2)])
@pytest.mark.parametrize("metric", ["sqeuclidean", "euclidean", "minkowski", "kl"])
def test_dist_batch_vs_oracle(dtype, shape, metric):
    """Ragged tiles (n, m not multiples of 64), several coordinate chunks, unaligned rows, config 5's shape."""
    B, n, m, d = shape
    rng = np.random.RandomState(n + m)
    X, Y = rng.rand(B, n, d).astype(dtype), (rng.rand(B, m, d) + 0.5).astype(dtype)
    if metric == "kl":
        X, Y = X / X.sum(-1, keepdims=True), Y / Y.sum(-1, keepdims=True)
    p = 3 if metric == "minkowski" else 2
    ref = orc.dist_batch(X.astype(np.float64), Y.astype(np.float64), metric, p)
    M = dist_batch(dev(X), dev(Y), metric=metric, p=p)
    tol = 1e-10 if dtype == np.float64 else 1e-4
    close(M, ref, tol, f"dist_batch {metric}")
    # X2 = None means X2 = X1 (:218); no diagonal zeroing and no clamp in the batched version
    M11 = dist_batch(dev(X), metric=metric, p=p)
    ref11 = orc.dist_batch(X.astype(np.float64), None, metric, p)
    if metric == "euclidean":  # the diagonal is sqrt(round-off): nan or tiny in the reference too; compare off-diagonal
        mask = ~np.eye(n, dtype=bool)[None].repeat(B, 0)
        a, r = npf(M11)[mask], ref11[mask]
        assert np.abs(a - r).max() <= tol * max(np.abs(r).max(), 1e-300) if a.size else True
    else:
        close(M11, ref11, tol, f"dist_batch {metric} (X2 None)")


def test_dist_batch_errors():
    X = torch.rand(2, 4, 3, device=DEV)
    with pytest.raises(ValueError, match="Unknown metric"):
        dist_batch(X, X, metric="l2")
    with pytest.raises(ValueError, match="Unknown metric"):
        loss_linear_samples_batch(X, X, torch.rand(2, 4, 4, device=DEV))  # the reference's default metric "l2"
    with pytest.raises(ValueError):
        dist_batch(X, torch.rand(3, 4, 3, device=DEV))
    with pytest.raises(TypeError):
        dist_batch(np.zeros((2, 4, 3)))
    with pytest.raises(TypeError):
        solve_batch(torch.rand(2, 4, 4, device=DEV), 0.1, foo=1)  # no extra keywords: the reference's signature


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(1, 1, 1), (7, 33, 65), (3, 1000, 999), (4096, 16, 16), (2, 2048, 2048)])
def test_reductions_vs_oracle(dtype, shape):
    """Split planes, unaligned planes (per % 4 != 0), many small problems."""
    B, n, m = shape
    rng = np.random.RandomState(B + n)
    M = rng.rand(B, n, m).astype(dtype)
    T = rng.rand(B, n, m).astype(dtype)
    T /= T.sum(axis=(1, 2), keepdims=True)
    T[0, 0, 0] = 0  # log(0 + eps)
    tol = 1e-10 if dtype == np.float64 else 1e-4
    close(loss_linear_batch(dev(M), dev(T)), orc.loss_linear_batch(M.astype(np.float64), T.astype(np.float64)), tol,
          "loss_linear_batch")
    close(entropy_batch(dev(T)), orc.entropy_batch(T.astype(np.float64)), tol, "entropy_batch")
    close(entropy_batch(dev(T), eps=1e-3), orc.entropy_batch(T.astype(np.float64), 1e-3), tol, "entropy_batch eps")
    # non-contiguous views are accepted
    Mt = dev(np.ascontiguousarray(M.transpose(0, 2, 1))).transpose(1, 2)
    close(loss_linear_batch(Mt, dev(T)), orc.loss_linear_batch(M.astype(np.float64), T.astype(np.float64)), tol,
          "loss_linear_batch (strided)")
    # deterministic
    assert torch.equal(entropy_batch(dev(T)), entropy_batch(dev(T)))


@pytest.mark.parametrize("metric", ["sqeuclidean", "euclidean", "minkowski", "kl"])
def test_metrics(metric):
    """Port of test/batch/test_solve_batch.py:82-112."""
    batchsize, n, d = 2, 4, 2
    rng = np.random.RandomState(0)
    X = rng.rand(batchsize, n, d)
    if metric == "kl":
        X = np.abs(X) + 1e-6
        X = X / np.sum(X, axis=-1, keepdims=True)
    X = dev(X)
    M = dist_batch(X, X, metric=metric)
    if metric != "euclidean":  # sqrt of the un-clamped round-off on the diagonal is nan in the reference as well
        assert bool((M >= -1e-12).all())
    M = torch.nan_to_num(M, nan=0.0).clamp_(min=0)
    res = solve_batch(M, reg=0.1, max_iter=10, tol=1e-5)
    assert torch.isfinite(res.plan).all()
    if metric != "euclidean":
        res = solve_sample_batch(X, X, reg=0.1, max_iter=10, tol=1e-5, metric=metric)
        loss = res.value_linear
        Ms = dist_batch(X, X, metric=metric)
        np.testing.assert_allclose(npf(loss), npf(loss_linear_batch(Ms, res.plan)), atol=1e-5)
        np.testing.assert_allclose(npf(loss), npf(loss_linear_samples_batch(X, X, res.plan, metric=metric)), atol=1e-5)


@pytest.mark.parametrize("grad", ["detach", "envelope", "autodiff", "last_step"])
def test_gradients_torch(grad):
    """Port of test/batch/test_solve_batch.py:114-134: every gradient mode runs and reaches the samples."""
    batchsize, n, d = 2, 4, 2
    for solver in ["sinkhorn", "log_sinkhorn"]:
        X = torch.randn((batchsize, n, d), requires_grad=True, device=DEV, dtype=torch.float64)
        M = dist_batch(X, X)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            res = solve_batch(M, reg=0.1, max_iter=10, tol=1e-5, grad=grad, solver=solver)
        loss = res.value_linear.sum()
        loss_plan = res.plan.sum()
        if grad == "detach":
            assert not loss.requires_grad and not res.plan.requires_grad
        elif grad == "envelope":
            loss.backward()
            assert X.grad is not None and torch.isfinite(X.grad).all()
            assert not res.plan.requires_grad
        else:
            loss_plan.backward()
            assert X.grad is not None and torch.isfinite(X.grad).all()


@pytest.mark.parametrize("solver", ["sinkhorn", "log_sinkhorn"])
def test_last_step_gradient_matches_torch_restatement(solver):
    """grad="last_step": the plan / value graph is the reference's last differentiable step (ot/batch/_utils.py:151-155,
    :279-283) on top of the device iterations: compare d value / dM with a pure-torch restatement of the same loop."""
    rng = np.random.RandomState(3)
    B, n, m, reg, its = 3, 12, 9, 0.2, 7
    M0 = rng.rand(B, n, m)
    a = np.full((B, n), 1.0 / n)
    b = np.full((B, m), 1.0 / m)
    M = dev(M0).requires_grad_(True)
    res = solve_batch(M, reg, max_iter=its, tol=0, solver=solver, grad="last_step")
    res.value.sum().backward()
    g = M.grad.clone()
    Mt = torch.tensor(M0, requires_grad=True)
    at, bt = torch.tensor(a), torch.tensor(b)
    if solver == "log_sinkhorn":
        K_, K = -Mt / reg, (-Mt / reg).detach()
        u, v = torch.zeros(B, n, dtype=torch.float64), torch.zeros(B, m, dtype=torch.float64)
        for _ in range(its):
            u = torch.log(at) - torch.logsumexp(K + v[:, None, :], dim=2)
            v = torch.log(bt) - torch.logsumexp(K + u[:, :, None], dim=1)
        u = torch.log(at) - torch.logsumexp(K_ + v[:, None, :], dim=2)
        v = torch.log(bt) - torch.logsumexp(K_ + u[:, :, None], dim=1)
        T = torch.exp(K + u[:, :, None] + v[:, None, :])
    else:
        K_, K = torch.exp(-Mt / reg), torch.exp(-Mt / reg).detach()
        f, gg = torch.ones(B, n, dtype=torch.float64), torch.ones(B, m, dtype=torch.float64)
        for _ in range(its):
            f = at / torch.sum(K * gg[:, None, :], dim=2)
            gg = bt / torch.sum(K * f[:, :, None], dim=1)
        f = at / torch.sum(K_ * gg[:, None, :], dim=2)
        gg = bt / torch.sum(K_ * f[:, :, None], dim=1)
        T = K * f[:, :, None] * gg[:, None, :]
    val = torch.sum(Mt * T, dim=(1, 2)) + reg * torch.sum(T * torch.log(T + 1e-16), dim=(1, 2))
    val.sum().backward()
    close(res.plan, T.detach().numpy(), 1e-10, "last-step plan")
    close(g, Mt.grad.numpy(), 1e-9, "d value / dM (last_step)")


def test_reduction_gradients():
    rng = np.random.RandomState(0)
    M = dev(rng.rand(2, 5, 6)).requires_grad_(True)
    T = dev(rng.rand(2, 5, 6)).requires_grad_(True)
    w = dev(rng.rand(2))
    (loss_linear_batch(M, T) * w).sum().backward()
    close(M.grad, npf(w)[:, None, None] * npf(T), 1e-12, "d loss / dM")
    close(T.grad, npf(w)[:, None, None] * npf(M), 1e-12, "d loss / dT")
    T2 = dev(rng.rand(2, 5, 6)).requires_grad_(True)
    (entropy_batch(T2) * w).sum().backward()
    T3 = torch.tensor(npf(T2), requires_grad=True)
    (-(T3 * torch.log(T3 + 1e-16)).sum(dim=(1, 2)) * torch.tensor(npf(w))).sum().backward()
    close(T2.grad, T3.grad.numpy(), 1e-12, "d entropy / dT")
