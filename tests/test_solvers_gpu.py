"""GPU parity tests of ot.solve / ot.solve_sample / empirical_sinkhorn_divergence
against the golden vectors of the reference (tests/golden/golden_solvers.npz) and the CPU oracle.
Everything goes through libb2ot.so (C ABI) via pot_b200.

Tolerances (north_star): 1e-10 relative in fp64, 1e-4 relative in fp32, fixed iteration count.
"""
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot
from oracle import solvers_oracle as sorc

DEV = "cuda:0"
SOLVE_CASES = [f"solve_{t}_{s}" for t in ("f64", "f32") for s in ("40x52", "128x96")]


def tol_of(case):
    return 1e-10 if "f64" in case else 1e-4


def to_dev(*xs):
    return [torch.as_tensor(np.ascontiguousarray(x), device=DEV) for x in xs]


def npf(x):
    return x.detach().cpu().numpy().astype(np.float64)


def assert_plan(G, Gref, tol):
    G, Gref = npf(G), np.asarray(Gref, dtype=np.float64)
    err = np.abs(G - Gref) / (np.abs(Gref) + 1e-6 * Gref.max())
    assert err.max() < tol, f"plan rel err {err.max():.3e} >= {tol}"


def assert_vec(x, ref, tol):
    x, ref = npf(x), np.asarray(ref, dtype=np.float64)
    err = np.abs(x - ref).max() / max(1.0, np.abs(ref).max())
    assert err < tol, f"vector err {err:.3e} >= {tol}"


def assert_scalar(x, ref, tol, scale=None):
    x, ref = float(x), float(ref)
    s = abs(ref) if scale is None else scale
    assert abs(x - ref) <= tol * s, f"scalar {x} vs {ref} (tol {tol}, scale {s})"


@pytest.mark.parametrize("case", SOLVE_CASES)
@pytest.mark.parametrize("rt", ["KL", "entropy"])
def test_solve_balanced_golden(gs, case, rt):
    M, a, b = to_dev(gs(case, "M"), gs(case, "a"), gs(case, "b"))
    reg, its, tol = float(gs(case, "reg")), int(gs(case, "its")), tol_of(case)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r = ot.solve(M, a, b, reg=reg, reg_type=rt, max_iter=its, tol=0)
    c = f"{case}/{rt}"
    assert r.plan.dtype == M.dtype and r.plan.device == M.device
    assert_plan(r.plan, gs(c, "plan"), tol)
    assert_vec(r.potentials[0], gs(c, "f"), tol)
    assert_vec(r.potentials[1], gs(c, "g"), tol)
    lin = float(gs(c, "value_linear"))
    assert_scalar(r.value_linear, lin, tol)
    # the regularisation term is O(reg * log n) next to value_linear: compare on that scale
    assert_scalar(r.value, gs(c, "value"), tol, scale=max(abs(lin), abs(float(gs(c, "value")))))
    assert r.value.dtype == M.dtype


@pytest.mark.parametrize("case", SOLVE_CASES)
@pytest.mark.parametrize("rt", ["KL", "entropy"])
def test_solve_envelope_gradients(gs, case, rt):
    M, a, b = to_dev(gs(case, "M"), gs(case, "a"), gs(case, "b"))
    for t in (M, a, b):
        t.requires_grad_(True)
    reg, its, tol = float(gs(case, "reg")), int(gs(case, "its")), tol_of(case)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r = ot.solve(M, a, b, reg=reg, reg_type=rt, max_iter=its, tol=0, grad="envelope")
    r.value.backward()
    c = f"{case}/{rt}_envelope"
    assert M.grad.shape == M.shape and a.grad.shape == a.shape and b.grad.shape == b.shape
    assert_plan(M.grad, gs(c, "gM"), tol)
    assert_vec(a.grad, gs(c, "ga"), tol)
    assert_vec(b.grad, gs(c, "gb"), tol)
    # detach: no graph
    r = ot.solve(M, a, b, reg=reg, reg_type=rt, max_iter=its, tol=0, grad="detach")
    with pytest.raises(RuntimeError):
        r.value.backward()


@pytest.mark.parametrize("case", SOLVE_CASES[:1] + SOLVE_CASES[2:3])
def test_solve_autodiff_substitutes_envelope(gs, case):
    M, a, b = to_dev(gs(case, "M"), gs(case, "a"), gs(case, "b"))
    M.requires_grad_(True)
    reg, its, tol = float(gs(case, "reg")), int(gs(case, "its")), tol_of(case)
    with pytest.warns(UserWarning, match="forward-only"):
        r = ot.solve(M, a, b, reg=reg, max_iter=its, tol=0)
    r.value.backward()
    assert_plan(M.grad, gs(f"{case}/KL_envelope", "gM"), tol)
    assert a.grad is None


@pytest.mark.parametrize("case", SOLVE_CASES)
@pytest.mark.parametrize("rt", ["KL", "entropy"])
def test_solve_last_step(gs, case, rt):
    M, a, b = to_dev(gs(case, "M"), gs(case, "a"), gs(case, "b"))
    reg, its, tol = float(gs(case, "reg")), int(gs(case, "its")), tol_of(case)
    c = f"{case}/{rt}_last_step"
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r = ot.solve(M, a, b, reg=reg, reg_type=rt, max_iter=its, tol=0, grad="last_step")
        assert_plan(r.plan, gs(c, "plan"), tol)
        assert_vec(r.potentials[0], gs(c, "f"), tol)
        lin = float(gs(c, "value_linear"))
        assert_scalar(r.value, gs(c, "value"), tol, scale=max(abs(lin), abs(float(gs(c, "value")))))
        # differentiable variant: same numbers, gradients flow through the last step only
        for t in (M, a, b):
            t.requires_grad_(True)
        r = ot.solve(M, a, b, reg=reg, reg_type=rt, max_iter=its, tol=0, grad="last_step")
    assert_plan(r.plan, gs(c, "plan"), tol)
    assert_scalar(r.value, gs(c, "value"), tol, scale=max(abs(lin), abs(float(gs(c, "value")))))
    r.value.backward()
    assert M.grad.shape == M.shape and torch.isfinite(M.grad).all()
    assert a.grad.shape == a.shape and b.grad.shape == b.shape
    with pytest.raises(ValueError):
        ot.solve(M, a, b, grad="last_step", max_iter=0, reg=reg)


@pytest.mark.parametrize("case", SOLVE_CASES)
def test_solve_default_weights_converged(gs, case):
    (M,) = to_dev(gs(case, "M"))
    reg = float(gs(case, "reg"))
    r = ot.solve(M, reg=reg, tol=1e-9 if "f64" in case else 1e-6, max_iter=3000)
    c = f"{case}/default"
    t = 1e-7 if "f64" in case else 1e-4  # converged runs: the stop iteration may differ by one check
    assert_plan(r.plan, gs(c, "plan"), t)
    assert_scalar(r.value_linear, gs(c, "value_linear"), t)
    ma, mb = r.marginals
    assert_vec(ma, np.full(M.shape[0], 1.0 / M.shape[0]), 1e-5)
    assert_vec(mb, np.full(M.shape[1], 1.0 / M.shape[1]), 1e-5)
    r.sparse_plan, r.status, r.potentials  # attributes the reference test touches (test_solvers.py:125-128)


def test_solve_errors():
    M = torch.rand(8, 6, device=DEV, dtype=torch.float64)
    with pytest.raises(NotImplementedError):
        ot.solve(M, reg=1, reg_type="cryptic divergence")  # test_solvers.py:143-144
    with pytest.raises(NotImplementedError):
        ot.solve(M, reg=1, unbalanced=1, unbalanced_type="cryptic divergence")  # :139-140
    with pytest.raises(NotImplementedError):
        ot.solve(M)  # exact OT: outside the path
    with pytest.raises((TypeError, ValueError)):
        ot.solve(M.cpu(), reg=1.0)
    with pytest.raises(TypeError):
        ot.solve(M.cpu().numpy(), reg=1.0)


@pytest.mark.parametrize("n,m,dtype", [(1024, 2048, torch.float32), (777, 1031, torch.float32),
                                       (512, 640, torch.float64), (301, 1001, torch.float64)])
@pytest.mark.parametrize("rt", ["KL", "entropy"])
def test_solve_vs_oracle_larger(n, m, dtype, rt):
    """Aligned (vector path) and ragged (scalar path) shapes against the oracle on seeded inputs."""
    rng = np.random.RandomState(n + m)
    npd = np.float32 if dtype == torch.float32 else np.float64
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = ((xs ** 2).sum(1)[:, None] + (xt ** 2).sum(1)[None, :] - 2 * xs @ xt.T).clip(0)
    M = (M / M.max()).astype(npd)
    a = rng.rand(n) + 0.1
    b = rng.rand(m) + 0.1
    a, b = (a / a.sum()).astype(npd), (b / b.sum()).astype(npd)
    ref = sorc.solve(M, a, b, reg=0.1, reg_type=rt, max_iter=30, tol=0)
    Mt, at, bt = to_dev(M, a, b)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r = ot.solve(Mt, at, bt, reg=0.1, reg_type=rt, max_iter=30, tol=0)
    tol = 1e-10 if dtype == torch.float64 else 1e-4
    assert_plan(r.plan, ref["plan"], tol)
    assert_scalar(r.value_linear, ref["value_linear"], tol)
    assert_scalar(r.value, ref["value"], tol, scale=max(abs(ref["value"]), abs(ref["value_linear"])))


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_empirical_divergence_and_solve_sample(gs, tag):
    base = f"emp2_{tag}"
    xs, xt, a, b = to_dev(*(gs(base, k) for k in ("xs", "xt", "a", "b")))
    reg, its = float(gs(base, "reg")), int(gs(base, "its"))
    tol = 1e-10 if tag == "f64" else 1e-4
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        D, lg = ot.bregman.empirical_sinkhorn_divergence(xs, xt, reg, a, b, numIterMax=its, stopThr=-1, isLazy=True,
                                                         batchSize=32, log=True)
        scale = abs(float(gs(base + "/div", "l_ab")))
        assert_scalar(lg["sinkhorn_loss_ab"], gs(base + "/div", "l_ab"), tol)
        assert_scalar(lg["sinkhorn_loss_a"], gs(base + "/div", "l_a"), tol)
        assert_scalar(lg["sinkhorn_loss_b"], gs(base + "/div", "l_b"), tol)
        assert_scalar(D, gs(base + "/div", "div"), tol, scale=scale)
        D2 = ot.bregman.empirical_sinkhorn_divergence(xs, xt, reg, a, b, numIterMax=its, stopThr=-1, isLazy=True)
        assert_scalar(D2, gs(base + "/div", "div"), tol, scale=scale)
        # solve_sample, lazy
        r = ot.solve_sample(xs, xt, a, b, reg=reg, lazy=True, max_iter=its, tol=-1, batch_size=32)
        c = base + "/solve_sample"
        assert r.plan is None and r.value is None
        assert_scalar(r.value_linear, gs(c, "value_linear"), tol)
        assert_vec(r.potentials[0], gs(c, "f"), tol)
        assert_vec(r.potentials[1], gs(c, "g"), tol)
        assert_plan(r.lazy_plan[3:9, 5:12], gs(c, "block"), tol)
        # solve_sample, dense
        r = ot.solve_sample(xs, xt, a, b, reg=reg, max_iter=its, tol=0)
        c = base + "/solve_sample_dense"
        assert_plan(r.plan, gs(c, "plan"), tol)
        assert_scalar(r.value_linear, gs(c, "value_linear"), tol)
        assert_scalar(r.value, gs(c, "value"), tol, scale=max(abs(float(gs(c, "value"))), abs(float(gs(c, "value_linear")))))
    with pytest.raises(NotImplementedError):
        ot.solve_sample(xs, xt, lazy=True)
    with pytest.raises(NotImplementedError):
        ot.solve_sample(xs, xt, reg=1.0, method="lowrank")


# ---- unbalanced scaling-domain Sinkhorn ----
UNB_CASES = [f"unb_{t}_{s}" for t in ("f64", "f32") for s in ("33x47", "128x160")]
UNB_VARIANTS = {"scalar": dict(reg_m=1.0), "pair": dict(reg_m=(0.5, 2.0)), "semi": dict(reg_m=(float("inf"), 1.0)),
                "entropy": dict(reg_m=1.0, reg_type="entropy")}


@pytest.mark.parametrize("case", UNB_CASES)
@pytest.mark.parametrize("variant", list(UNB_VARIANTS))
def test_unbalanced_fixed_golden(gs, case, variant):
    M, a, b = to_dev(gs(case, "M"), gs(case, "a"), gs(case, "b"))
    reg, its, tol = float(gs(case, "reg")), int(gs(case, "its")), tol_of(case)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, reg, numItermax=its, stopThr=0, log=True,
                                                        **UNB_VARIANTS[variant])
        P2 = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, reg, numItermax=its, stopThr=0, **UNB_VARIANTS[variant])
    c = f"{case}/{variant}"
    assert P.dtype == M.dtype and P.device == M.device
    assert_plan(P, gs(c, "plan"), tol)
    assert_plan(P2, gs(c, "plan"), tol)
    assert_vec(lg["logu"], gs(c, "logu"), tol)
    assert_vec(lg["logv"], gs(c, "logv"), tol)
    assert_scalar(lg["cost"], gs(c, "cost"), tol)
    assert_scalar(lg["total_cost"], gs(c, "total_cost"), tol,
                  scale=max(abs(float(gs(c, "cost"))), abs(float(gs(c, "total_cost")))))
    eref = np.asarray(gs(c, "err"), dtype=np.float64)
    e = np.array([float(x) for x in lg["err"]])
    assert e.shape == eref.shape
    # err is a difference of successive iterates: in fp32 its noise floor is a few ulp of max|u| per iteration
    floor = 1e-13 if "f64" in case else 1e-6
    assert np.all(np.abs(e - eref) <= 1e-3 * eref + floor), (e[-3:], eref[-3:])


@pytest.mark.parametrize("case", UNB_CASES)
def test_unbalanced_conv_c_warm_multib(gs, case):
    M, a, b = to_dev(gs(case, "M"), gs(case, "a"), gs(case, "b"))
    reg, its, tol = float(gs(case, "reg")), int(gs(case, "its")), tol_of(case)
    P, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, reg, 1.0, numItermax=5000,
                                                    stopThr=1e-6 if "f32" in case else 1e-9, log=True)
    nref = int(gs(f"{case}/conv", "nerr"))
    if "f64" in case:
        assert len(lg["err"]) == nref
    else:
        assert abs(len(lg["err"]) - nref) <= max(2, nref // 20)
    assert_plan(P, gs(f"{case}/conv", "plan"), 1e-7 if "f64" in case else 1e-4)
    c = f"{case}/c_warm"
    cm, wu, wv = to_dev(gs(c, "c"), gs(c, "ws_u"), gs(c, "ws_v"))
    P, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, reg, 1.0, c=cm, warmstart=(wu, wv), numItermax=its,
                                                    stopThr=0, log=True)
    assert_plan(P, gs(c, "plan"), tol)
    assert_scalar(lg["total_cost"], gs(c, "total_cost"), tol,
                  scale=max(abs(float(gs(c, "cost"))), abs(float(gs(c, "total_cost")))))
    c = f"{case}/multib"
    (b2,) = to_dev(gs(c, "b2"))
    L, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b2, M, reg, 1.0, numItermax=its, stopThr=0, log=True)
    np.testing.assert_allclose(npf(L), gs(c, "loss"), rtol=tol)
    assert_vec(lg["logu"], gs(c, "logu"), tol)
    assert_vec(lg["logv"], gs(c, "logv"), tol)


def test_unbalanced_guard(gs):
    M, a, b = to_dev(gs("unb_guard", "M"), gs("unb_guard", "a"), gs("unb_guard", "b"))
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        P, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, float(gs("unb_guard", "reg")), 1.0, numItermax=50,
                                                        stopThr=0, log=True)
    assert int(any("Numerical errors" in str(x.message) for x in w)) == int(gs("unb_guard", "warned"))
    assert len(lg["err"]) == int(gs("unb_guard", "nerr"))
    assert_vec(lg["logu"], gs("unb_guard", "logu"), 1e-10)  # back at the previous (initial) scalings
    np.testing.assert_allclose(npf(P), gs("unb_guard", "plan"), rtol=1e-10, atol=0)  # K underflowed: plan is zero


@pytest.mark.parametrize("case", SOLVE_CASES)
def test_solve_unbalanced_golden(gs, case):
    M, a, b = to_dev(gs(case, "M"), gs(case, "a"), gs(case, "b"))
    reg, its, tol = float(gs(case, "reg")), int(gs(case, "its")), tol_of(case)
    for utag, um in (("unb1", 1.0), ("unb_pair", (0.5, 2.0))):
        r = ot.solve(M, a, b, reg=reg, unbalanced=um, max_iter=its, tol=0)
        c = f"{case}/{utag}"
        assert_plan(r.plan, gs(c, "plan"), tol)
        assert_vec(r.potentials[0], gs(c, "f"), tol)
        assert_vec(r.potentials[1], gs(c, "g"), tol)
        assert_scalar(r.value_linear, gs(c, "value_linear"), tol)
        assert_scalar(r.value, gs(c, "value"), tol,
                      scale=max(abs(float(gs(c, "value"))), abs(float(gs(c, "value_linear")))))


@pytest.mark.parametrize("n,m,dtype,variant", [(2048, 4096, torch.float32, "fused"), (1000, 1531, torch.float32, "split"),
                                               (1024, 768, torch.float64, "fused"), (333, 1001, torch.float64, "split")])
def test_unbalanced_vs_oracle_larger(n, m, dtype, variant):
    """Fused-sweep shapes (aligned) and split-sweep shapes (ragged) against the oracle on seeded inputs."""
    rng = np.random.RandomState(n + m)
    npd = np.float32 if dtype == torch.float32 else np.float64
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = ((xs ** 2).sum(1)[:, None] + (xt ** 2).sum(1)[None, :] - 2 * xs @ xt.T).clip(0)
    M = (M / M.max()).astype(npd)
    a = rng.rand(n) + 0.1
    b = rng.rand(m) + 0.1
    a, b = (1.2 * a / a.sum()).astype(npd), (b / b.sum()).astype(npd)
    Pr, lr = sorc.sinkhorn_knopp_unbalanced(a, b, M, 0.1, (1.0, 0.7), numItermax=25, stopThr=0, log=True)
    Mt, at, bt = to_dev(M, a, b)
    P, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(at, bt, Mt, 0.1, (1.0, 0.7), numItermax=25, stopThr=0, log=True)
    tol = 1e-10 if dtype == torch.float64 else 1e-4
    assert_plan(P, Pr, tol)
    assert_vec(lg["logu"], lr["logu"], tol)
    assert_scalar(lg["total_cost"], lr["total_cost"], tol, scale=max(abs(lr["cost"]), abs(lr["total_cost"])))
    assert len(lg["err"]) == 25


def test_unbalanced_errors():
    M = torch.rand(8, 6, device=DEV, dtype=torch.float64)
    a = torch.full((8,), 0.125, device=DEV, dtype=torch.float64)
    b = torch.full((6,), 1 / 6, device=DEV, dtype=torch.float64)
    with pytest.raises(ValueError):
        ot.sinkhorn_unbalanced(a, b, M, 1.0, 1.0, method="nope")
    with pytest.raises(NotImplementedError):
        ot.sinkhorn_unbalanced(a, b, M, 1.0, 1.0, method="sinkhorn_stabilized")
    with pytest.raises(ValueError):
        ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, 1.0, (1.0, 2.0, 3.0))
    P = ot.sinkhorn_unbalanced([], [], M, 1.0, 1.0)  # empty weights -> uniform (:706-709)
    P2 = ot.sinkhorn_unbalanced(a, b, M, 1.0, 1.0)
    assert torch.allclose(P, P2)
