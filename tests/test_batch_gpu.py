"""GPU parity tests of the batched on-chip solver (ot.batch path) against the golden vectors of the reference
and the CPU oracle.  Everything goes through libb2ot.so (b2ot_batch_solve) via pot_b200.batch.

Tolerances (north_star): 1e-10 relative in fp64, 1e-4 relative in fp32 at a fixed iteration count; plan entries
are compared as |x - ref| <= tol * (|ref| + 1e-6 * max(ref)).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot
    from pot_b200 import _lib
    from pot_b200 import batch as otb
from oracle import sinkhorn_oracle as orc

DEV = "cuda:0"


def tol_of(tag):
    return 1e-10 if tag in ("f64", np.float64) else 1e-4


def dev(x):
    return torch.as_tensor(np.ascontiguousarray(x), device=DEV)


def close(x, ref, tol, what, floor_rel=1e-6):
    x = x.detach().cpu().numpy().astype(np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert x.shape == ref.shape, (what, x.shape, ref.shape)
    err = np.abs(x - ref) / (np.abs(ref) + floor_rel * np.abs(ref).max())
    assert err.max() < tol, f"{what}: rel err {err.max():.3e} >= {tol}"


def logclose(x, ref, tol, what):
    x = x.detach().cpu().numpy().astype(np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    err = np.abs(x - ref).max() / max(1.0, np.abs(ref).max())
    assert err < tol, f"{what}: err {err:.3e} >= {tol}"


@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("solver", ["log_sinkhorn", "sinkhorn"])
@pytest.mark.parametrize("reg_type", ["entropy", "kl"])
def test_solve_batch_golden(golden, tag, solver, reg_type):
    """Fixed iteration count and converged runs against the unmodified reference's outputs."""
    base = f"batch_{tag}"
    M, a, b = golden(base, "M"), golden(base, "a"), golden(base, "b")
    tol = tol_of(tag)
    r = ot.solve_batch(dev(M), 0.1, dev(a), dev(b), max_iter=40, tol=0, solver=solver, reg_type=reg_type)
    c = f"{base}/{solver}_{reg_type}_fixed"
    assert r.log["n_iter"] == int(golden(c, "n_iter")) == 39
    assert r.plan.dtype == torch.as_tensor(M).dtype and r.plan.is_cuda
    close(r.plan, golden(c, "plan"), tol, "plan")
    close(r.value, golden(c, "value"), tol, "value")
    close(r.value_linear, golden(c, "value_linear"), tol, "value_linear")
    if solver == "sinkhorn":
        close(r.potentials[0], golden(c, "pa"), tol, "f")
        close(r.potentials[1], golden(c, "pb"), tol, "g")
    else:
        logclose(r.potentials[0], golden(c, "pa"), tol, "u")
        logclose(r.potentials[1], golden(c, "pb"), tol, "v")
    # convergence run: the device-side stop test must fire at the same check as the reference's
    r = ot.solve_batch(dev(M), 0.1, dev(a), dev(b), max_iter=1000, tol=1e-5, solver=solver, reg_type=reg_type)
    c = f"{base}/{solver}_{reg_type}_conv"
    assert r.log["n_iter"] == int(golden(c, "n_iter"))
    close(r.plan, golden(c, "plan"), tol, "plan (converged)")


def test_solve_batch_uniform_defaults(golden):
    M = golden("batch_f64", "M")
    r = ot.solve_batch(dev(M), 0.1, max_iter=25, tol=0)
    close(r.plan, golden("batch_f64/uniform", "plan"), 1e-10, "plan")
    close(r.value, golden("batch_f64/uniform", "value"), 1e-10, "value")
    logclose(r.potentials[0], golden("batch_f64/uniform", "pa"), 1e-10, "u")
    ma, mb = r.marginals
    assert ma.shape == (5, 24) and mb.shape == (5, 31)


def _problem(B, n, m, dtype, seed=0, d=2):
    rng = np.random.RandomState(seed)
    X = rng.randn(B, n, d)
    Y = rng.randn(B, m, d) + 0.5
    M = orc.dist_sqeuclidean_batch(X, Y)
    M = (M / M.max()).astype(dtype)
    a = rng.rand(B, n) + 0.5
    b = rng.rand(B, m) + 0.5
    a = (a / a.sum(1, keepdims=True)).astype(dtype)
    b = (b / b.sum(1, keepdims=True)).astype(dtype)
    return M, a, b


SHAPES = [(3, 7, 5), (4, 33, 47), (2, 128, 96), (3, 64, 300), (2, 60, 1000), (2, 256, 256), (2, 300, 280),
          (1, 150, 480)]
# full tiles: the register-resident fast path of batch_solve_kernel (NTRIP = 1 / 2, 4 and 8 columns per lane, clusters
# of 1 to 8 CTAs), and near misses that must fall back to the generic loop
FAST_SHAPES = [(3, 64, 256), (2, 128, 256), (5, 256, 256), (2, 512, 256), (3, 64, 128), (2, 128, 128),
               (2, 256, 128), (2, 32, 256), (2, 255, 256), (2, 256, 255), (2, 192, 256)]


@pytest.mark.parametrize("shape", SHAPES + FAST_SHAPES)
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("solver", ["log_sinkhorn", "sinkhorn"])
def test_solve_batch_vs_oracle(shape, dtype, solver):
    """Ragged / non-aligned shapes, 1-CTA and multi-CTA (cluster) problems, fixed iteration count."""
    B, n, m = shape
    if dtype == np.float64 and m > 512 and n > 8:
        # the fp64 cross-warp column scratch alone is 128 KB at m > 512: documented limit, clean error
        with pytest.raises(_lib.B2otError, match="does not fit"):
            otb.batch_config("log", torch.float64, n, m)
        return
    M, a, b = _problem(B, n, m, dtype, seed=n + m)
    its = 30
    ref = orc.solve_batch(M, 0.1, a, b, max_iter=its, tol=0, solver=solver, reg_type="entropy")
    r = ot.solve_batch(dev(M), 0.1, dev(a), dev(b), max_iter=its, tol=0, solver=solver, reg_type="entropy")
    tol = tol_of(dtype)
    close(r.plan, ref["plan"], tol, "plan")
    close(r.value_linear, ref["value_linear"], tol, "value_linear")
    close(r.value, ref["value"], 10 * tol, "value")
    mode = "knopp" if solver == "sinkhorn" else "log"
    cs, smem, cpl = otb.batch_config(mode, torch.as_tensor(M).dtype, n, m)
    assert 1 <= cs <= 8 and cpl * 32 >= m
    # marginal violation of the returned plan (reported, and bounded loosely: 30 iterations at reg = 0.1)
    viol = float(torch.abs(r.plan.sum(2) - dev(a)).max())
    assert viol < 1e-3, viol


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_projection_functions(dtype):
    B, n, m = 3, 40, 50
    M, a, b = _problem(B, n, m, dtype, seed=5)
    K = np.exp(-M / 0.1).astype(dtype)
    tol = tol_of(dtype)
    ref = orc.bregman_projection_batch(K, a, b, max_iter=20, tol=0)
    out = otb.bregman_projection_batch(dev(K), dev(a), dev(b), max_iter=20, tol=0)
    assert out["n_iters"] == ref["n_iters"] == 19
    close(out["T"], ref["T"], tol, "T")
    close(out["potentials"][0], ref["potentials"][0], tol, "f")
    close(out["potentials"][1], ref["potentials"][1], tol, "g")
    Kl = (-M / 0.1).astype(dtype)
    ref = orc.bregman_log_projection_batch(Kl, a, b, max_iter=20, tol=0)
    out = otb.bregman_log_projection_batch(dev(Kl), dev(a), dev(b), max_iter=20, tol=0)
    assert out["n_iters"] == 19
    close(out["T"], ref["T"], tol, "T")
    logclose(out["log_T"], ref["log_T"], tol, "log_T")
    logclose(out["potentials"][0], ref["potentials"][0], tol, "u")
    logclose(out["potentials"][1], ref["potentials"][1], tol, "v")
    # scaling and log domain agree (reference test_bregman_batch: atol 1e-3 at reg 0.01, 50 iterations)
    K2, Kl2 = np.exp(-M / 0.01).astype(dtype), (-M / 0.01).astype(dtype)
    o1 = otb.bregman_projection_batch(dev(K2), dev(a), dev(b), max_iter=50, tol=0)
    o2 = otb.bregman_log_projection_batch(dev(Kl2), dev(a), dev(b), max_iter=50, tol=0)
    assert float(torch.abs(o1["T"] - o2["T"]).max()) < 1e-3


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(3, 40, 50), (2, 300, 280)])
def test_log_exact_column_pass(dtype, shape):
    """The exact (online-max) column pass, forced, gives the same iterates as the row-frame fast path."""
    B, n, m = shape
    M, a, b = _problem(B, n, m, dtype, seed=9)
    ref = orc.solve_batch(M, 0.05, a, b, max_iter=15, tol=0, solver="log_sinkhorn")
    from pot_b200 import batch as pbatch
    with pbatch._solver_flags(_lib.BATCH_FORCE_EXACT):
        r = ot.solve_batch(dev(M), 0.05, dev(a), dev(b), max_iter=15, tol=0, solver="log_sinkhorn")
    close(r.plan, ref["plan"], tol_of(dtype), "plan (exact column pass)")


def test_log_underflow_falls_back():
    """fp32, tiny reg: row-frame column sums underflow for far-away columns; the solver must stay finite and
    match the oracle (which uses scipy's exact logsumexp)."""
    B, n, m = 2, 48, 40
    rng = np.random.RandomState(3)
    X = rng.randn(B, n, 2)
    Y = rng.randn(B, m, 2)
    Y[:, 0, :] += 40.0  # one target point far from every source: its column is e^-hundreds in every row frame
    M = orc.dist_sqeuclidean_batch(X, Y).astype(np.float32)
    a = np.full((B, n), 1.0 / n, np.float32)
    b = np.full((B, m), 1.0 / m, np.float32)
    ref = orc.solve_batch(M, 1.0, a, b, max_iter=20, tol=0, solver="log_sinkhorn")
    r = ot.solve_batch(dev(M), 1.0, dev(a), dev(b), max_iter=20, tol=0, solver="log_sinkhorn")
    assert torch.isfinite(r.plan).all() and torch.isfinite(r.potentials[1]).all()
    logclose(r.potentials[1], ref["potentials"][1], 1e-4, "v")
    # K, u, v are O(1e3) here: one fp32 ulp of the exponent K + u + v is ~1e-4, so two correctly rounded fp32
    # evaluations of the plan differ by ~1e-3 relative.  Compare with the fp64 oracle at that conditioning.
    ref64 = orc.solve_batch(M.astype(np.float64), 1.0, a.astype(np.float64), b.astype(np.float64), max_iter=20,
                            tol=0, solver="log_sinkhorn")
    close(r.plan, ref64["plan"], 5e-3, "plan")


def test_stop_test_couples_the_batch():
    """The reference stops when the WORST problem of the batch is below tol (max over the batch): an easy problem
    next to a hard one keeps iterating; n_iter equals the oracle's."""
    M1, a1, b1 = _problem(1, 30, 30, np.float64, seed=1)
    M = np.concatenate([M1 * 0.05, M1 * 3.0])  # problem 0 converges almost at once, problem 1 slowly
    a, b = np.concatenate([a1, a1]), np.concatenate([b1, b1])
    for solver in ("sinkhorn", "log_sinkhorn"):
        ref = orc.solve_batch(M, 0.1, a, b, max_iter=2000, tol=1e-7, solver=solver)
        r = ot.solve_batch(dev(M), 0.1, dev(a), dev(b), max_iter=2000, tol=1e-7, solver=solver)
        assert 10 < ref["n_iter"] < 1999
        assert r.log["n_iter"] == ref["n_iter"]
        close(r.plan, ref["plan"], 1e-10, "plan")
        # max_iter reached without convergence: last index reported, silently (no warning in the reference)
        r = ot.solve_batch(dev(M), 0.1, dev(a), dev(b), max_iter=15, tol=1e-12, solver=solver)
        assert r.log["n_iter"] == 14


def test_errors_and_grad_modes():
    M, a, b = _problem(2, 10, 12, np.float64)
    with pytest.raises(ValueError, match="Unknown solver"):
        ot.solve_batch(dev(M), 0.1, solver="bogus")
    with pytest.raises(ValueError, match="no CPU fallback"):
        ot.solve_batch(torch.as_tensor(M), 0.1)
    with pytest.raises(TypeError):
        ot.solve_batch(M, 0.1)
    Mt = dev(M).requires_grad_(True)
    r = ot.solve_batch(Mt, 0.1, dev(a), dev(b), max_iter=50, tol=0, grad="envelope")
    r.value.sum().backward()  # envelope gradient: d value / d M = T
    close(Mt.grad, r.plan.detach().cpu().numpy(), 1e-12, "envelope gradient")
    r = ot.solve_batch(dev(M), 0.1, dev(a), dev(b), max_iter=50, tol=0, grad="detach")
    assert not r.value.requires_grad


def test_config5_shape_properties():
    """BASELINE config 5 shape (256 x 256 fp32, 2-CTA clusters) at a reduced batch: marginals, and problems
    are independent (solving a sub-batch gives bit-identical results)."""
    B, n, m = 64, 256, 256
    g = torch.Generator(device=DEV).manual_seed(0)
    X = torch.randn(B, n, 2, device=DEV, generator=g)
    Y = torch.randn(B, m, 2, device=DEV, generator=g) + 1.0
    M = ((X * X).sum(-1)[:, :, None] + (Y * Y).sum(-1)[:, None, :] - 2 * torch.einsum("bid,bjd->bij", X, Y))
    M = (M / M.max()).float().contiguous()
    cs, smem, cpl = otb.batch_config("log", torch.float32, n, m)
    assert cs == 2 and cpl == 8
    for solver in ("log_sinkhorn", "sinkhorn"):
        r = ot.solve_batch(M, 0.1, max_iter=100, tol=0, solver=solver)
        assert torch.isfinite(r.plan).all()
        assert float(torch.abs(r.plan.sum(1) - 1.0 / m).max()) < 1e-6       # g-update is the last one: exact columns
        assert float(torch.abs(r.plan.sum(2) - 1.0 / n).max()) < 1e-4
        r2 = ot.solve_batch(M[10:20].contiguous(), 0.1, max_iter=100, tol=0, solver=solver)
        assert torch.equal(r2.plan, r.plan[10:20])
        ref = orc.solve_batch(M[:2].cpu().numpy(), 0.1, max_iter=100, tol=0, solver=solver)
        close(r.plan[:2], ref["plan"], 1e-4, "plan")


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("solver", ["log_sinkhorn", "sinkhorn"])
def test_fast_path_converged_and_kl(dtype, solver):
    """Full-tile problems (register-resident fast path) with the batch-wide stop test active (state travels through
    the potentials between launches) and the KL value epilogue: n_iter and results against the oracle."""
    B, n, m = 6, 256, 256
    M, a, b = _problem(B, n, m, dtype, seed=11)
    tol_stop = 1e-6 if dtype == np.float64 else 1e-4
    ref = orc.solve_batch(M, 0.2, a, b, max_iter=500, tol=tol_stop, solver=solver, reg_type="kl")
    r = ot.solve_batch(dev(M), 0.2, dev(a), dev(b), max_iter=500, tol=tol_stop, solver=solver, reg_type="kl")
    if dtype == np.float64:
        assert r.log["n_iter"] == ref["n_iter"], (r.log["n_iter"], ref["n_iter"])
        tol = tol_of(dtype)
    else:  # fp32: the error can cross the threshold one check (10 iterations) earlier or later
        assert abs(r.log["n_iter"] - ref["n_iter"]) <= 10, (r.log["n_iter"], ref["n_iter"])
        tol = 1e-3 if r.log["n_iter"] != ref["n_iter"] else tol_of(dtype)
    close(r.plan, ref["plan"], tol, "plan")
    close(r.value, ref["value"], 10 * tol, "value")


# ------------------------------------------------------------------------------------------------
# one-problem-per-CTA kernel (batch_hyb_kernels.cuh): fp32, m = 256, n = 192 / 256 (BASELINE config 5's shape)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [192, 256])
def test_hybrid_exact_column_pass_and_underflow(n):
    """The two-pass exact column log-sum-exp of the hybrid kernel: forced (test hook) it must reproduce the fast
    row-frame update; with a far-away target point the row-frame column totals underflow and it runs for real."""
    B, m = 3, 256
    M, a, b = _problem(B, n, m, np.float32, seed=21)
    ref = orc.solve_batch(M, 0.05, a, b, max_iter=15, tol=0, solver="log_sinkhorn")
    with otb._solver_flags(_lib.BATCH_FORCE_EXACT):
        r = ot.solve_batch(dev(M), 0.05, dev(a), dev(b), max_iter=15, tol=0, solver="log_sinkhorn")
    close(r.plan, ref["plan"], 1e-4, "plan (exact column pass, hybrid)")
    rng = np.random.RandomState(5)
    X, Y = rng.randn(B, n, 2), rng.randn(B, m, 2)
    Y[:, 0, :] += 40.0
    M = orc.dist_sqeuclidean_batch(X, Y).astype(np.float32)
    ref64 = orc.solve_batch(M.astype(np.float64), 1.0, a.astype(np.float64), b.astype(np.float64), max_iter=20, tol=0,
                            solver="log_sinkhorn")
    r = ot.solve_batch(dev(M), 1.0, dev(a), dev(b), max_iter=20, tol=0, solver="log_sinkhorn")
    assert torch.isfinite(r.plan).all() and torch.isfinite(r.potentials[1]).all()
    logclose(r.potentials[1], ref64["potentials"][1], 1e-4, "v (underflow fallback, hybrid)")
    close(r.plan, ref64["plan"], 5e-3, "plan (underflow fallback, hybrid)")


@pytest.mark.parametrize("solver", ["log_sinkhorn", "sinkhorn"])
def test_hybrid_stale_reference_survives_a_jump(solver):
    """The log-domain hot loop uses last iteration's row log-sum-exp as the exponent reference.  A warm start far from
    the fixed point (v jumps by hundreds between the first iterations) must take the exact-maximum redo, not overflow:
    bregman_*_projection_batch on a badly scaled kernel, against the oracle; also an unaligned problem pointer (the
    cooperative load path) and the log-plan output."""
    B, n, m = 2, 256, 256
    M, a, b = _problem(B, n, m, np.float32, seed=31)
    M = (M * 400.0).astype(np.float32)  # K = -M / reg spans 0 .. -4000: e^-4000 underflows every row frame at first
    if solver == "log_sinkhorn":
        K = (-M / 0.1).astype(np.float32)
        ref = orc.bregman_log_projection_batch(K.astype(np.float64), a.astype(np.float64), b.astype(np.float64),
                                               max_iter=12, tol=0)
        buf = torch.zeros(B * n * m + 1, dtype=torch.float32, device=DEV)
        Kd = buf[1:].view(B, n, m)  # 4-byte aligned only
        Kd.copy_(dev(K))
        out = otb.bregman_log_projection_batch(Kd, dev(a), dev(b), max_iter=12, tol=0)
        assert torch.isfinite(out["potentials"][0]).all()
        logclose(out["potentials"][0], ref["potentials"][0], 1e-4, "u")
        logclose(out["potentials"][1], ref["potentials"][1], 1e-4, "v")
        logclose(out["log_T"], ref["log_T"], 1e-4, "log T")
    else:
        K = np.exp(-(M / 400.0) / 0.1).astype(np.float32)
        ref = orc.bregman_projection_batch(K.astype(np.float64), a.astype(np.float64), b.astype(np.float64),
                                           max_iter=12, tol=0)
        out = otb.bregman_projection_batch(dev(K), dev(a), dev(b), max_iter=12, tol=0)
        close(out["T"], ref["T"], 1e-4, "T")
        close(out["potentials"][0], ref["potentials"][0], 1e-4, "f")
