"""GPU parity tests of the online-cost (empirical) Sinkhorn path, ot.dist and the lazy plan against the golden
vectors of the reference and the CPU oracle.  Everything goes through libb2ot.so (b2ot_emp_*, b2ot_dist).

Tolerances (north_star): potentials / plan within 1e-10 relative in fp64 and 1e-4 relative in fp32 at a fixed
iteration count (stopThr = -1 disables the reference's `<=` stop test).
"""
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot
    from pot_b200 import _lib
from oracle import sinkhorn_oracle as orc

DEV = "cuda:0"
EMP = [f"emp_{t}_{s}" for t in ("f64", "f32") for s in ("40x50x3", "64x32x8", "37x41x2")]


def tol_of(x):
    return 1e-10 if x in ("f64", np.float64, torch.float64) else 1e-4


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


def errs_close(errs, ref, dtype):
    e = np.array([float(x) for x in errs])
    r = np.asarray(ref, dtype=np.float64)
    assert e.shape == r.shape, (e, r)
    floor = 1e-13 if dtype == "f64" else 3e-6  # L1 error over n rows bottoms out at n * eps
    assert np.all(np.abs(e - r) <= 2e-3 * r + floor), (e, r)


@pytest.mark.parametrize("case", EMP)
def test_lazy_golden_fixed_iterations(golden, case):
    tag = case.split("_")[1]
    xs, xt, a, b = (golden(case, k) for k in ("xs", "xt", "a", "b"))
    reg, its = float(golden(case, "reg")), int(golden(case, "its"))
    tol = tol_of(tag)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g, lg = ot.empirical_sinkhorn(dev(xs), dev(xt), reg, dev(a), dev(b), numIterMax=its, stopThr=-1,
                                         isLazy=True, batchSize=tuple(int(x) for x in golden(case, "bs")), log=True)
    assert f.dtype == torch.as_tensor(xs).dtype and f.is_cuda
    assert lg["niter"] == int(golden(case, "niter")) == its - 1
    logclose(f, golden(case, "f"), tol, "f")
    logclose(g, golden(case, "g"), tol, "g")
    errs_close(lg["err"], golden(case, "err"), tag)
    assert lg["u"] is f and lg["v"] is g
    # lazy plan: full and block reads
    lp = lg["lazy_plan"]
    assert lp.shape == (xs.shape[0], xt.shape[0])
    close(lp[:], golden(case, "plan"), 10 * tol, "lazy_plan[:]")
    close(lp[3:11, 5:20], golden(case, "plan_block"), 10 * tol, "block")  # the fixture's block (make_golden.py)
    row = lp[5]
    assert row.shape == (1, xt.shape[0])
    close(row, golden(case, "plan")[5:6], 10 * tol, "lazy_plan[5]")


@pytest.mark.parametrize("case", EMP)
def test_lazy_golden_converged(golden, case):
    """The device-side stop test (`err <= stopThr` at (it + 1) % 10 == 0) fires at the reference's iteration."""
    tag = case.split("_")[1]
    xs, xt, a, b = (golden(case, k) for k in ("xs", "xt", "a", "b"))
    reg, thr = float(golden(case, "reg")), float(golden(case + "/conv", "thr"))
    f, g, lg = ot.empirical_sinkhorn(dev(xs), dev(xt), reg, dev(a), dev(b), numIterMax=3000, stopThr=thr, isLazy=True,
                                     batchSize=7, log=True)
    assert lg["niter"] == int(golden(case + "/conv", "niter"))
    assert len(lg["err"]) == len(golden(case + "/conv", "err"))
    logclose(f, golden(case + "/conv", "f"), 10 * tol_of(tag), "f")
    logclose(g, golden(case + "/conv", "g"), 10 * tol_of(tag), "g")


def test_default_weights_and_nonlazy(golden):
    xs, xt = golden("emp_default", "xs"), golden("emp_default", "xt")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g = ot.empirical_sinkhorn(dev(xs), dev(xt), 0.7, numIterMax=50, stopThr=-1, isLazy=True, batchSize=7)
        G = ot.empirical_sinkhorn(dev(xs), dev(xt), 0.7, numIterMax=50, stopThr=0)
    logclose(f, golden("emp_default", "f"), 1e-10, "f")
    logclose(g, golden("emp_default", "g"), 1e-10, "g")
    close(G, golden("emp_default", "plan"), 1e-10, "plan (non-lazy)")
    # doctest of the reference (_empirical.py:159-167)
    x = torch.arange(2, dtype=torch.float64, device=DEV).reshape(2, 1)
    P = ot.empirical_sinkhorn(x, x.clone(), 0.1)
    close(P, golden("doctest", "emp_2x2"), 1e-10, "doctest plan")


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(1, 1, 1), (5, 3, 1), (70, 130, 2), (129, 64, 17), (300, 257, 64), (1000, 777, 5)])
def test_dist(dtype, shape):
    n, m, d = shape
    rng = np.random.RandomState(n + m)
    X, Y = rng.randn(n, d).astype(dtype), (rng.randn(m, d) + 0.5).astype(dtype)
    tol = 1e-12 if dtype == np.float64 else 2e-5
    C = ot.dist(dev(X), dev(Y))
    ref = orc.sqeuclidean(X.astype(np.float64), Y.astype(np.float64))
    assert C.shape == (n, m) and float(C.min()) >= 0
    np.testing.assert_allclose(C.cpu().numpy(), ref, rtol=tol, atol=tol * max(1.0, ref.max()))
    E = ot.dist(dev(X), dev(Y), metric="euclidean")
    np.testing.assert_allclose(E.cpu().numpy(), np.sqrt(ref), rtol=1e3 * tol, atol=3e-3 if dtype == np.float32 else 1e-7)
    S = ot.dist(dev(X))
    assert torch.all(torch.diagonal(S) == 0) and S.shape == (n, n)
    with pytest.raises(NotImplementedError):
        ot.dist(dev(X), dev(Y), metric="cityblock")


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(1, 9, 2), (65, 64, 3), (200, 333, 16), (513, 190, 64), (257, 1030, 7)])
def test_lazy_vs_oracle(dtype, shape):
    """Ragged sizes across the 64 x 64 tiling and the reduction splits; non-uniform weights; warm start."""
    n, m, d = shape
    rng = np.random.RandomState(7 * n + m)
    xs = (rng.randn(n, d) / np.sqrt(d)).astype(dtype)
    xt = (rng.randn(m, d) / np.sqrt(d) + 0.3).astype(dtype)
    a = rng.rand(n) + 0.2
    b = rng.rand(m) + 0.2
    a, b = (a / a.sum()).astype(dtype), (b / b.sum()).astype(dtype)
    ws = (rng.randn(n).astype(dtype) * 0.1, rng.randn(m).astype(dtype) * 0.1)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        fr, gr, lr = orc.empirical_sinkhorn_lazy(xs, xt, 0.5, a, b, numIterMax=20, stopThr=-1, batchSize=50, log=True,
                                                 warmstart=ws)
        f, g, lg = ot.empirical_sinkhorn(dev(xs), dev(xt), 0.5, dev(a), dev(b), numIterMax=20, stopThr=-1, isLazy=True,
                                         log=True, warmstart=(dev(ws[0]), dev(ws[1])))
    tol = tol_of(dtype)
    logclose(f, fr, tol, "f")
    logclose(g, gr, tol, "g")
    assert lg["niter"] == lr["niter"] == 19 and len(lg["err"]) == 2
    # marginal violation of the lazy plan (g was updated last: column marginal exact up to round-off)
    P = lg["lazy_plan"][:]
    assert float(torch.abs(P.sum(0) - dev(b)).max()) < (1e-12 if dtype == np.float64 else 1e-6)
    loss = ot.empirical_sinkhorn2(dev(xs), dev(xt), 0.5, dev(a), dev(b), numIterMax=20, stopThr=-1, isLazy=True,
                                  warn=False, warmstart=(dev(ws[0]), dev(ws[1])))
    C = orc.sqeuclidean(xs, xt)
    ref_loss = float(np.sum(C * np.exp(fr[:, None] + gr[None, :] - C / 0.5)))
    assert abs(float(loss) - ref_loss) <= 10 * tol * abs(ref_loss)


def test_warnings_errors_and_dtype():
    rng = np.random.RandomState(0)
    xs, xt = dev(rng.randn(20, 2)), dev(rng.randn(30, 2))
    with pytest.warns(UserWarning, match="did not converge"):
        ot.empirical_sinkhorn(xs, xt, 1.0, numIterMax=5, stopThr=0, isLazy=True)
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        ot.empirical_sinkhorn(xs, xt, 1.0, numIterMax=5, stopThr=0, isLazy=True, warn=False)
    with pytest.raises(ValueError, match="Batch size"):
        ot.empirical_sinkhorn(xs, xt, 1.0, isLazy=True, batchSize=(1, 2, 3))
    with pytest.raises(ValueError, match="no CPU fallback"):
        ot.empirical_sinkhorn(xs.cpu(), xt.cpu(), 1.0, isLazy=True)
    with pytest.raises(NotImplementedError):
        ot.empirical_sinkhorn(xs, xt, 1.0, isLazy=True, metric="cosine")
    f, g = ot.empirical_sinkhorn(xs.float(), xt.float(), 1.0, numIterMax=30, isLazy=True, warn=False)
    assert f.dtype == torch.float32 and g.dtype == torch.float32 and f.device == xs.device
    # euclidean metric runs through the same sweep
    f, g, lg = ot.empirical_sinkhorn(xs, xt, 1.0, numIterMax=200, isLazy=True, metric="euclidean", log=True)
    P = lg["lazy_plan"][:]
    assert float(torch.abs(P.sum(1) - 1.0 / 20).sum()) < 1e-8
    assert lg["niter"] < 199


def test_lazy_equals_dense_log_solver():
    """Reference test_lazy_empirical_sinkhorn: the lazy plan has the marginals of the dense solution (atol 1e-5)."""
    rng = np.random.RandomState(3)
    xs, xt = rng.rand(100, 2), rng.rand(120, 2)
    M = orc.sqeuclidean(xs, xt)
    f, g, lg = ot.empirical_sinkhorn(dev(xs), dev(xt), 1.0, numIterMax=1000, isLazy=True, batchSize=(1, 3), log=True)
    G = ot.sinkhorn([], [], dev(M), 1.0, stopThr=1e-10)
    P = lg["lazy_plan"][:]
    np.testing.assert_allclose(P.sum(1).cpu().numpy(), G.sum(1).cpu().numpy(), atol=1e-5)
    np.testing.assert_allclose(P.sum(0).cpu().numpy(), G.sum(0).cpu().numpy(), atol=1e-5)
    np.testing.assert_allclose(P.cpu().numpy(), G.cpu().numpy(), atol=1e-5)


# ------------------------------------------------------------------------------------------------
# tensor-core sweep (tcgen05, 3xTF32): fp32, sqeuclidean, d <= 64
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("operand", ["f16", "tf32"])
@pytest.mark.parametrize("case", [c for c in EMP if "_f32_" in c])
def test_tc_golden(golden, case, operand):
    xs, xt, a, b = (golden(case, k) for k in ("xs", "xt", "a", "b"))
    reg, its = float(golden(case, "reg")), int(golden(case, "its"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g, lg = ot.empirical_sinkhorn(dev(xs), dev(xt), reg, dev(a), dev(b), numIterMax=its, stopThr=-1, isLazy=True,
                                         log=True, b2ot_variant="tc", b2ot_operand=operand)
    logclose(f, golden(case, "f"), 1e-4, "f (tc)")
    logclose(g, golden(case, "g"), 1e-4, "g (tc)")
    assert lg["niter"] == its - 1
    errs_close(lg["err"], golden(case, "err"), "f32")


@pytest.mark.parametrize("operand", ["f16", "tf32"])
@pytest.mark.parametrize("shape", [(1, 9, 2), (130, 127, 3), (200, 333, 16), (513, 190, 64), (257, 1030, 33), (1500, 900, 64)])
def test_tc_vs_oracle(shape, operand):
    """The tcgen05 sweep against the fp64 oracle: ragged tiles, 1 and 2 K-chunks, reduction splits, weights,
    warm start.  Also the auto variant and the register-tiled sweep must agree with it."""
    n, m, d = shape
    rng = np.random.RandomState(11 * n + m)
    xs = (rng.randn(n, d) / np.sqrt(d)).astype(np.float32)
    xt = (rng.randn(m, d) / np.sqrt(d) + 0.3).astype(np.float32)
    a = rng.rand(n) + 0.2
    b = rng.rand(m) + 0.2
    a, b = (a / a.sum()).astype(np.float32), (b / b.sum()).astype(np.float32)
    ws = (rng.randn(n).astype(np.float32) * 0.1, rng.randn(m).astype(np.float32) * 0.1)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        fr, gr, lr = orc.empirical_sinkhorn_lazy(xs.astype(np.float64), xt.astype(np.float64), 0.5, a.astype(np.float64),
                                                 b.astype(np.float64), numIterMax=20, stopThr=-1, batchSize=256, log=True,
                                                 warmstart=(ws[0].astype(np.float64), ws[1].astype(np.float64)))
        out = {}
        for variant in ("tc", "simt", "auto"):
            f, g, lg = ot.empirical_sinkhorn(dev(xs), dev(xt), 0.5, dev(a), dev(b), numIterMax=20, stopThr=-1,
                                             isLazy=True, log=True, warmstart=(dev(ws[0]), dev(ws[1])),
                                             b2ot_variant=variant, b2ot_operand=operand)
            logclose(f, fr, 1e-4, f"f ({variant})")
            logclose(g, gr, 1e-4, f"g ({variant})")
            assert lg["niter"] == 19 and len(lg["err"]) == 2
            out[variant] = (f, g)
    # the two sweeps differ only by round-off (3xTF32 keeps the dot product at fp32 level)
    assert float(torch.abs(out["tc"][0] - out["simt"][0]).max()) < 2e-5 * max(1.0, float(np.abs(fr).max()))


@pytest.mark.parametrize("operand", ["f16", "tf32"])
@pytest.mark.parametrize("kind", ["outliers", "tiny_coords", "offcentre"])
def test_tc_wide_dynamic_range(kind, operand):
    """Clouds the operand split could be sensitive to (the 3xFP16 split has one power-of-two scale per cloud, so its
    low halves lose bits for coordinates far below the cloud maximum; 3xTF32 has an 8-bit exponent):
      outliers    : a few points at 2^10 x the typical norm;
      tiny_coords : some coordinates at 2^-12 x the typical size;
      offcentre   : the whole cloud shifted by 30 standard deviations (|x|^2 + |y|^2 - 2 x.y cancels 3 digits).
    Checked against the fp64 oracle on the same fp32 inputs at the fp32 tolerance, and against the fp32 SIMT sweep."""
    n, m, d = 700, 520, 48
    rng = np.random.RandomState(97)
    xs = (rng.randn(n, d) / np.sqrt(d)).astype(np.float32)
    xt = (rng.randn(m, d) / np.sqrt(d) + 0.2).astype(np.float32)
    if kind == "outliers":
        xs[::97] *= 1024.0
        xt[5::131] *= 1024.0
    elif kind == "tiny_coords":
        xs[:, ::5] *= 2.0 ** -12
        xt[:, 1::7] *= 2.0 ** -12
        xs[3::11] *= 2.0 ** -12
    else:
        xs += 30.0 / np.sqrt(d)
        xt += 30.0 / np.sqrt(d)
    reg = 0.5
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        fr, gr = orc.empirical_sinkhorn_lazy(xs.astype(np.float64), xt.astype(np.float64), reg, numIterMax=20,
                                             stopThr=-1, batchSize=256)
        f, g = ot.empirical_sinkhorn(dev(xs), dev(xt), reg, numIterMax=20, stopThr=-1, isLazy=True,
                                     b2ot_variant="tc", b2ot_operand=operand)
        fs, gs = ot.empirical_sinkhorn(dev(xs), dev(xt), reg, numIterMax=20, stopThr=-1, isLazy=True,
                                       b2ot_variant="simt")
    assert torch.isfinite(f).all() and torch.isfinite(g).all()
    logclose(f, fr, 1e-4, f"f ({kind}, {operand})")
    logclose(g, gr, 1e-4, f"g ({kind}, {operand})")
    # the typical points must be right too, not only the outlier-dominated maximum: compare where |ref| is moderate
    fn, gn = f.cpu().numpy().astype(np.float64), g.cpu().numpy().astype(np.float64)
    for x, r, nm in ((fn, fr, "f"), (gn, gr, "g")):
        typ = np.abs(r) <= 10 * np.median(np.abs(r)) + 1.0
        scale = max(1.0, np.abs(r[typ]).max())
        # fp32 inputs: the cost itself carries eps * (|x|^2 + |y|^2) / reg of rounding; allow the SIMT sweep's own error
        simt = (fs if nm == "f" else gs).cpu().numpy().astype(np.float64)
        allow = max(1e-4 * scale, 4.0 * np.abs(simt[typ] - r[typ]).max())
        assert np.abs(x[typ] - r[typ]).max() <= allow, (nm, kind, operand, np.abs(x[typ] - r[typ]).max(), allow)


@pytest.mark.parametrize("operand", ["f16", "tf32"])
@pytest.mark.parametrize("reg", [0.5, 0.02])
def test_tc_coincident_points(reg, operand):
    """The tensor-core sweep leaves out the reference's clamp C >= 0 (ot/utils.py:299; emp_tc_kernels.cuh header): it
    matters only where the computed cost is negative, i.e. for (near-)coincident points.  Worst case on purpose: the
    target cloud IS the source cloud (every row has an exact zero of the cost on the diagonal) plus duplicated points,
    small reg so that the diagonal term dominates each log-sum-exp.  Reference: the fp64 oracle, which clamps."""
    n, d = 900, 64
    rng = np.random.RandomState(5)
    xs = (rng.randn(n, d) / np.sqrt(d)).astype(np.float32)
    xs[1::50] = xs[::50][: len(xs[1::50])]  # duplicates off the diagonal too
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        fr, gr = orc.empirical_sinkhorn_lazy(xs.astype(np.float64), xs.astype(np.float64), reg, numIterMax=20,
                                             stopThr=-1, batchSize=256)
        X = dev(xs)
        f, g = ot.empirical_sinkhorn(X, X, reg, numIterMax=20, stopThr=-1, isLazy=True, b2ot_variant="tc",
                                     b2ot_operand=operand)
    assert torch.isfinite(f).all() and torch.isfinite(g).all()
    # (f, g) share one additive gauge and on a symmetric problem g ~ 0 while |f| ~ log n: one joint scale.  What is
    # left at small reg is not the clamp (the clamping build gives the same digits, scripts/diag_coincident.py:
    # 2.6e-4 absolute at reg = 0.02 with and without it) but the fp32 accumulation of the tensor pipe on x.x ~ |x|^2,
    # dC ~ 5e-6 |x|^2, divided by reg.
    scale = max(1.0, float(np.abs(fr).max()), float(np.abs(gr).max()))
    for got, ref, nm in ((f, fr, "f"), (g, gr, "g")):
        err = float(np.abs(got.cpu().numpy().astype(np.float64) - ref).max())
        assert err <= 1e-4 * scale, (nm, reg, operand, err, scale)


@pytest.mark.parametrize("n", [16384])
def test_tc_multiwave_vs_fp64(n):
    """Config-4 shaped sweep at a size with several waves of CTAs and reduction splits (16384^2 x 64: 128 row tiles x
    splits over 148 SMs): 3 iterations of the tensor-core solver against a chunked fp64 torch restatement of the same
    recursion (_empirical.py:218-233: f = log a - LSE_j(g_j - C_ij/reg), then g = log b - LSE_i(f_i - C_ij/reg)), 1e-4."""
    d, reg, its = 64, 0.1, 3
    gen = torch.Generator(device="cuda").manual_seed(3)
    xs = torch.randn(n, d, device="cuda", generator=gen) / d ** 0.5
    xt = torch.randn(n, d, device="cuda", generator=gen) / d ** 0.5 + 1.0 / d ** 0.5
    f, g = ot.empirical_sinkhorn(xs, xt, reg, numIterMax=its, stopThr=-1, isLazy=True, warn=False, b2ot_variant="tc")
    X, Y = xs.double(), xt.double()
    la = torch.full((n,), -np.log(n), dtype=torch.float64, device="cuda")
    fr = torch.zeros(n, dtype=torch.float64, device="cuda")
    gr = torch.zeros(n, dtype=torch.float64, device="cuda")

    def half(P, Q, pot):  # log(1/n) - LSE_j(pot_j - |P_i - Q_j|^2 / reg), row chunks of 2048
        out = torch.empty(P.shape[0], dtype=torch.float64, device="cuda")
        qq = (Q * Q).sum(1)
        for c0 in range(0, P.shape[0], 2048):
            Pc = P[c0:c0 + 2048]
            C = ((Pc * Pc).sum(1)[:, None] + qq[None, :] - 2.0 * Pc @ Q.T).clamp_min_(0.0)
            out[c0:c0 + 2048] = -torch.logsumexp(pot[None, :] - C / reg, dim=1)
        return out + la

    for _ in range(its):
        fr = half(X, Y, gr)
        gr = half(Y, X, fr)
    for got, ref, nm in ((f, fr, "f"), (g, gr, "g")):
        err = float((got.double() - ref).abs().max())
        scale = max(1.0, float(ref.abs().max()))
        assert err <= 1e-4 * scale, (nm, err, scale)


def test_tc_rejects_ineligible():
    rng = np.random.RandomState(0)
    xs, xt = dev(rng.randn(20, 70).astype(np.float32)), dev(rng.randn(30, 70).astype(np.float32))
    with pytest.raises(_lib.B2otError, match="tensor-core sweep needs"):
        ot.empirical_sinkhorn(xs, xt, 1.0, isLazy=True, numIterMax=5, b2ot_variant="tc")
    with pytest.raises(_lib.B2otError, match="tensor-core sweep needs"):
        ot.empirical_sinkhorn(xs[:, :8].double(), xt[:, :8].double(), 1.0, isLazy=True, numIterMax=5, b2ot_variant="tc")
    f, g = ot.empirical_sinkhorn(xs, xt, 1.0, isLazy=True, numIterMax=5, warn=False)  # auto -> register-tiled sweep
    assert torch.isfinite(f).all()
