"""GPU parity tests of the dense Sinkhorn path (scaling + log domain) against the CPU oracle and the
golden vectors of the reference.  Everything goes through libb2ot.so (C ABI) via pot_b200.

Tolerances (north_star): plan and potentials within 1e-10 relative in fp64 and 1e-4 relative in fp32 at a
fixed iteration count.  "Relative" for plan entries spanning many decades = |x - ref| <= tol * (|ref| + floor)
with floor = 1e-6 * max(ref); for log-potentials relative to max(1, max|ref|).
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

DENSE = [f"dense_{t}_{s}" for t in ("f64", "f32") for s in ("50x60", "33x47", "128x96")]
DEV = "cuda:0"


def tol_of(dtype):
    return 1e-10 if dtype in (np.float64, torch.float64) else 1e-4


def to_dev(*xs):
    return [torch.as_tensor(np.ascontiguousarray(x), device=DEV) for x in xs]


def assert_plan(G, Gref, tol):
    G = G.detach().cpu().numpy().astype(np.float64)
    Gref = np.asarray(Gref, dtype=np.float64)
    err = np.abs(G - Gref) / (np.abs(Gref) + 1e-6 * Gref.max())
    assert err.max() < tol, f"plan rel err {err.max():.3e} >= {tol}"


def assert_scaling(x, ref, tol):
    x = x.detach().cpu().numpy().astype(np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    err = np.abs(x - ref) / np.abs(ref)
    assert err.max() < tol, f"scaling rel err {err.max():.3e} >= {tol}"


def assert_logpot(x, ref, tol):
    x = x.detach().cpu().numpy().astype(np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    err = np.abs(x - ref).max() / max(1.0, np.abs(ref).max())
    assert err < tol, f"log-potential err {err:.3e} >= {tol}"


def assert_errs(errs, ref, dtype):
    e = np.array([float(x) for x in errs])
    r = np.asarray(ref, dtype=np.float64)
    assert e.shape == r.shape
    floor = 1e-13 if dtype == np.float64 else 2e-7   # marginal error bottoms out at round-off level
    assert np.all(np.abs(e - r) <= 1e-3 * r + floor), (e, r)


@pytest.mark.parametrize("variant", ["fused", "split", "resident"])
@pytest.mark.parametrize("case", DENSE)
def test_knopp_golden_fixed(golden, case, variant):
    a, b, M = golden(case, "a"), golden(case, "b"), golden(case, "M")
    if variant == "fused" and M.shape[1] % (4 if M.dtype == np.float32 else 2):
        pytest.skip("fused sweep needs 16-byte aligned rows")
    reg, its = float(golden(case, "reg")), int(golden(case, "its"))
    at, bt, Mt = to_dev(a, b, M)
    G, lg = ot.bregman.sinkhorn_knopp(at, bt, Mt, reg, numItermax=its, stopThr=0, log=True, warn=False,
                                      b2ot_variant=variant)
    tol = tol_of(M.dtype)
    assert G.dtype == Mt.dtype and G.device == Mt.device and G.shape == Mt.shape
    assert lg["niter"] == int(golden(case + "/knopp_fixed", "niter")) == its - 1
    assert_scaling(lg["u"], golden(case + "/knopp_fixed", "u"), tol)
    assert_scaling(lg["v"], golden(case + "/knopp_fixed", "v"), tol)
    assert_plan(G, golden(case + "/knopp_fixed", "plan"), tol)
    assert_errs(lg["err"], golden(case + "/knopp_fixed", "err"), M.dtype)
    assert all(e.dtype == Mt.dtype and e.device == Mt.device for e in lg["err"])


@pytest.mark.parametrize("variant", ["fused", "split", "resident"])
@pytest.mark.parametrize("case", DENSE)
def test_log_golden_fixed(golden, case, variant):
    a, b, M = golden(case, "a"), golden(case, "b"), golden(case, "M")
    if variant == "fused" and M.shape[1] % (4 if M.dtype == np.float32 else 2):
        pytest.skip("fused sweep needs 16-byte aligned rows")
    reg, its = float(golden(case, "reg")), int(golden(case, "its"))
    at, bt, Mt = to_dev(a, b, M)
    G, lg = ot.bregman.sinkhorn_log(at, bt, Mt, reg, numItermax=its, stopThr=0, log=True, warn=False,
                                    b2ot_variant=variant)
    tol = tol_of(M.dtype)
    assert G.dtype == Mt.dtype and G.device == Mt.device
    assert lg["niter"] == its - 1
    assert_logpot(lg["log_u"], golden(case + "/log_fixed", "log_u"), tol)
    assert_logpot(lg["log_v"], golden(case + "/log_fixed", "log_v"), tol)
    assert_plan(G, golden(case + "/log_fixed", "plan"), tol)
    assert_errs(lg["err"], golden(case + "/log_fixed", "err"), M.dtype)
    np.testing.assert_allclose(lg["u"].cpu().numpy(), np.exp(lg["log_u"].cpu().numpy()), rtol=1e-6)


@pytest.mark.parametrize("case", DENSE)
def test_converged_and_warmstart(golden, case):
    a, b, M = golden(case, "a"), golden(case, "b"), golden(case, "M")
    reg = float(golden(case, "reg"))
    at, bt, Mt = to_dev(a, b, M)
    f64 = M.dtype == np.float64
    thr = float(golden(case + "/knopp_conv", "thr"))
    for name, fn in (("knopp", ot.bregman.sinkhorn_knopp), ("log", ot.bregman.sinkhorn_log)):
        G, lg = fn(at, bt, Mt, reg, numItermax=2000, stopThr=thr, log=True)
        ref_it = int(golden(f"{case}/{name}_conv", "niter"))
        if f64:
            assert lg["niter"] == ref_it   # same stop iteration as the reference
        else:
            assert abs(lg["niter"] - ref_it) <= 10
        np.testing.assert_allclose(G.cpu().numpy(), golden(f"{case}/{name}_conv", "plan"), atol=1e-5 if not f64 else 1e-12,
                                   rtol=1e-8 if f64 else 1e-3)
        assert float(lg["err"][-1]) < thr
    ws = to_dev(golden(case + "/knopp_warm", "ws_u"), golden(case + "/knopp_warm", "ws_v"))
    tol = tol_of(M.dtype)
    G, lg = ot.sinkhorn(at, bt, Mt, reg, numItermax=20, stopThr=0, log=True, warn=False, warmstart=tuple(ws))
    assert_plan(G, golden(case + "/knopp_warm", "plan"), tol)
    G, lg = ot.sinkhorn(at, bt, Mt, reg, method="sinkhorn_log", numItermax=20, stopThr=0, log=True, warn=False,
                        warmstart=tuple(ws))
    assert_plan(G, golden(case + "/log_warm", "plan"), tol)


@pytest.mark.parametrize("case", DENSE)
def test_sinkhorn2_loss(golden, case):
    a, b, M = golden(case, "a"), golden(case, "b"), golden(case, "M")
    reg, its = float(golden(case, "reg")), int(golden(case, "its"))
    at, bt, Mt = to_dev(a, b, M)
    tol = 1e-10 if M.dtype == np.float64 else 1e-4
    L = ot.sinkhorn2(at, bt, Mt, reg, numItermax=its, stopThr=0)
    assert L.dtype == Mt.dtype and L.dim() == 0
    np.testing.assert_allclose(float(L), float(golden(case + "/loss", "knopp")), rtol=tol)
    L, lg = ot.sinkhorn2(at, bt, Mt, reg, method="sinkhorn_log", numItermax=its, stopThr=0, log=True)
    np.testing.assert_allclose(float(L), float(golden(case + "/loss", "log")), rtol=tol)
    assert "log_u" in lg


def test_multi_histogram(golden):
    a, b, M = golden("multib", "a"), golden("multib", "b"), golden("multib", "M")
    at, bt, Mt = to_dev(a, b, M)
    L, lg = ot.bregman.sinkhorn_knopp(at, bt, Mt, 0.1, numItermax=100, stopThr=0, log=True, warn=False)
    np.testing.assert_allclose(L.cpu().numpy(), golden("multib", "knopp_loss"), rtol=1e-10)
    np.testing.assert_allclose(lg["u"].cpu().numpy(), golden("multib", "knopp_u"), rtol=1e-10)
    L, lg = ot.bregman.sinkhorn_log(at, bt, Mt, 0.1, numItermax=100, stopThr=0, log=True, warn=False)
    np.testing.assert_allclose(L.cpu().numpy(), golden("multib", "log_loss"), rtol=1e-10)
    np.testing.assert_allclose(lg["log_v"].cpu().numpy(), golden("multib", "log_v"), rtol=1e-9, atol=1e-10)
    L2 = ot.sinkhorn2(at, bt, Mt, 0.1, numItermax=100, stopThr=0)
    np.testing.assert_allclose(L2.cpu().numpy(), golden("multib", "knopp_loss"), rtol=1e-10)


def test_numerical_guard_reverts_like_reference(golden):
    """reg so small that K underflows: the reference warns, comes back to the previous (u, v) and stops
    (ot/bregman/_sinkhorn.py:633-645).  The device-side guard must stop at the same iteration."""
    a, b, M = golden("guard", "a"), golden("guard", "b"), golden("guard", "M")
    at, bt, Mt = to_dev(a, b, M)
    for variant in ("fused", "split", "resident"):
        with warnings.catch_warnings(record=True) as w:
            warnings.simplefilter("always")
            G, lg = ot.bregman.sinkhorn_knopp(at, bt, Mt, 1e-3, numItermax=50, stopThr=0, log=True,
                                              b2ot_variant=variant)
        assert int(golden("guard", "warned")) == 1
        assert any("numerical errors at iteration %d" % int(golden("guard", "niter")) in str(x.message) for x in w)
        assert lg["niter"] == int(golden("guard", "niter"))
        np.testing.assert_allclose(lg["u"].cpu().numpy(), golden("guard", "u"), rtol=1e-10)
        np.testing.assert_allclose(lg["v"].cpu().numpy(), golden("guard", "v"), rtol=1e-10)
        np.testing.assert_allclose(G.cpu().numpy(), golden("guard", "plan"), rtol=1e-10, atol=0)
        assert len(lg["err"]) == len(golden("guard", "err"))


def test_empty_weights_lists_and_errors(golden):
    M = golden("empty", "M")
    (Mt,) = to_dev(M)
    G = ot.sinkhorn([], [], Mt, 0.2, numItermax=30, stopThr=0, warn=False)
    np.testing.assert_allclose(G.cpu().numpy(), golden("empty", "plan"), rtol=1e-10)
    n, m = M.shape
    G2 = ot.sinkhorn([1.0 / n] * n, [1.0 / m] * m, Mt, 0.2, numItermax=30, stopThr=0, warn=False)
    np.testing.assert_allclose(G2.cpu().numpy(), golden("empty", "plan"), rtol=1e-10)
    with pytest.raises(ValueError, match="Unknown method"):
        ot.sinkhorn([], [], Mt, 0.2, method="nope")
    with pytest.raises(ValueError):
        ot.sinkhorn([], [], Mt.cpu(), 0.2)              # no CPU fallback
    with pytest.raises(TypeError):
        ot.sinkhorn([], [], M, 0.2)                     # numpy input: no other backend


def test_convergence_warning(golden):
    M = golden("empty", "M")
    (Mt,) = to_dev(M)
    with pytest.warns(UserWarning, match="did not converge"):
        ot.sinkhorn([], [], Mt, 0.2, numItermax=1, stopThr=0)
    with pytest.warns(UserWarning, match="did not converge"):
        ot.bregman.sinkhorn_log([], [], Mt, 0.2, numItermax=1, stopThr=0)
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        ot.sinkhorn([], [], Mt, 0.2, numItermax=1, stopThr=0, warn=False)
        ot.sinkhorn([], [], Mt, 0.2, numItermax=500, stopThr=1e-7)


def test_doctest_known_answers():
    a, b = [0.5, 0.5], [0.5, 0.5]
    M = torch.tensor([[0.0, 1.0], [1.0, 0.0]], dtype=torch.float64, device=DEV)
    want = np.array([[0.36552929, 0.13447071], [0.13447071, 0.36552929]])      # _sinkhorn.py:127-129
    for fn in (ot.sinkhorn, ot.bregman.sinkhorn_knopp, ot.bregman.sinkhorn_log):
        np.testing.assert_allclose(fn(a, b, M, 1).cpu().numpy(), want, atol=1e-8)
    np.testing.assert_allclose(float(ot.sinkhorn2(a, b, M, 1)), 0.26894142136999516, rtol=1e-9)   # :346-347


def test_config1_full(golden_c1):
    """BASELINE config 1: n=m=1000 fp64, reg=0.05, 1000 fixed iterations, against the reference's own output."""
    rng = np.random.RandomState(0)
    x = rng.randn(1000, 2)
    y = rng.randn(1000, 2) + 1
    M = orc.sqeuclidean(x, y)
    a = np.ones(1000) / 1000
    at, bt, Mt = to_dev(a, a.copy(), M)
    G, lg = ot.sinkhorn(at, bt, Mt, 0.05, numItermax=1000, stopThr=0, log=True, warn=False)
    assert lg["niter"] == 999
    np.testing.assert_allclose(lg["u"].cpu().numpy(), golden_c1["u"], rtol=1e-10)
    np.testing.assert_allclose(lg["v"].cpu().numpy(), golden_c1["v"], rtol=1e-10)
    np.testing.assert_allclose(G[0].cpu().numpy(), golden_c1["plan_row0"], rtol=1e-10, atol=1e-300)
    np.testing.assert_allclose(G[:, 0].cpu().numpy(), golden_c1["plan_col0"], rtol=1e-10, atol=1e-300)
    np.testing.assert_allclose(float((G * Mt).sum()), float(golden_c1["loss"]), rtol=1e-10)
    assert len(lg["err"]) == len(golden_c1["err"]) == 100


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("mode", ["knopp", "log"])
@pytest.mark.parametrize("shape", [(3, 4), (1, 8), (7, 1), (257, 1031), (130, 4100), (37, 8196), (50, 20000),
                                   (33, 65536), (40, 32768)])
def test_oracle_parity_shapes(mode, shape, dtype):
    """Ragged / tiny / cluster-sized shapes: every sweep configuration (single CTA, clusters of 2/4/8,
    masked tails, split fallback for unaligned rows) against the oracle at a fixed iteration count."""
    n, m = shape
    rng = np.random.RandomState(n * 7 + m)
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = orc.sqeuclidean(xs, xt)
    M = (M / M.max()).astype(dtype)
    a = rng.rand(n) + 0.1
    b = rng.rand(m) + 0.1
    a, b = (a / a.sum()).astype(dtype), (b / b.sum()).astype(dtype)
    at, bt, Mt = to_dev(a, b, M)
    tol = tol_of(dtype)
    fn, ofn = ((ot.bregman.sinkhorn_knopp, orc.sinkhorn_knopp) if mode == "knopp"
               else (ot.bregman.sinkhorn_log, orc.sinkhorn_log))
    G, lg = fn(at, bt, Mt, 0.1, numItermax=12, stopThr=0, log=True, warn=False)
    Gr, lgr = ofn(a, b, M, 0.1, numItermax=12, stopThr=0, log=True, warn=False)
    assert lg["niter"] == lgr["niter"] == 11
    assert_plan(G, Gr, tol)
    if mode == "knopp":
        assert_scaling(lg["u"], lgr["u"], tol)
        assert_scaling(lg["v"], lgr["v"], tol)
    else:
        assert_logpot(lg["log_u"], lgr["log_u"], tol)
        assert_logpot(lg["log_v"], lgr["log_v"], tol)
    assert_errs(lg["err"], lgr["err"], dtype)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(1, 1), (3, 4), (7, 1), (148, 256), (149, 257), (257, 1031), (1000, 1000), (2000, 999),
                                   (4000, 1500), (300, 4096)])
def test_resident_solver(shape, dtype):
    """Small problems on the resident solver (K in the grid's shared memory, one cooperative launch per chunk of
    iterations, dense_resident_kernels.cuh): ragged row / column splits over the CTAs against the oracle and against the
    streaming sweep; convergence stop at the reference's iteration; chunked iterate calls; what does not fit is refused."""
    n, m = shape
    rng = np.random.RandomState(n * 5 + m)
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = orc.sqeuclidean(xs, xt)
    M = (M / M.max()).astype(dtype)
    a = rng.rand(n) + 0.1
    b = rng.rand(m) + 0.1
    a, b = (a / a.sum()).astype(dtype), (b / b.sum()).astype(dtype)
    at, bt, Mt = to_dev(a, b, M)
    tol = tol_of(dtype)
    grid = min(148, max(n, m))  # plan_resident (dense.cu): a row block and a column block of K per CTA, <= 200 KB
    R_, Q_ = -(-n // grid), -(-m // grid)
    fits = (R_ * m + Q_ * n + 2 * m + n + 2 * (R_ + Q_) + 64) * M.itemsize + 64 <= 200 * 1024
    if not fits:
        with pytest.raises(_lib.B2otError, match="resident solver needs"):
            ot.bregman.sinkhorn_knopp(at, bt, Mt, 0.1, numItermax=5, stopThr=0, warn=False, b2ot_variant="resident")
        return
    G, lg = ot.bregman.sinkhorn_knopp(at, bt, Mt, 0.1, numItermax=23, stopThr=0, log=True, warn=False,
                                      b2ot_variant="resident")
    Gr, lgr = orc.sinkhorn_knopp(a, b, M, 0.1, numItermax=23, stopThr=0, log=True, warn=False)
    assert lg["niter"] == lgr["niter"] == 22
    assert_plan(G, Gr, tol)
    assert_scaling(lg["u"], lgr["u"], tol)
    assert_scaling(lg["v"], lgr["v"], tol)
    assert_errs(lg["err"], lgr["err"], dtype)
    Gs, lgs = ot.bregman.sinkhorn_knopp(at, bt, Mt, 0.1, numItermax=23, stopThr=0, log=True, warn=False,
                                        b2ot_variant="split")
    assert_scaling(lg["u"], lgs["u"].cpu().numpy(), 1e-12 if dtype == np.float64 else 1e-5)
    # the default picks it, and a converging solve stops where the reference stops
    Gc, lgc = ot.bregman.sinkhorn_knopp(at, bt, Mt, 0.1, numItermax=3000, stopThr=1e-6, log=True, warn=False)
    Grc, lgrc = orc.sinkhorn_knopp(a, b, M, 0.1, numItermax=3000, stopThr=1e-6, log=True, warn=False)
    if dtype == np.float64:
        assert lgc["niter"] == lgrc["niter"]
    else:
        assert abs(lgc["niter"] - lgrc["niter"]) <= 10
    assert len(lgc["err"]) == len(lgrc["err"]) or dtype == np.float32


def test_resident_handle_chunks_and_info():
    """The handle API on the resident path: info reports it, iterate calls of any size continue the same solve (state
    written back by the kernel), sinkhorn_log refuses the variant."""
    from pot_b200 import _dense
    n, m = 500, 640
    rng = np.random.RandomState(2)
    K = torch.as_tensor(np.exp(-rng.rand(n, m) / 0.3), device=DEV)
    a = torch.full((n,), 1.0 / n, dtype=torch.float64, device=DEV)
    b = torch.full((m,), 1.0 / m, dtype=torch.float64, device=DEV)
    outs = []
    for variant, chunks in (("resident", (1, 2, 7, 10, 17)), ("resident", (37,)), ("fused", (37,))):
        s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=1000, check_every=10, stop_thr=0.0, variant=variant)
        assert s.info()["variant"] == variant
        s.init(a.clone(), b.clone())
        for c in chunks:
            s.iterate(c)
        st, it = s.poll()
        assert st == 0 and it == 37
        u, v, errs = s.export()
        outs.append((u, v, errs))
        s.close()
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])  # chunking changes nothing
    assert outs[0][2] == outs[1][2] and len(outs[0][2]) == 4
    np.testing.assert_allclose(outs[0][0].cpu().numpy(), outs[2][0].cpu().numpy(), rtol=1e-12)
    np.testing.assert_allclose(outs[0][2], outs[2][2], rtol=1e-6, atol=1e-15)
    s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=1000, check_every=10, stop_thr=0.0)
    assert s.info()["variant"] == "resident"   # what "auto" resolves to at this size
    s.close()
    s = _dense.DenseSolver("log", K.log(), a.log(), b, b.log(), max_iter=10, check_every=10, stop_thr=0.0)
    assert s.info()["variant"] == "resident"   # the log domain has its resident form too
    s.close()


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(1, 1), (3, 4), (7, 1), (149, 257), (257, 1031), (1000, 1000), (300, 2100)])
def test_resident_log_solver(shape, dtype):
    """Log-domain resident solver (exact log-sum-exps on the row and column blocks) against the oracle's sinkhorn_log
    (ot/bregman/_sinkhorn.py:880-906) and the streaming sweep."""
    n, m = shape
    rng = np.random.RandomState(n * 3 + m)
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = orc.sqeuclidean(xs, xt)
    M = (M / M.max()).astype(dtype)
    a = rng.rand(n) + 0.1
    b = rng.rand(m) + 0.1
    a, b = (a / a.sum()).astype(dtype), (b / b.sum()).astype(dtype)
    at, bt, Mt = to_dev(a, b, M)
    tol = tol_of(dtype)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G, lg = ot.bregman.sinkhorn_log(at, bt, Mt, 0.05, numItermax=23, stopThr=0, log=True, warn=False,
                                        b2ot_variant="resident")
        Gr, lgr = orc.sinkhorn_log(a, b, M, 0.05, numItermax=23, stopThr=0, log=True, warn=False)
        Gs, lgs = ot.bregman.sinkhorn_log(at, bt, Mt, 0.05, numItermax=23, stopThr=0, log=True, warn=False,
                                          b2ot_variant="split")
    assert lg["niter"] == lgr["niter"] == 22
    assert_plan(G, Gr, tol)
    fin = np.isfinite(lgr["log_u"])
    assert np.array_equal(np.isfinite(lg["log_u"].cpu().numpy()), fin)
    assert_logpot(lg["log_u"][torch.as_tensor(fin, device=DEV)], lgr["log_u"][fin], tol)
    assert_logpot(lg["log_v"], lgr["log_v"], tol)
    assert_errs(lg["err"], lgr["err"], dtype)
    assert_plan(G, Gs.cpu().numpy(), 10 * tol)
    # converging solve: same stop iteration as the reference in fp64
    Gc, lgc = ot.bregman.sinkhorn_log(at, bt, Mt, 0.05, numItermax=3000, stopThr=1e-6, log=True, warn=False)
    Grc, lgrc = orc.sinkhorn_log(a, b, M, 0.05, numItermax=3000, stopThr=1e-6, log=True, warn=False)
    if dtype == np.float64:
        assert lgc["niter"] == lgrc["niter"]
    else:
        assert abs(lgc["niter"] - lgrc["niter"]) <= 10


def test_noncontiguous_and_strided_inputs():
    rng = np.random.RandomState(3)
    n, m = 96, 128
    big = rng.rand(n, 2 * m)
    M = big[:, :m] / big[:, :m].max()          # row stride 2m: leading dimension != m
    a, b = np.ones(n) / n, np.ones(m) / m
    bigt = torch.as_tensor(big / big[:, :m].max(), device=DEV)
    Mt = bigt[:, :m]
    assert not Mt.is_contiguous()
    at, bt = to_dev(a, b)
    G = ot.sinkhorn(at, bt, Mt, 0.5, numItermax=25, stopThr=0, warn=False)
    Gr = orc.sinkhorn_knopp(a, b, np.ascontiguousarray(M), 0.5, numItermax=25, stopThr=0, warn=False)
    assert_plan(G, Gr, 1e-10)
    Gt = ot.sinkhorn(bt, at, Mt.t(), 0.5, numItermax=25, stopThr=0, warn=False)   # column-major view
    Grt = orc.sinkhorn_knopp(b, a, np.ascontiguousarray(M.T), 0.5, numItermax=25, stopThr=0, warn=False)
    assert_plan(Gt, Grt, 1e-10)


def test_log_underflow_fallback_matches_oracle():
    """A target point so far away that its row-frame column sum underflows in fp32: the device asks for the
    exact column log-sum-exp pass and the result still matches the reference's logsumexp formulation."""
    rng = np.random.RandomState(5)
    n, m = 64, 65
    xs, xt = rng.randn(n, 2) * 0.05, rng.randn(m, 2) * 0.05
    xt[-1] += 30.0
    M = orc.sqeuclidean(xs, xt)
    M = (M / M.max()).astype(np.float32)
    a, b = (np.ones(n) / n).astype(np.float32), (np.ones(m) / m).astype(np.float32)
    at, bt, Mt = to_dev(a, b, M)
    reg = 0.002
    G, lg = ot.bregman.sinkhorn_log(at, bt, Mt, reg, numItermax=15, stopThr=0, log=True, warn=False)
    Gr, lgr = orc.sinkhorn_log(a, b, M, reg, numItermax=15, stopThr=0, log=True, warn=False)
    assert lg["niter"] == lgr["niter"]
    assert torch.isfinite(lg["log_v"]).all()
    assert_logpot(lg["log_u"], lgr["log_u"], 1e-4)
    assert_logpot(lg["log_v"], lgr["log_v"], 1e-4)
    assert_plan(G, Gr, 1e-4)


@pytest.mark.parametrize("mode", ["knopp", "log"])
def test_large_fused_vs_split_and_marginals(mode):
    """BASELINE-scale property checks (n=m=16384 fp32, config 2 shape): the fused single-pass sweep and the
    independent two-pass split implementation agree, the row marginal of the plan equals a (the u update is
    the last one applied), and the reported error equals the recomputed column-marginal violation."""
    n = m = 16384
    g = torch.Generator(device=DEV).manual_seed(0)
    xs = torch.randn(n, 2, device=DEV, generator=g)
    xt = torch.randn(m, 2, device=DEV, generator=g) + 1.0
    M = torch.cdist(xs, xt).pow_(2)
    M /= M.max()
    a = torch.full((n,), 1.0 / n, device=DEV)
    b = torch.full((m,), 1.0 / m, device=DEV)
    reg = 0.05 if mode == "knopp" else 0.01
    fn = ot.bregman.sinkhorn_knopp if mode == "knopp" else ot.bregman.sinkhorn_log
    its = 21
    G1, l1 = fn(a, b, M, reg, numItermax=its, stopThr=0, log=True, warn=False, b2ot_variant="fused")
    G2, l2 = fn(a, b, M, reg, numItermax=its, stopThr=0, log=True, warn=False, b2ot_variant="split")
    ku = "u" if mode == "knopp" else "log_u"
    if mode == "knopp":
        assert ((l1[ku] - l2[ku]).abs() / l2[ku].abs()).max() < 1e-4
    else:
        assert (l1[ku] - l2[ku]).abs().max() / l2[ku].abs().max() < 1e-4
    row = G1.sum(1, dtype=torch.float64)
    assert ((row - a.double()).abs() / a.double()).max() < 1e-4
    col = G1.sum(0, dtype=torch.float64)
    err = (col - b.double()).norm().item()
    assert abs(float(l1["err"][-1]) - err) <= 1e-3 * err + 1e-7
    assert l1["niter"] == l2["niter"] == its - 1


def test_launch_counter_and_variant_info():
    from pot_b200 import _dense
    lib = _lib.load()
    K = torch.rand(256, 512, device=DEV)
    a = torch.full((256,), 1 / 256, device=DEV)
    b = torch.full((512,), 1 / 512, device=DEV)
    s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=100, variant="fused")
    assert s.info()["variant"] == "fused"
    c0 = lib.b2ot_launch_count()
    s.init(a.clone(), b.clone())
    s.iterate(10)
    st, it = s.poll()
    assert (st, it) == (0, 10)
    assert lib.b2ot_launch_count() - c0 == 2 + 2 * 10
    s.close()
    s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=100)   # small problem: one launch per iterate call
    assert s.info()["variant"] == "resident"
    c0 = lib.b2ot_launch_count()
    s.init(a.clone(), b.clone())
    s.iterate(10)
    assert s.poll() == (0, 10)
    assert lib.b2ot_launch_count() - c0 == 2 + 1
    s.close()
