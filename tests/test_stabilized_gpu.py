"""GPU parity tests of ot.bregman.sinkhorn_stabilized / sinkhorn_epsilon_scaling against the
golden vectors of the reference (tests/golden/golden_stabilized.npz) and the CPU oracle.
Everything goes through libb2ot.so (C ABI) via pot_b200.

Tolerances (north_star): 1e-10 relative in fp64, 1e-4 relative in fp32, fixed iteration count.
"""
import ctypes as C
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pot_b200 as ot
    from pot_b200 import _dense, _lib, stabilized
from oracle import solvers_oracle as sorc
from test_oracle_stabilized import STAB, STABM, kwargs
from test_solvers_gpu import assert_plan, assert_scalar, assert_vec, npf, to_dev

DEV = "cuda:0"


def tol_of(case):
    return 1e-10 if "f64" in case else 1e-4


def fixed_count(kw):
    """stopThr = 0 means "fixed iteration count" in the golden cases.  The reference stops on err <= stopThr (:1316) and
    the device path can reach err == 0.0 exactly in fp32 (potentials accumulated in double), which that rule counts as
    converged; a negative threshold keeps the iteration count fixed."""
    return {**kw, "stopThr": -1.0} if kw.get("stopThr") == 0.0 else kw


def check_errs(errs, ref, f64):
    errs = [float(e) for e in errs]
    assert len(errs) == len(ref)
    np.testing.assert_allclose(errs, ref, rtol=1e-5 if f64 else 5e-2, atol=1e-12 if f64 else 2e-6)


@pytest.mark.parametrize("name", STAB)
def test_stabilized_golden(gstab, name):
    case = f"stab_{name}"
    M, a, b = to_dev(gstab(case, "M"), gstab(case, "a"), gstab(case, "b"))
    reg, tol = float(gstab(case, "reg")), tol_of(name)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G, log = ot.bregman.sinkhorn_stabilized(a, b, M, reg, log=True, **fixed_count(kwargs(gstab, case)))
    assert G.dtype == M.dtype and G.device == M.device
    assert log["n_iter"] == int(gstab(case, "n_iter"))
    assert_plan(G, gstab(case, "plan"), tol)
    for k in ("logu", "logv", "alpha", "beta"):
        assert_vec(log[k], gstab(case, k), tol)
    assert_vec(log["warmstart"][0], gstab(case, "alpha"), tol)
    check_errs(log["err"], gstab(case, "err"), "f64" in name)
    # marginal violation of the returned plan, reported next to the reference's
    viol = float((G.sum(0) - b).norm())
    ref_viol = float(np.linalg.norm(gstab(case, "plan").sum(0) - gstab(case, "b")))
    assert viol <= ref_viol + (1e-12 if "f64" in name else 1e-5)


@pytest.mark.parametrize("name", [n for n in STAB if "absorb" in n])
def test_absorption_happens_on_the_device_path(gstab, name):
    """The golden cases named *absorb* must really pause for absorptions (as often as the oracle's loop does)."""
    case = f"stab_{name}"
    M, a, b = to_dev(gstab(case, "M"), gstab(case, "a"), gstab(case, "b"))
    kw = fixed_count(kwargs(gstab, case))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        _, ref = sorc.sinkhorn_stabilized(gstab(case, "a"), gstab(case, "b"), gstab(case, "M"), float(gstab(case, "reg")),
                                          log=True, **kw)
        out = stabilized._stabilized_core(a, b, M, float(gstab(case, "reg")), kw["numItermax"], kw.get("tau", 1e3),
                                          kw["stopThr"], None, kw.get("print_period", 20))
    assert ref["n_absorb"] >= 1
    assert out[8] == ref["n_absorb"]
    assert out[6] == ref["n_iter"]


def test_warmstart_and_loss_through_ot_sinkhorn(gstab):
    case = "stab_absorb_f64"
    M, a, b, al, be = to_dev(*(gstab(case, k) for k in ("M", "a", "b", "alpha", "beta")))
    reg = float(gstab(case, "reg"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G, log = ot.sinkhorn(a, b, M, reg, method="sinkhorn_stabilized", warmstart=(al, be), numItermax=40,
                             stopThr=0.0, log=True)
        val = ot.sinkhorn2(a, b, M, reg, method="sinkhorn_stabilized", numItermax=200, stopThr=0.0)
        val2, log2 = ot.sinkhorn2(a, b, M, reg, method="sinkhorn_stabilized", numItermax=200, stopThr=0.0, log=True)
    assert_plan(G, gstab(case + "_warm", "plan"), 1e-10)
    assert_vec(log["alpha"], gstab(case + "_warm", "alpha"), 1e-10)
    assert_vec(log["beta"], gstab(case + "_warm", "beta"), 1e-10)
    assert log["n_iter"] == int(gstab(case + "_warm", "n_iter"))
    assert_scalar(val, gstab(case + "_loss", "value"), 1e-10)
    assert_scalar(val2, gstab(case + "_loss", "value"), 1e-10)
    assert val.dim() == 0 and val.dtype == M.dtype
    assert_vec(log2["alpha"], gstab(case, "alpha"), 1e-10)


@pytest.mark.parametrize("name", ["f64", "short_f64"])
def test_epsilon_scaling_golden(gstab, name):
    case = f"eps_{name}"
    M, a, b = to_dev(gstab(case, "M"), gstab(case, "a"), gstab(case, "b"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G, log = ot.bregman.sinkhorn_epsilon_scaling(a, b, M, float(gstab(case, "reg")), log=True,
                                                     **kwargs(gstab, case))
        G2 = ot.sinkhorn(a, b, M, float(gstab(case, "reg")), method="sinkhorn_epsilon_scaling",
                         **{k: v for k, v in kwargs(gstab, case).items()})
    assert log["niter"] == int(gstab(case, "niter"))
    # Tolerance: the stages start at reg = epsilon0 = 1e4, where the absorbed potentials are alpha ~ -eps0 log n,
    # beta ~ +eps0 log m (3.7e4 here); every later get_K cancels them in M - alpha - beta, an absolute rounding noise of
    # ~4e-12 that the last stages divide by reg = 0.01.  The reference's own result is therefore defined to ~5e-10
    # relative only (tests/diag_eps.py: a single stage from identical potentials already differs by 1e-11 / reg);
    # 40 chained stages are compared at 5e-8.
    tol = 5e-8 if name == "f64" else 1e-9
    assert_plan(G, gstab(case, "plan"), tol)
    assert_plan(G2, gstab(case, "plan"), tol)
    assert_vec(log["alpha"], gstab(case, "alpha"), tol)
    assert_vec(log["beta"], gstab(case, "beta"), tol)
    ref = gstab(case, "err")
    errs = [float(e) for e in log["err"]]
    assert len(errs) == len(ref)
    np.testing.assert_allclose(errs, ref, rtol=1e-2, atol=1e-17)
    # marginals of the returned plan
    assert float((G.sum(0) - b).norm()) < 1e-8 and float((G.sum(1) - a).norm()) < 1e-8 or name != "f64"


@pytest.mark.parametrize("dtype,n,m,reg,its,tau", [
    (np.float64, 1500, 1100, 0.004, 120, 1e3),   # fused sweep, many absorptions
    (np.float64, 700, 2300, 0.02, 60, 1e2),
    (np.float32, 1200, 1000, 0.02, 100, 1e2),
])
def test_stabilized_vs_oracle_larger(dtype, n, m, reg, its, tau):
    rng = np.random.RandomState(n + m)
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = ((xs ** 2).sum(1)[:, None] + (xt ** 2).sum(1)[None, :] - 2 * xs @ xt.T).clip(0)
    M = (M / M.max()).astype(dtype)
    a = (np.ones(n) / n).astype(dtype)
    b = rng.rand(m) + 0.1
    b = (b / b.sum()).astype(dtype)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        Gr, lr = sorc.sinkhorn_stabilized(a, b, M, reg, numItermax=its, tau=tau, stopThr=0.0, log=True)
        Md, ad, bd = to_dev(M, a, b)
        G, lg = ot.bregman.sinkhorn_stabilized(ad, bd, Md, reg, numItermax=its, tau=tau, stopThr=0.0, log=True)
    assert lr["n_absorb"] >= 1
    assert lg["n_iter"] == lr["n_iter"]
    if dtype == np.float32:
        # fp32: the reference's own result is noisy here (alpha += reg log u is rounded in fp32 at every one of the ~100
        # absorptions and divided by reg in get_K); the device path accumulates the absorbed potentials in double.
        # 1e-4 against the fp64 oracle on the same fp32 inputs, 1e-3 against the fp32 oracle.
        assert_plan(G, Gr, 1e-3)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            Gr, lr = sorc.sinkhorn_stabilized(a.astype(np.float64), b.astype(np.float64), M.astype(np.float64), reg,
                                              numItermax=its, tau=tau, stopThr=0.0, log=True)
    tol = 1e-10 if dtype == np.float64 else 1e-4
    assert_plan(G, Gr, tol)
    assert_vec(lg["alpha"], lr["alpha"], tol)
    assert_vec(lg["beta"], lr["beta"], tol)
    check_errs(lg["err"], np.asarray(lr["err"], dtype=np.float64), dtype == np.float64)
    # n or m above tau: every iteration absorbs, the last one included, so the reference returns Gamma on the reset
    # scalings (the plan scaled by 1/(n m), :1283-1287 then :1349); same here
    row, row_ref = npf(G.sum(1)), Gr.sum(1)
    assert np.abs(row - row_ref).max() / row_ref.max() < (1e-9 if dtype == np.float64 else 1e-3)


def test_convergence_stop_matches_reference_rule(gstab):
    """err <= stopThr (not <) and the check period: n_iter of the converging golden cases already pins it; here the
    boundary: stopThr equal to a recorded error must stop at that very check."""
    case = "stab_conv_f64"
    M, a, b = to_dev(gstab(case, "M"), gstab(case, "a"), gstab(case, "b"))
    kw = kwargs(gstab, case)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        _, log = ot.bregman.sinkhorn_stabilized(a, b, M, float(gstab(case, "reg")), log=True,
                                                **{**kw, "stopThr": 0.0, "numItermax": 41})
        e = float(log["err"][3])  # check of iteration 15 (print_period 5)
        _, log2 = ot.bregman.sinkhorn_stabilized(a, b, M, float(gstab(case, "reg")), log=True,
                                                 **{**kw, "stopThr": e, "numItermax": 41})
    assert log2["n_iter"] == 15 and len(log2["err"]) == 4


def test_errors():
    M = torch.rand(8, 9, device=DEV, dtype=torch.float64)
    a = torch.full((8,), 1 / 8, device=DEV, dtype=torch.float64)
    b2 = torch.full((9, 2), 1 / 9, device=DEV, dtype=torch.float64)
    with pytest.raises(TypeError):
        ot.bregman.sinkhorn_stabilized(a.cpu().numpy(), b2[:, 0].cpu().numpy(), M.cpu().numpy(), 0.1)
    # C ABI misuse: log-domain handles cannot be switched, tau must be positive, buffers must be given
    lib = _lib.load()
    b = b2[:, 0].contiguous()
    Mr = _dense.cost_transform(M, 0.1, "negscale")
    al, be = torch.zeros_like(a, dtype=torch.float64), torch.zeros_like(b, dtype=torch.float64)
    args = (Mr.data_ptr(), Mr.stride(0), al.data_ptr(), be.data_ptr())
    s = _dense.DenseSolver("log", Mr, torch.log(a), b, torch.log(b), max_iter=5)
    assert lib.b2ot_dense_set_stabilized(s.h, 1e3, *args) != 0
    s.close()
    K = _dense.cost_transform(M, 0.1, "kexp")
    s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=5)
    assert lib.b2ot_dense_set_stabilized(s.h, -1.0, *args) != 0
    assert lib.b2ot_dense_set_stabilized(s.h, 1e3, None, Mr.stride(0), al.data_ptr(), be.data_ptr()) != 0
    assert lib.b2ot_dense_set_stabilized(s.h, 1e3, *args) == 0
    cnt = C.c_int(-1)
    assert lib.b2ot_dense_absorb_count(s.h, C.byref(cnt)) == 0
    s.close()


def test_odd_layouts_take_the_scalar_rebuild():
    """m not a multiple of the vector width and a strided M: the rebuild kernel's scalar variant."""
    rng = np.random.RandomState(5)
    n, m = 130, 77
    Mfull = rng.rand(n, m + 3)
    a = np.ones(n) / n
    b = rng.rand(m) + 0.1
    b /= b.sum()
    Md = torch.as_tensor(Mfull, device=DEV)[:, 1:m + 1]  # unaligned rows, leading dimension m + 3
    M = Mfull[:, 1:m + 1]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        Gr, lr = sorc.sinkhorn_stabilized(a, b, M, 0.01, numItermax=80, tau=5.0, stopThr=0.0, log=True)
        G, lg = ot.bregman.sinkhorn_stabilized(*to_dev(a, b), Md, 0.01, numItermax=80, tau=5.0, stopThr=0.0, log=True)
    assert lr["n_absorb"] >= 1
    assert_plan(G, Gr, 1e-10)
    assert_vec(lg["alpha"], lr["alpha"], 1e-10)
    check_errs(lg["err"], np.asarray(lr["err"], dtype=np.float64), True)


# ---- several target histograms (2-D b): joint absorption / joint error on one kernel (:1230-1234, :1275-1279) ----
@pytest.mark.parametrize("name", STABM)
def test_stabilized_multi_golden(gstabm, name):
    case = f"stabm_{name}"
    M, a, b = to_dev(gstabm(case, "M"), gstabm(case, "a"), gstabm(case, "b"))
    reg, tol = float(gstabm(case, "reg")), tol_of(name)
    kw = fixed_count(kwargs(gstabm, case))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        res, log = ot.bregman.sinkhorn_stabilized(a, b, M, reg, log=True, **kw)
    assert res.dtype == M.dtype and res.device == M.device and tuple(res.shape) == (b.shape[1],)
    # fp32: the reference can hit u == uprev exactly (err = 0.0 <= stopThr = 0) and leave early; the device path sums in
    # double, keeps the fixed count and sits on the same fixed point
    same_count = int(gstabm(case, "n_iter")) == kw["numItermax"] - 1 or kw["stopThr"] > 0
    assert same_count or "f32" in name
    if same_count:
        assert log["n_iter"] == int(gstabm(case, "n_iter"))
        check_errs(log["err"], gstabm(case, "err"), "f64" in name)
    assert_vec(res, gstabm(case, "losses"), tol)
    for k in ("logu", "logv", "alpha", "beta"):
        assert tuple(log[k].shape) == gstabm(case, k).shape
        assert_vec(log[k], gstabm(case, k), tol)
    assert_vec(log["warmstart"][1], gstabm(case, "beta"), tol)


def test_stabilized_multi_warmstart_and_sinkhorn2(gstabm):
    case = "stabm_absorb_f64"
    M, a, b = to_dev(gstabm(case, "M"), gstabm(case, "a"), gstabm(case, "b"))
    reg = float(gstabm(case, "reg"))
    ws = tuple(to_dev(gstabm(case + "_warm", "ws_alpha"), gstabm(case + "_warm", "ws_beta")))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        res, log = ot.sinkhorn(a, b, M, reg, method="sinkhorn_stabilized", warmstart=ws, numItermax=40, stopThr=-1.0,
                               log=True)
        val = ot.sinkhorn2(a, b, M, reg, method="sinkhorn_stabilized", numItermax=200, stopThr=-1.0)
    assert log["n_iter"] == int(gstabm(case + "_warm", "n_iter"))
    assert_vec(res, gstabm(case + "_warm", "losses"), 1e-10)
    assert_vec(log["alpha"], gstabm(case + "_warm", "alpha"), 1e-10)
    assert_vec(log["beta"], gstabm(case + "_warm", "beta"), 1e-10)
    check_errs(log["err"], gstabm(case + "_warm", "err"), True)
    assert_vec(val, gstabm(case + "_loss", "value"), 1e-10)


@pytest.mark.parametrize("dtype,n,m,h,reg,its,tau", [
    (np.float64, 300, 257, 5, 0.01, 120, 10.0),
    (np.float64, 1000, 130, 2, 0.05, 60, 1e3),
    (np.float32, 200, 333, 6, 0.05, 80, 1e3),
])
def test_stabilized_multi_vs_oracle_larger(dtype, n, m, h, reg, its, tau):
    rng = np.random.RandomState(n + m + h)
    M = rng.rand(n, m).astype(dtype)
    a = (rng.rand(n) + 0.1).astype(dtype)
    a = (a / a.sum()).astype(dtype)
    b = (rng.rand(m, h) + 0.1).astype(dtype)
    b = (b / b.sum(0)).astype(dtype)
    f64 = dtype == np.float64
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        rr, lr = sorc.sinkhorn_stabilized(a.astype(np.float64), b.astype(np.float64), M.astype(np.float64), reg,
                                          numItermax=its, tau=tau, stopThr=-1.0, log=True)
        rg, lg = ot.bregman.sinkhorn_stabilized(*to_dev(a, b, M), reg, numItermax=its, tau=tau, stopThr=-1.0, log=True)
    tol = 1e-10 if f64 else 1e-4
    assert lg["n_iter"] == lr["n_iter"] == its - 1
    assert_vec(rg, rr, tol)
    assert_vec(lg["logu"], lr["logu"], tol)
    assert_vec(lg["logv"], lr["logv"], tol)
    check_errs(lg["err"], np.asarray(lr["err"], dtype=np.float64), f64)


def test_stabilized_multi_convergence_and_guard():
    rng = np.random.RandomState(11)
    n, m, h = 50, 60, 3
    M = rng.rand(n, m)
    a = np.ones(n) / n
    b = rng.rand(m, h) + 0.1
    b /= b.sum(0)
    # stop rule (:1316): same iteration as the oracle, no warning
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        rr, lr = sorc.sinkhorn_stabilized(a, b, M, 0.1, numItermax=5000, stopThr=1e-9, print_period=3, log=True)
        rg, lg = ot.bregman.sinkhorn_stabilized(*to_dev(a, b, M), 0.1, numItermax=5000, stopThr=1e-9, print_period=3,
                                                log=True)
    assert 0 < lr["n_iter"] < 4999 and lg["n_iter"] == lr["n_iter"]
    assert_vec(rg, rr, 1e-10)
    # not converged: the reference's warning
    with pytest.warns(UserWarning, match="did not converge"):
        ot.bregman.sinkhorn_stabilized(*to_dev(a, b, M), 0.1, numItermax=3, stopThr=1e-30)
    # numerical guard (:1319-1325): K underflows to 0 -> v = inf, u = nan at the first iteration
    with warnings.catch_warnings(record=True) as wr:
        warnings.simplefilter("always")
        _, lr = sorc.sinkhorn_stabilized(a, b, M * 1e6, 1e-3, numItermax=10, stopThr=0.0, log=True)
    with warnings.catch_warnings(record=True) as wg:
        warnings.simplefilter("always")
        _, lg = ot.bregman.sinkhorn_stabilized(*to_dev(a, b, M * 1e6), 1e-3, numItermax=10, stopThr=0.0, log=True)
    assert any("Numerical errors at iteration" in str(w.message) for w in wr)
    assert [str(w.message) for w in wg if "Numerical errors" in str(w.message)] == \
        [str(w.message) for w in wr if "Numerical errors" in str(w.message)]
    assert lg["n_iter"] == lr["n_iter"]


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_multi_step_cabi(dtype):
    """b2ot_multi_scaling_step / b2ot_multi_absorb against fp64 torch on a ragged, strided problem (several chunks of
    histograms, fewer rows than row blocks in one case)."""
    lib = _lib.load()
    gen = torch.Generator().manual_seed(3)
    for n, m, h in [(130, 77, 6), (9, 700, 1), (515, 300, 9)]:
        Kfull = (torch.rand(n, m + 5, generator=gen, dtype=torch.float64) + 0.05).to(dtype).to(DEV)
        K = Kfull[:, 2:m + 2]
        a = (torch.rand(n, generator=gen, dtype=torch.float64) + 0.1).to(dtype).to(DEV)
        B = (torch.rand(m, h, generator=gen, dtype=torch.float64) + 0.1).to(dtype).to(DEV)
        up = (torch.rand(n, h, generator=gen, dtype=torch.float64) + 0.1).to(dtype).to(DEV)
        vp = (torch.rand(m, h, generator=gen, dtype=torch.float64) + 0.1).to(dtype).to(DEV)
        un, vn = torch.full_like(up, float("nan")), torch.full_like(vp, float("nan"))
        stats = torch.zeros(16, dtype=torch.float64, device=DEV)
        scratch = torch.empty(lib.b2ot_multi_scratch_bytes(n, m, h), dtype=torch.uint8, device=DEV)
        rc = lib.b2ot_multi_scaling_step(_dense.dtype_code(dtype), _lib.MULTI_STATS, K.data_ptr(), n, m, K.stride(0), h,
                                         a.data_ptr(), B.data_ptr(), up.data_ptr(), vp.data_ptr(), un.data_ptr(),
                                         vn.data_ptr(), 0, 1, 0.0, 1.0, 1.0, stats.data_ptr(), None, 0,
                                         scratch.data_ptr(), _dense.stream_ptr(torch.device(DEV)))
        assert rc == 0, lib.b2ot_last_error()
        Kd = K.double()
        vref = B.double() / (Kd.T @ up.double())
        uref = a.double()[:, None] / (Kd @ vn.double())
        rt = 1e-12 if dtype == torch.float64 else 2e-6
        torch.testing.assert_close(vn.double(), vref, rtol=rt, atol=0)
        torch.testing.assert_close(un.double(), uref, rtol=rt, atol=0)
        want = torch.stack([un.abs().max(), vn.abs().max(), (un - up).abs().max(), up.abs().max(),
                            (vn - vp).abs().max(), vp.abs().max()]).double()
        torch.testing.assert_close(stats[:6], want, rtol=1e-15, atol=0)
        assert stats[6].item() == 0.0 and stats[7].item() == 0.0
        # a nan in the input shows up in the flags, the maxima skip it
        up2 = up.clone()
        up2[n // 2, 0] = float("nan")
        rc = lib.b2ot_multi_scaling_step(_dense.dtype_code(dtype), _lib.MULTI_STATS, K.data_ptr(), n, m, K.stride(0), h,
                                         a.data_ptr(), B.data_ptr(), up2.data_ptr(), vp.data_ptr(), un.data_ptr(),
                                         vn.data_ptr(), 0, 1, 0.0, 1.0, 1.0, stats.data_ptr(), None, 0,
                                         scratch.data_ptr(), _dense.stream_ptr(torch.device(DEV)))
        assert rc == 0
        assert stats[6].item() == 1.0 and stats[7].item() == 1.0 and bool(torch.isfinite(stats[:2]).all())
        # absorption
        al = torch.rand(n, generator=gen, dtype=torch.float64).to(DEV)
        be = torch.rand(m, generator=gen, dtype=torch.float64).to(DEV)
        al0, be0 = al.clone(), be.clone()
        rc = lib.b2ot_multi_absorb(_dense.dtype_code(dtype), n, m, h, up.data_ptr(), vp.data_ptr(), al.data_ptr(),
                                   be.data_ptr(), _dense.stream_ptr(torch.device(DEV)))
        assert rc == 0
        torch.testing.assert_close(al, al0 + torch.log(up.double()).amax(1), rtol=1e-14, atol=1e-15)
        torch.testing.assert_close(be, be0 + torch.log(vp.double()).max(), rtol=1e-14, atol=1e-15)
    # misuse: leading dimension, in-place update, solver mode without its state block
    def call(mode, ld, u_new, st, eh, cap):
        return lib.b2ot_multi_scaling_step(_dense.dtype_code(dtype), mode, K.data_ptr(), n, m, ld, h, a.data_ptr(),
                                           B.data_ptr(), up.data_ptr(), vp.data_ptr(), u_new.data_ptr(), vn.data_ptr(),
                                           0, 10, 0.0, 1.0, 1.0, st, eh, cap, scratch.data_ptr(), None)
    assert call(_lib.MULTI_STATS, m - 1, un, stats.data_ptr(), None, 0) != 0
    assert call(_lib.MULTI_STATS, K.stride(0), up, stats.data_ptr(), None, 0) != 0
    assert call(_lib.MULTI_KNOPP, K.stride(0), un, stats.data_ptr(), None, 0) != 0
    assert call(7, K.stride(0), un, stats.data_ptr(), None, 0) != 0
