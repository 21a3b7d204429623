"""BASELINE config 3 at FULL size (n = m = 65536 fp32, K = 16 GiB) on one B200: size-independent properties.

The numpy oracle cannot run at this size (64 GiB of temporaries, seconds per iteration), so parity is checked through
properties (pitfall 3):
  * the fused single-pass sweep and the independent two-pass split implementation agree on (u, v);
  * u (K v) = a, i.e. the row marginal of the plan is exact after the last (u) update -- recomputed in fp64 by a
    chunked torch restatement that shares no code with the kernels;
  * the reported error equals the recomputed column-marginal violation ||v (K^T u) - b||_2 of the iterate it was
    measured on;
  * the unbalanced iteration (same sweep kernel, row exponent) agrees between the fused and the split path.
Everything goes through libb2ot.so (C ABI) via pot_b200.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from pot_b200 import _dense, _lib

DEV = "cuda:0"
N = 65536


def _need_memory(gib):
    free, _ = torch.cuda.mem_get_info()
    if free < gib * 2 ** 30:
        pytest.skip(f"needs {gib} GiB of free device memory")


@pytest.fixture(scope="module")
def gibbs_kernel():
    """K = exp(-M/reg) for config 3 (2-D Gaussian clouds, M / max M, reg = 0.05), built in place over M."""
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    _need_memory(60)
    g = torch.Generator(device=DEV).manual_seed(0)
    xs = torch.randn(N, 2, device=DEV, generator=g)
    xt = torch.randn(N, 2, device=DEV, generator=g) + 1.0
    M = torch.cdist(xs, xt).pow_(2)
    M /= M.max()
    lib = _lib.load()
    _lib.check(lib.b2ot_kexp(M.data_ptr(), M.data_ptr(), N, N, M.stride(0), M.stride(0), 0.05, _lib.F32,
                             _dense.stream_ptr(M.device)), "b2ot_kexp")
    torch.cuda.synchronize()
    yield M
    del M
    torch.cuda.empty_cache()


def _matvec64(K, x, transpose=False, chunk=2048):
    """fp64 K x or K^T x in row chunks (checker only)."""
    x = x.double()
    if not transpose:
        return torch.cat([K[i:i + chunk].double() @ x for i in range(0, K.shape[0], chunk)])
    out = torch.zeros(K.shape[1], dtype=torch.float64, device=K.device)
    for i in range(0, K.shape[0], chunk):
        out += K[i:i + chunk].double().t() @ x[i:i + chunk]
    return out


def _run(K, a, b, variant, its, unbalanced=None):
    lib = _lib.load()
    s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=its, check_every=1 if unbalanced else 10, stop_thr=0.0,
                           variant=variant)
    if unbalanced:
        _lib.check(lib.b2ot_dense_set_unbalanced(s.h, *unbalanced), "b2ot_dense_set_unbalanced")
        u0, v0 = torch.ones_like(a), torch.ones_like(b)
    else:
        u0, v0 = a.clone(), b.clone()
    status, niter, _ = s.solve(u0, v0, poll_every=0)
    u, v, errs = s.export()
    info = s.info()
    s.close()
    return u, v, errs, niter, status, info


def test_config3_fused_vs_split_and_marginals(gibbs_kernel):
    K = gibbs_kernel
    a = torch.full((N,), 1.0 / N, device=DEV)
    b = torch.full((N,), 1.0 / N, device=DEV)
    its = 11  # error recorded at ii = 0 and ii = 10 (the last iteration)
    u1, v1, e1, n1, st1, info = _run(K, a, b, "fused", its)
    assert info["variant"] == "fused" and info["cluster_size"] == 8
    u2, v2, e2, n2, st2, info2 = _run(K, a, b, "split", its)
    assert info2["variant"] == "split"
    assert st1 == st2 == _lib.ST_MAXITER and n1 == n2 == its - 1
    assert ((u1 - u2).abs() / u2.abs()).max().item() < 1e-4
    assert ((v1 - v2).abs() / v2.abs()).max().item() < 1e-4
    assert torch.isfinite(u1).all() and torch.isfinite(v1).all()
    # row marginal of diag(u) K diag(v) equals a (u was updated last), fp64 restatement
    row = u1.double() * _matvec64(K, v1)
    assert ((row - a.double()).abs() / a.double()).max().item() < 1e-4
    # the last check (ii = 10) is measured on the final iterate (u_10, v_10): recompute it in fp64; the two code paths
    # must report the same history
    assert len(e1) == len(e2) == 2
    np.testing.assert_allclose(e1, e2, rtol=1e-3, atol=1e-9)
    col = v1.double() * _matvec64(K, u1, transpose=True)
    final_err = (col - b.double()).norm().item()
    assert abs(final_err - e1[-1]) <= 1e-2 * final_err + 1e-9, (final_err, e1)
    assert e1[-1] < e1[0]


def test_config3_unbalanced_fused_vs_split(gibbs_kernel):
    K = gibbs_kernel
    a = torch.full((N,), 1.3 / N, device=DEV)
    b = torch.full((N,), 1.0 / N, device=DEV)
    fi = (1.0 / 1.05, 0.5 / 0.55)  # reg_m = (1, 0.5), reg = 0.05
    u1, v1, e1, n1, st1, _ = _run(K, a, b, "fused", 8, unbalanced=fi)
    u2, v2, e2, n2, st2, _ = _run(K, a, b, "split", 8, unbalanced=fi)
    assert st1 == st2 == _lib.ST_MAXITER and n1 == n2 == 7
    assert len(e1) == len(e2) == 8
    assert ((u1 - u2).abs() / u2.abs()).max().item() < 1e-4
    assert ((v1 - v2).abs() / v2.abs()).max().item() < 1e-4
    np.testing.assert_allclose(e1, e2, rtol=1e-3, atol=1e-7)
    # fixed-point residual of the last half-step in fp64: v = (b / K^T u)^fi_2
    c = _matvec64(K, u1, transpose=True)
    v_ref = (b.double() / c) ** fi[1]
    assert ((v1.double() - v_ref).abs() / v_ref).max().item() < 1e-4


def test_config3_stabilized_absorbing_iterations():
    """sinkhorn_stabilized's device loop at 65536^2 fp32 with tau below the scalings (1e-7), so that every iteration
    absorbs: fused sweep + K rebuild with 2^32-element index arithmetic.  Properties, checked by a chunked fp64 torch restatement:
      * every iteration absorbed;
      * exp(Mr + alpha/reg + beta/reg) -- the plan of the last iterate, u updated last -- has row marginal a;
      * the rebuilt K buffer equals exp(Mr + alpha/reg + beta/reg) scaled by nothing else (spot rows);
      * its column marginal violation equals the error of the last check un-scaled by the reference's 1/(nm) quirk."""
    _need_memory(70)
    from pot_b200 import stabilized

    g = torch.Generator(device=DEV).manual_seed(1)
    xs = torch.randn(N, 2, device=DEV, generator=g)
    xt = torch.randn(N, 2, device=DEV, generator=g) + 1.0
    M = torch.cdist(xs, xt).pow_(2)
    M /= M.max()
    a = torch.full((N,), 1.0 / N, device=DEV)
    b = torch.full((N,), 1.0 / N, device=DEV)
    reg, its = 0.05, 6
    Mr, alpha_r, beta_r, u, v, errs, niter, status, n_absorb = stabilized._stabilized_core(
        a, b, M, reg, its, 1e-7, -1.0, None, 1)
    del M
    assert status == _lib.ST_MAXITER and niter == its - 1 and n_absorb == its and len(errs) == its
    assert torch.allclose(u, torch.full_like(u, 1.0 / N)) and torch.allclose(v, torch.full_like(v, 1.0 / N))
    assert torch.isfinite(alpha_r).all() and torch.isfinite(beta_r).all()
    row = torch.empty(N, dtype=torch.float64, device=DEV)
    col = torch.zeros(N, dtype=torch.float64, device=DEV)
    chunk = 2048
    for i in range(0, N, chunk):
        P = torch.exp(Mr[i:i + chunk].double() + alpha_r[i:i + chunk, None] + beta_r[None, :])
        row[i:i + chunk] = P.sum(1)
        col += P.sum(0)
    assert ((row - a.double()).abs() / a.double()).max().item() < 1e-4
    # reported error of the last (absorbing) check: || col / (n m) - b ||_2 (reference quirk, :1283-1287 then :1303)
    ref_err = (col / (float(N) * float(N)) - b.double()).norm().item()
    assert abs(errs[-1] - ref_err) <= 1e-3 * ref_err
    # and the plan really is close to feasible: un-scaled column violation small and decreasing over the run
    assert (col - b.double()).norm().item() < 1e-4


# ------------------------------------------------------------------------------------------------------------------
# Config 3: the reference loop itself, restated in chunked fp64 torch (ot/bregman/_sinkhorn.py:626-631:
# KtransposeU = K^T u; v = b / KtransposeU; u = 1 / (Kp v)), 10 full iterations at 65536^2, compared on (u, v).
# ------------------------------------------------------------------------------------------------------------------
def test_config3_ten_iterations_vs_fp64_restatement(gibbs_kernel):
    K = gibbs_kernel
    a = torch.full((N,), 1.0 / N, device=DEV)
    b = torch.full((N,), 1.0 / N, device=DEV)
    its = 10
    u1, v1, e1, n1, st1, info = _run(K, a, b, "fused", its)
    assert st1 == _lib.ST_MAXITER and n1 == its - 1 and info["variant"] == "fused"
    u = torch.full((N,), 1.0 / N, dtype=torch.float64, device=DEV)  # reference init (:585-590)
    a64, b64 = a.double(), b.double()
    err0 = None
    for ii in range(its):
        ktu = _matvec64(K, u, transpose=True)
        v = b64 / ktu
        u = a64 / _matvec64(K, v)
        if ii == 0:  # :647-657, the check of ii % 10 == 0: tmp2 = einsum("i,ij,j->j", u, K, v)
            err0 = (v * _matvec64(K, u, transpose=True) - b64).norm().item()
    assert ((u1.double() - u).abs() / u.abs()).max().item() < 1e-4
    assert ((v1.double() - v).abs() / v.abs()).max().item() < 1e-4
    assert len(e1) == 1 and abs(e1[0] - err0) <= 1e-3 * err0 + 1e-9, (e1, err0)


# ------------------------------------------------------------------------------------------------------------------
# Golden vectors of the UNMODIFIED reference at the BASELINE sizes (tests/golden/make_golden_fullsize.py;
# inputs regenerated from the seeds by tests/fullsize_inputs.py).
# ------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def gfull():
    import os

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_fullsize.npz")
    return np.load(path, allow_pickle=False)


def _maxrel_log(x, ref):
    x, ref = np.asarray(x, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    return np.abs(x - ref).max() / max(1.0, np.abs(ref).max())


def test_config2_golden_16384(gfull):
    """BASELINE config 2: ot.bregman.sinkhorn_log, n = m = 16384 fp32, reg 0.01 -- the fp32 ping-pong sweep against the
    reference's own (log_u, log_v, err) after 20 fixed iterations, 1e-4 relative (north_star's fp32 tolerance)."""
    import warnings

    import pot_b200 as ot
    import fullsize_inputs as fin

    a, b, M = fin.config2_inputs()
    Md = torch.from_numpy(M).to(DEV)
    del M
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G, lg = ot.bregman.sinkhorn_log(torch.from_numpy(a).to(DEV), torch.from_numpy(b).to(DEV), Md, fin.C2_REG,
                                        numItermax=fin.C2_ITS, stopThr=0, log=True, warn=False)
    assert lg["niter"] == int(gfull["c2/niter"])
    assert _maxrel_log(lg["log_u"].cpu().numpy(), gfull["c2/log_u"]) < 1e-4
    assert _maxrel_log(lg["log_v"].cpu().numpy(), gfull["c2/log_v"]) < 1e-4
    e, r = np.asarray([float(x) for x in lg["err"]]), gfull["c2/err"]
    assert e.shape == r.shape and np.all(np.abs(e - r) <= 5e-3 * r + 2e-6), (e, r)
    # marginals of the returned plan: u was updated last, so rows are exact to fp32 round-off
    assert float((G.sum(1) - 1.0 / fin.C2_N).abs().max()) < 1e-4 / fin.C2_N


@pytest.mark.parametrize("operand", ["f16", "tf32"])
def test_config4_golden_16384(gfull, operand):
    """BASELINE config 4's path (online cost, log domain, d = 64, fp32, 128 x 128 tcgen05 tiles, several reduction
    splits) at n = m = 16384 against the reference's own (f, g, err) after 10 fixed iterations; both operand splits."""
    import warnings

    import pot_b200 as ot
    import fullsize_inputs as fin

    X, Y = fin.config4_inputs()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g, lg = ot.empirical_sinkhorn(torch.from_numpy(X).to(DEV), torch.from_numpy(Y).to(DEV), fin.C4_REG,
                                         numIterMax=fin.C4_ITS, stopThr=-1.0, isLazy=True, log=True, warn=False,
                                         b2ot_variant="tc", b2ot_operand=operand)
    assert _maxrel_log(f.cpu().numpy(), gfull["c4/f"]) < 1e-4
    assert _maxrel_log(g.cpu().numpy(), gfull["c4/g"]) < 1e-4
    e, r = np.asarray([float(x) for x in lg["err"]]), gfull["c4/err"]
    assert e.shape == r.shape and np.all(np.abs(e - r) <= 5e-3 * r + 2e-5), (e, r)


@pytest.mark.parametrize("operand", ["f16", "tf32"])
def test_config4_65536_vs_fp64(operand):
    """The tcgen05 sweep at n = m = 65536, d = 64 (multi-wave grid, 4-7 reduction splits) against (i) the library's
    fp64 register-tiled sweep on the same clouds and (ii) one half-iteration restated in chunked fp64 torch
    (ot/bregman/_empirical.py:218-233: f = -logsumexp(g - C/reg) + log a)."""
    import warnings

    import pot_b200 as ot

    n, d, reg, its = 65536, 64, 0.1, 4
    gen = torch.Generator(device=DEV).manual_seed(41)
    X = torch.randn(n, d, device=DEV, generator=gen) / d ** 0.5
    Y = torch.randn(n, d, device=DEV, generator=gen) / d ** 0.5 + 1.0 / d ** 0.5
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        f, g = ot.empirical_sinkhorn(X, Y, reg, numIterMax=its, stopThr=-1.0, isLazy=True, warn=False,
                                     b2ot_variant="tc", b2ot_operand=operand)
        f64, g64 = ot.empirical_sinkhorn(X.double(), Y.double(), reg, numIterMax=its, stopThr=-1.0, isLazy=True,
                                         warn=False, b2ot_variant="simt")
    assert _maxrel_log(f.cpu().numpy(), f64.cpu().numpy()) < 1e-4
    assert _maxrel_log(g.cpu().numpy(), g64.cpu().numpy()) < 1e-4
    # (ii) the last half-step of the solve (g from f) in chunked fp64 torch, from the fp32 f the kernel produced
    Xd, Yd = X.double(), Y.double()
    xx, yy = (Xd * Xd).sum(1), (Yd * Yd).sum(1)
    logb = -np.log(n)
    g_ref = torch.empty(n, dtype=torch.float64, device=DEV)
    fd = f.double()
    for j0 in range(0, n, 2048):
        C = (yy[j0:j0 + 2048, None] + xx[None, :] - 2.0 * Yd[j0:j0 + 2048] @ Xd.t()).clamp_(min=0)
        g_ref[j0:j0 + 2048] = logb - torch.logsumexp(fd[None, :] - C / reg, dim=1)
    assert _maxrel_log(g.cpu().numpy(), g_ref.cpu().numpy()) < 1e-4


@pytest.mark.parametrize("solver", ["log_sinkhorn", "sinkhorn"])
def test_config5_golden_full(gfull, solver):
    """BASELINE config 5 at FULL size (4096 problems of 256 x 256 fp32): ot.batch.solve_batch against the reference's own
    values for every problem and potentials of the first 8, 20 fixed iterations."""
    import pot_b200 as ot
    import fullsize_inputs as fin

    Mb = torch.from_numpy(fin.config5_inputs()).to(DEV)
    res = ot.batch.solve_batch(Mb, fin.C5_REG, max_iter=fin.C5_ITS, tol=0, solver=solver, reg_type="entropy")
    for key, got in (("value", res.value), ("value_linear", res.value_linear)):
        ref = gfull[f"c5/{solver}/{key}"].astype(np.float64)
        err = np.abs(got.cpu().numpy().astype(np.float64) - ref).max() / np.abs(ref).max()
        assert err < 1e-4, (key, err)
    fh, gh = res.potentials[0][:8].cpu().numpy(), res.potentials[1][:8].cpu().numpy()
    rf, rg = gfull[f"c5/{solver}/f_head"].astype(np.float64), gfull[f"c5/{solver}/g_head"].astype(np.float64)
    # same start (zeros / ones) and same update order as the reference: the potentials agree without a gauge shift
    if solver == "log_sinkhorn":
        assert _maxrel_log(fh, rf) < 1e-4 and _maxrel_log(gh, rg) < 1e-4
    else:  # scalings (f, g) of bregman_projection_batch, ot/batch/_utils.py:138-160
        assert (np.abs(fh - rf) / np.abs(rf)).max() < 1e-4 and (np.abs(gh - rg) / np.abs(rg)).max() < 1e-4
