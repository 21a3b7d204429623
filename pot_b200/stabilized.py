"""Log-stabilized scaling Sinkhorn and epsilon scaling, on the fused dense sweep.

Mirrors (same names, argument meaning, return values, warnings and errors):

    ot.bregman.sinkhorn_stabilized        ot/bregman/_sinkhorn.py:1105-1382   (1-D b: device loop; 2-D b: _stabilized_multi)
    ot.bregman.sinkhorn_epsilon_scaling   ot/bregman/_sinkhorn.py:1385-1573

The stabilized iteration is the balanced scaling iteration v = b / K^T u, u = a / K v (:1270-1271) on the kernel
K = exp(-(M - alpha - beta)/reg) (:1256-1258): ONE fused sweep over K per iteration, like sinkhorn_knopp.  What it adds
is the absorption (:1274-1288): when max|u| or max|v| exceeds tau, (u, v) are folded into the potentials and K is
rebuilt.  Because the scalings restart at 1/n, 1/m on a K that is then the current plan, the next update gives v ~ n:
after a first absorption every iteration absorbs as soon as n exceeds tau, so the absorption is part of the device
loop, not a host-side exception: the finalize kernel tests tau every iteration and
gated kernels in the enqueued iteration stream update alpha/reg, beta/reg, reset (u, v), rewrite K from Mr = -M/reg in
one pass that also produces the seeding column sums (`b2ot_dense_set_stabilized`).  Error (every print_period
iterations, :1303-1304), stop test `err <= stopThr` (:1316), numerical guard and iteration count live on the device;
the host only polls the status word.
Inputs must be torch CUDA tensors; there is no CPU path.
"""

import ctypes as C
import math
import warnings

import torch

from . import _autograd, _dense, _lib
from .bregman import _NOCONV, POLL_EVERY, _require_cuda_matrix, _rowmajor, _weights

__all__ = ["sinkhorn_stabilized", "sinkhorn_epsilon_scaling"]


def _stabilized_core(a, b, Mc, reg, numItermax, tau, stopThr, warmstart, print_period):
    """Runs the loop.  Returns (Mr, alpha/reg, beta/reg (float64), u, v, errs, n_iter, status, n_absorb)."""
    lib = _lib.load()
    n, m = Mc.shape
    dt, dev = Mc.dtype, Mc.device
    f64 = torch.float64  # the absorbed potentials live in double for both dtypes (include/b2ot.h)
    if warmstart is None:
        alpha_r = torch.zeros(n, dtype=f64, device=dev)
        beta_r = torch.zeros(m, dtype=f64, device=dev)
    else:
        alpha_r = (warmstart[0].detach().to(f64).reshape(-1) / reg).contiguous()
        beta_r = (warmstart[1].detach().to(f64).reshape(-1) / reg).contiguous()
    u0 = torch.full((n,), 1.0 / n, dtype=dt, device=dev)  # :1252-1254
    v0 = torch.full((m,), 1.0 / m, dtype=dt, device=dev)
    Mr = _dense.cost_transform(Mc, reg, "negscale")
    K = _dense.plan("log", Mr, alpha_r.to(dt), beta_r.to(dt))  # get_K (:1256-1258) = exp(Mr + alpha/reg + beta/reg)
    solver = _dense.DenseSolver("knopp", K, a, b, None, max_iter=numItermax, check_every=print_period,
                                stop_thr=stopThr)
    _lib.check(lib.b2ot_dense_set_stabilized(solver.h, float(tau), Mr.data_ptr(), Mr.stride(0), alpha_r.data_ptr(),
                                             beta_r.data_ptr()), "b2ot_dense_set_stabilized")
    status, niter, _ = solver.solve(u0, v0, poll_every=POLL_EVERY)
    u, v, errs = solver.export()
    cnt = C.c_int()
    _lib.check(lib.b2ot_dense_absorb_count(solver.h, C.byref(cnt)), "b2ot_dense_absorb_count")
    solver.close()
    return Mr, alpha_r, beta_r, u, v, errs, niter, status, cnt.value


def _stabilized_multi(a, b, Mc, reg, numItermax, tau, stopThr, warmstart, print_period):
    """2-D ``b`` (ot/bregman/_sinkhorn.py:1230-1234): H histograms on ONE kernel, joint absorption and joint error.

    The loop is the reference's (:1267-1325) with its per-iteration host decision (:1274): every iteration is one
    `b2ot_multi_scaling_step` (two passes over K for all histograms + the statistics of the host's tests, four
    launches) and one 64-byte read; an absorption is `b2ot_multi_absorb` + a rewrite of K from Mr = -M/reg.
    Returns (Mr, alpha/reg, beta/reg (float64), u, v (dim x H), errs, n_iter, status).
    """
    lib = _lib.load()
    n, m = Mc.shape
    H = b.shape[1]
    dt, dev = Mc.dtype, Mc.device
    code = _dense.dtype_code(dt)
    f64 = torch.float64
    if warmstart is None:
        alpha_r = torch.zeros(n, dtype=f64, device=dev)
        beta_r = torch.zeros(m, dtype=f64, device=dev)
    else:
        alpha_r = (warmstart[0].detach().to(f64).reshape(-1) / reg).contiguous()
        beta_r = (warmstart[1].detach().to(f64).reshape(-1) / reg).contiguous()
    u = torch.full((n, H), 1.0 / n, dtype=dt, device=dev)  # :1250-1251
    v = torch.full((m, H), 1.0 / m, dtype=dt, device=dev)
    un, vn = torch.empty_like(u), torch.empty_like(v)
    Mr = _dense.cost_transform(Mc, reg, "negscale")
    K = _dense.plan("log", Mr, alpha_r.to(dt), beta_r.to(dt))  # get_K, :1256-1258
    stats = torch.zeros(16, dtype=f64, device=dev)
    stats_host = torch.zeros(16, dtype=f64).pin_memory()
    scratch = torch.empty(lib.b2ot_multi_scratch_bytes(n, m, H), dtype=torch.uint8, device=dev)
    stream = _dense.stream_ptr(dev)
    nan = float("nan")
    errs, err, ii, status = [], 1.0, 0, _lib.ST_MAXITER
    for ii in range(numItermax):
        _lib.check(lib.b2ot_multi_scaling_step(code, _lib.MULTI_STATS, K.data_ptr(), n, m, K.stride(0), H,
                                               a.data_ptr(), b.data_ptr(), u.data_ptr(), v.data_ptr(), un.data_ptr(),
                                               vn.data_ptr(), ii, 1, 0.0, 1.0, 1.0, stats.data_ptr(), None, 0,
                                               scratch.data_ptr(), stream), "b2ot_multi_scaling_step")
        stats_host.copy_(stats, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        mu, mv, du, mup, dv, mvp, nan_u, nan_v = stats_host[:8].tolist()
        if nan_u:  # numpy's max propagates nan
            mu = du = nan
        if nan_v:
            mv = dv = nan
        uprev, vprev, u, v = u, v, un, vn
        un, vn = uprev, vprev  # the next step writes over the buffers of the iteration before this one
        if mu > tau or mv > tau:  # :1274-1288, n_hists branch: joint absorption, the scalings are kept
            _lib.check(lib.b2ot_multi_absorb(code, n, m, H, u.data_ptr(), v.data_ptr(), alpha_r.data_ptr(),
                                             beta_r.data_ptr(), stream), "b2ot_multi_absorb")
            _dense.plan("log", Mr, alpha_r.to(dt), beta_r.to(dt), out=K)
        if ii % print_period == 0:  # :1297-1301
            err = 0.5 * (du / max(mu, mup, 1.0) + dv / max(mv, mvp, 1.0))
            errs.append(err)
        if err <= stopThr:  # :1316
            status = _lib.ST_CONVERGED
            break
        if nan_u or nan_v:  # :1319-1325
            u, v = uprev, vprev
            status = _lib.ST_NUMERR
            break
    return Mr, alpha_r, beta_r, u, v, errs, ii, status


def _multi_result(a, b, Mc, reg, numItermax, tau, stopThr, warmstart, verbose, print_period, log, warn):
    """sinkhorn_stabilized for a 2-D b: the losses <M, Gamma_h> (:1342-1348) and the log dict of :1332-1341."""
    dt, dev = Mc.dtype, Mc.device
    with torch.cuda.device(dev):
        Mr, alpha_r, beta_r, u, v, errs, niter, status = _stabilized_multi(
            a, b, Mc, reg, int(numItermax), float(tau), float(stopThr), warmstart, int(print_period))
        if status == _lib.ST_NUMERR:
            warnings.warn("Numerical errors at iteration %d" % niter)
        elif status == _lib.ST_MAXITER and warn:
            warnings.warn(_NOCONV)
        if verbose:
            for k, e in enumerate(errs):
                ii = k * int(print_period)
                if ii % (int(print_period) * 20) == 0:
                    print("{:5s}|{:12s}".format("It.", "Err") + "\n" + "-" * 19)
                print("{:5d}|{:8e}|".format(ii, e))
        logu = (alpha_r[:, None] + torch.log(u.double())).to(dt)  # :1333-1336
        logv = (beta_r[:, None] + torch.log(v.double())).to(dt)
        res = torch.stack([_dense.loss("log", Mr, Mc, logu[:, h].contiguous(), logv[:, h].contiguous())
                           for h in range(b.shape[1])]).to(dt)
    if not log:
        return res
    alpha_out, beta_out = reg * logu, reg * logv
    return res, {"err": [torch.tensor(e, dtype=dt, device=dev) for e in errs], "n_iter": niter, "logu": logu,
                 "logv": logv, "alpha": alpha_out, "beta": beta_out, "warmstart": (alpha_out, beta_out)}


def sinkhorn_stabilized(a, b, M, reg, numItermax=1000, tau=1e3, stopThr=1e-9, warmstart=None, verbose=False,
                        print_period=20, log=False, warn=True, **kwargs):
    """Drop-in for ot.bregman.sinkhorn_stabilized (ot/bregman/_sinkhorn.py:1105-1382).

    Returns the plan (1-D ``b``) or the losses per histogram (2-D ``b``, :1342-1348); with ``log=True`` also
    {"err", "n_iter", "logu", "logv", "alpha", "beta", "warmstart"} (:1332-1340).
    """
    if not kwargs.get("b2ot_loss", False):
        _autograd.warn_detached("sinkhorn_stabilized", M, a, b)
    M = _require_cuda_matrix(M)
    Mc = _rowmajor(M.detach())
    a = _weights(a, Mc.shape[0], Mc, "a").detach()
    b = _weights(b, Mc.shape[1], Mc, "b").detach()
    if b.dim() > 1:
        return _multi_result(a, b, Mc, reg, numItermax, tau, stopThr, warmstart, verbose, print_period, log, warn)
    dt, dev = Mc.dtype, Mc.device
    with torch.cuda.device(dev):
        Mr, alpha_r, beta_r, u, v, errs, niter, status, _ = _stabilized_core(
            a, b, Mc, reg, int(numItermax), float(tau), float(stopThr), warmstart, int(print_period))
        if status == _lib.ST_NUMERR:
            warnings.warn("Numerical errors at iteration %d" % niter)  # :1322
        elif status == _lib.ST_MAXITER and warn:
            warnings.warn(_NOCONV)
        if verbose:
            for k, e in enumerate(errs):  # :1308-1313
                ii = k * int(print_period)
                if ii % (int(print_period) * 20) == 0:
                    print("{:5s}|{:12s}".format("It.", "Err") + "\n" + "-" * 19)
                print("{:5d}|{:8e}|".format(ii, e))
        logu = (alpha_r + torch.log(u.double())).to(dt)  # :1335-1336
        logv = (beta_r + torch.log(v.double())).to(dt)
        if kwargs.get("b2ot_loss", False):  # sinkhorn2: <M, Gamma> reduced straight from (Mr, M), no plan written
            G = _dense.loss("log", Mr, Mc, logu, logv).to(dt)
        else:
            G = _dense.plan("log", Mr, logu, logv)  # get_Gamma, :1260-1266
    if not log:
        return G
    alpha_out = reg * logu  # alpha + reg log u, :1339-1340
    beta_out = reg * logv
    return G, {"err": [torch.tensor(e, dtype=dt, device=dev) for e in errs], "n_iter": niter, "logu": logu,
               "logv": logv, "alpha": alpha_out, "beta": beta_out, "warmstart": (alpha_out, beta_out)}


def sinkhorn_epsilon_scaling(a, b, M, reg, numItermax=100, epsilon0=1e4, numInnerItermax=100, tau=1e3, stopThr=1e-9,
                             warmstart=None, verbose=False, print_period=10, log=False, warn=True, **kwargs):
    """Drop-in for ot.bregman.sinkhorn_epsilon_scaling (ot/bregman/_sinkhorn.py:1385-1573).

    Outer loop on the host (at most numItermax stabilized solves, reg_i = (epsilon0 - reg) exp(-i) + reg, :1509-1510);
    every inner solve is the device loop above; the outer error ||G^T 1 - b||^2 + ||G 1 - a||^2 (:1536-1541) comes from
    the epilogue kernel's row marginal and the column-sum kernel.
    """
    M = _require_cuda_matrix(M)
    Mc = _rowmajor(M.detach())
    a = _weights(a, Mc.shape[0], Mc, "a")
    b = _weights(b, Mc.shape[1], Mc, "b")
    if b.dim() > 1:
        raise NotImplementedError("pot_b200: sinkhorn_epsilon_scaling needs a 1-D b")
    dt, dev = Mc.dtype, Mc.device
    n, m = Mc.shape
    num_min = 35  # :1495-1496
    numItermax = max(num_min, int(numItermax))
    if warmstart is None:
        alpha = torch.zeros(n, dtype=dt, device=dev)
        beta = torch.zeros(m, dtype=dt, device=dev)
    else:
        alpha, beta = warmstart[0].detach().to(dt), warmstart[1].detach().to(dt)
    errs = []
    err = 1.0
    G = None
    converged = False
    ii = 0
    with torch.cuda.device(dev):
        for ii in range(numItermax):
            regi = (epsilon0 - reg) * math.exp(-ii) + reg
            Mr, al_r, be_r, u, v, _, _, _, _ = _stabilized_core(a, b, Mc, regi, int(numInnerItermax), float(tau),
                                                                float(stopThr), (alpha, beta), 20)
            logu = (al_r + torch.log(u.double())).to(dt)
            logv = (be_r + torch.log(v.double())).to(dt)
            check = ii % int(print_period) == 0
            alpha = regi * logu  # logi["alpha"], logi["beta"] (:1339-1340, :1528-1529)
            beta = regi * logv
            G = None
            if check:  # the stage's plan (:1514-1526) is only READ here and after the loop
                G, _, rows = _dense.plan_value("log", logu, logv, K=Mr, want_rowsum=True)
                cols = _dense.colsum(G)
                err = float((torch.linalg.norm(cols - b) ** 2 + torch.linalg.norm(rows - a) ** 2).item())
                errs.append(err)
                if verbose:
                    if ii % (int(print_period) * 10) == 0:
                        print("{:5s}|{:12s}".format("It.", "Err") + "\n" + "-" * 19)
                    print("{:5d}|{:8e}|".format(ii, err))
            if err <= stopThr and ii > num_min:
                converged = True
                break
        if G is None:
            G = _dense.plan("log", Mr, logu, logv)
    if not converged and warn:
        warnings.warn(_NOCONV)
    if not log:
        return G
    return G, {"err": [torch.tensor(e, dtype=dt, device=dev) for e in errs], "alpha": alpha, "beta": beta,
               "warmstart": (alpha, beta), "niter": ii}
