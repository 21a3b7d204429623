"""Unbalanced entropic OT in the scaling domain, on the same fused sweep kernel as the balanced solver.

Mirrors (same names, argument meaning, return values, warnings and errors):

    ot.unbalanced.sinkhorn_knopp_unbalanced   ot/unbalanced/_sinkhorn.py:565-818
    ot.unbalanced.sinkhorn_unbalanced         ot/unbalanced/_sinkhorn.py:19-249   (method "sinkhorn")

The reference iteration is u = (a / Kv)^fi_1, v = (b / K^T u)^fi_2 (:756-759) -- already in the u-then-K^T u order of the
fused sweep, so one iteration is ONE pass over K (the reference makes two GEMVs): the sweep kernel raises the row
scaling to fi_1, the finalize kernel does the v update, the numerical guard, the error
(|u - uprev|_inf / max(|u|_inf, |uprev|_inf, 1) + same for v) / 2 every iteration and the stop test on the device.
K = exp(-M/reg) * c is built once by the epilogue kernel (the plan of u = v = 1); the same kernel produces the plan, the
linear cost, KL(plan, c) and the row marginal at the end; the column marginal is one more reduction kernel.
Inputs must be torch CUDA tensors; there is no CPU path.
"""

import warnings

import torch

from . import _dense, _lib
from .bregman import _require_cuda_matrix, _rowmajor, _weights

__all__ = ["sinkhorn_knopp_unbalanced", "sinkhorn_unbalanced"]

POLL_EVERY = 50


def _parameter_pair(parameter):
    """ot.utils.get_parameter_pair (ot/utils.py:846-883)."""
    if isinstance(parameter, (float, int)):
        return parameter, parameter
    if len(parameter) == 1:
        return parameter[0], parameter[0]
    if len(parameter) > 2:
        raise ValueError("Parameter must be either a scalar, or an indexable object of length 1 or 2.")
    return parameter[0], parameter[1]


def _kl_vec(p, q):
    """nx.kl_div(p, q) for marginal vectors (ot/backend.py:2594-2598): n- or m-sized, host-side glue."""
    return torch.sum(p * torch.log(p / q + 1e-16))


def _solve_one(K, a, b, fi_1, fi_2, u0, v0, numItermax, stopThr):
    lib = _lib.load()
    solver = _dense.DenseSolver("knopp", K, a, b, None, max_iter=numItermax, check_every=1, stop_thr=stopThr)
    _lib.check(lib.b2ot_dense_set_unbalanced(solver.h, float(fi_1), float(fi_2)), "b2ot_dense_set_unbalanced")
    status, niter, _ = solver.solve(u0, v0, poll_every=POLL_EVERY)
    u, v, errs = solver.export()
    solver.close()
    return u, v, errs, niter, status


def sinkhorn_knopp_unbalanced(a, b, M, reg, reg_m, reg_type="kl", c=None, warmstart=None, numItermax=1000,
                              stopThr=1e-6, verbose=False, log=False, **kwargs):
    """Drop-in for ot.unbalanced.sinkhorn_knopp_unbalanced (ot/unbalanced/_sinkhorn.py:565-818).

    Returns the plan (1-D b) or the per-histogram losses (2-D b); with log=True also the dict
    {"err", "logu", "logv"} (+ "cost", "total_cost" for 1-D b).
    """
    M = _require_cuda_matrix(M)
    Mc = _rowmajor(M.detach())
    dim_a, dim_b = Mc.shape
    a = _weights(a, dim_a, Mc, "a")
    b = _weights(b, dim_b, Mc, "b")
    n_hists = b.shape[1] if b.dim() > 1 else 0
    reg_m1, reg_m2 = _parameter_pair(reg_m)
    fi_1 = reg_m1 / (reg_m1 + reg) if reg_m1 != float("inf") else 1  # :743-744
    fi_2 = reg_m2 / (reg_m2 + reg) if reg_m2 != float("inf") else 1
    dt, dev = Mc.dtype, Mc.device
    one_n = torch.ones(dim_a, dtype=dt, device=dev)
    one_m = torch.ones(dim_b, dtype=dt, device=dev)

    entropy = reg_type.lower() == "entropy"
    if entropy:
        warnings.warn("If reg_type = entropy, then the matrix c is overwritten by the one matrix.")  # :729-733
    if c is not None and not entropy and not n_hists:
        c = _rowmajor(_require_cuda_matrix(c, "c").detach().to(dt))
    else:
        c = None

    with torch.cuda.device(dev):
        # K = exp(-M/reg) [* c] (:736-740): the epilogue kernel with u = v = 1 (one pass, M read once)
        if n_hists or entropy:
            K = _dense.cost_transform(Mc, reg, "kexp")
        else:
            K, _, _ = _dense.plan_value("knopp", one_n, one_m, M=Mc, reg=reg, a=a, b=b, c=c, kmulq=True)

        if n_hists:
            return _multi_hist(a, b, Mc, K, fi_1, fi_2, warmstart, numItermax, stopThr, verbose, log)

        if warmstart is None:
            u0, v0 = one_n, one_m
        else:
            u0 = torch.exp(warmstart[0].detach().to(dt)).contiguous()
            v0 = torch.exp(warmstart[1].detach().to(dt)).contiguous()
        u, v, errs, niter, status = _solve_one(K, a, b, fi_1, fi_2, u0, v0, int(numItermax), float(stopThr))
        if status == _lib.ST_NUMERR:
            warnings.warn("Numerical errors at iteration %s" % niter)  # :768
        if log and verbose:
            _print_errs(errs)
        if not log:
            return _dense.plan("knopp", K, u, v)
        # plan, <plan, M>, KL(plan, c) and plan 1 in one pass over (K, M); c = a b^T unless given / entropy (ones)
        ones_a, ones_b = (one_n, one_m) if entropy else (a, b)
        plan, vals, rows = _dense.plan_value("knopp", u, v, K=K, M=Mc, a=ones_a, b=ones_b, c=c,
                                             terms=("linear", "kl"), want_rowsum=True)
        cost = vals[0].to(dt)
        total = vals[0] + reg * vals[2]
        if reg_m1 != float("inf"):
            total = total + reg_m1 * _kl_vec(rows, a).double()
        if reg_m2 != float("inf"):
            total = total + reg_m2 * _kl_vec(_dense.colsum(plan), b).double()
    dict_log = {"err": [torch.tensor(e, dtype=dt, device=dev) for e in errs],
                "logu": torch.log(u + 1e-300), "logv": torch.log(v + 1e-300), "cost": cost,
                "total_cost": total.to(dt)}
    return plan, dict_log


def _print_errs(errs):
    for i, e in enumerate(errs):  # :778-783
        if i % 50 == 0:
            print("{:5s}|{:12s}".format("It.", "Err") + "\n" + "-" * 19)
        print("{:5d}|{:8e}|".format(i, e))


def _multi_hist(a, b, M, K, fi_1, fi_2, warmstart, numItermax, stopThr, verbose, log):
    """2-D b: losses only (:789-795).  All histograms iterate on ONE K = exp(-M/reg) with the joint guard and the joint
    error / stop test of the reference loop (:749-787): `_dense.multi_hist_solve` (u-then-v with the exponents fi,
    two passes over K per iteration whatever the number of histograms, decisions on the device)."""
    dt, dev = M.dtype, M.device
    n, m = M.shape
    n_h = b.shape[1]
    if warmstart is None:
        u0 = torch.ones((n, n_h), dtype=dt, device=dev)  # :713-715 (the (dim_a, 1) column broadcasts)
        v0 = torch.ones((m, n_h), dtype=dt, device=dev)
    else:
        u0 = torch.exp(warmstart[0].detach().to(dt)).reshape(n, -1).expand(n, n_h)
        v0 = torch.exp(warmstart[1].detach().to(dt)).reshape(m, -1).expand(m, n_h)
    u, v, errs, niter, status = _dense.multi_hist_solve("unbalanced", K, a, b.contiguous(), u0, v0, int(numItermax),
                                                        float(stopThr), fi=(fi_1, fi_2), poll_every=POLL_EVERY)
    if status == _lib.ST_NUMERR:
        warnings.warn("Numerical errors at iteration %s" % niter)
    if log and verbose:
        _print_errs(errs)
    losses = []
    for k in range(n_h):  # einsum("ik,ij,jk,ij->k", u, K, v, M)
        _, vals, _ = _dense.plan_value("knopp", u[:, k].contiguous(), v[:, k].contiguous(), K=K, M=M, terms=("linear",),
                                       want_plan=False)
        losses.append(vals[0].to(dt))
    res = torch.stack(losses)
    if not log:
        return res
    return res, {"err": [torch.tensor(e, dtype=dt, device=dev) for e in errs], "logu": torch.log(u + 1e-300),
                 "logv": torch.log(v + 1e-300)}


def sinkhorn_unbalanced(a, b, M, reg, reg_m, method="sinkhorn", reg_type="kl", c=None, warmstart=None,
                        numItermax=1000, stopThr=1e-6, verbose=False, log=False, **kwargs):
    """Drop-in for ot.unbalanced.sinkhorn_unbalanced (ot/unbalanced/_sinkhorn.py:19-249), method "sinkhorn"."""
    name = method.lower()
    if name == "sinkhorn":
        return sinkhorn_knopp_unbalanced(a, b, M, reg, reg_m, reg_type, c, warmstart, numItermax=numItermax,
                                         stopThr=stopThr, verbose=verbose, log=log, **kwargs)
    if name in ("sinkhorn_stabilized", "sinkhorn_translation_invariant", "sinkhorn_reg_scaling"):
        raise NotImplementedError(f"pot_b200: method '{method}' is outside the accelerated hot path")
    raise ValueError("Unknown method '%s'." % method)
