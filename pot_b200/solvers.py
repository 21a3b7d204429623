"""`ot.solve` / `ot.solve_sample` for the entropic (Sinkhorn) branches, on the libb2ot.so kernels.

Mirrors (same names, argument meaning, defaults, return object and errors):

    ot.solve          ot/solvers.py:57-604    balanced entropic branch :429-494 (reg > 0, reg_type "KL" | "entropy"),
                                              unbalanced KL branch :524-556 (reg > 0, unbalanced = reg_m)
    ot.solve_sample   ot/solvers.py:1354-2060 non-lazy: dist + solve (:1727-1755); lazy regularised: :1977-2002

Everything else `solve` can do in the reference (exact LP, L2 / generic regularisers, L-BFGS-B, low-rank, GeomLoss,
Gaussian ...) is outside the accelerated path and raises NotImplementedError.

After the log-domain solve the plan, value_linear = <M, plan> and the entropy / KL term are produced by ONE kernel
pass that reads M once and writes the plan once (b2ot_dense_plan_value) instead of the reference's n x m temporaries
(:476-483).

Gradients (`grad`), reference :433-494:
    "detach"    : everything detached.
    "envelope"  : d value / d(M, a, b) = (plan, reg (f - mean f), reg (g - mean g)) attached to `value` (:485-494).
    "last_step" : max_iter - 1 iterations on the device, then the reference's one differentiable step in torch (:463-474).
    "autodiff"  : the kernels are forward-only, so when an input requires grad `value` carries the envelope gradient
                  (identical to unrolled autodiff at convergence, test_solve_envelope in the reference) and a
                  UserWarning says so; plan / potentials are detached.
"""

import warnings

import torch

from . import _autograd, _dense, _lib
from .bregman import _NOCONV, _print_table, _require_cuda_matrix, _rowmajor, _solve_dense, _weights
from .utils import OTResult

__all__ = ["solve", "solve_sample"]


_set_gradients = _autograd.set_gradients
_any_requires_grad = _autograd.any_requires_grad


def solve(M, a=None, b=None, reg=None, c=None, reg_type="KL", unbalanced=None, unbalanced_type="KL", method=None,
          n_threads=1, max_iter=None, plan_init=None, potentials_init=None, tol=None, verbose=False, grad="autodiff"):
    """Drop-in for ot.solve (ot/solvers.py:57-604) on the entropic branches; returns an OTResult.

    reg > 0, unbalanced None, reg_type "KL" | "entropy": log-domain Sinkhorn (:429-494).
    reg > 0, unbalanced = reg_m, reg_type "KL", unbalanced_type "KL": sinkhorn_knopp_unbalanced (:524-556).
    """
    M = _require_cuda_matrix(M)
    n, m = M.shape
    M0, a0, b0 = M, a, b
    Md = _rowmajor(M.detach())
    # uniform weights if not given (:301-304)
    a_ = _weights(a.detach() if isinstance(a, torch.Tensor) else ([] if a is None else a), n, Md, "a")
    b_ = _weights(b.detach() if isinstance(b, torch.Tensor) else ([] if b is None else b), m, Md, "b")
    if reg is None:
        reg = 0
    if reg == 0:
        raise NotImplementedError(
            "pot_b200: exact (unregularised) OT is outside the accelerated Sinkhorn path; pass reg > 0")

    if unbalanced is None:
        if isinstance(reg_type, tuple):
            raise NotImplementedError("pot_b200: generic regularisers (f, df) are outside the accelerated path")
        rt = reg_type.lower()
        if rt == "l2":
            raise NotImplementedError("pot_b200: reg_type='l2' is outside the accelerated Sinkhorn path")
        if rt not in ("entropy", "kl"):
            raise NotImplementedError('Not implemented reg_type="{}"'.format(reg_type))  # :516-520
        if max_iter is None:
            max_iter = 1000
        if tol is None:
            tol = 1e-9
        if grad == "last_step":
            if max_iter == 0:
                raise ValueError(
                    "The maximum number of iterations must be greater than 0 when using grad=last_step.")
            max_iter = max_iter - 1
        ws = None
        if max_iter >= 1:
            Mr, f, g, errs, niter, status = _solve_dense("log", a_, b_, Md, reg, int(max_iter), float(tol), ws)
            del Mr  # the epilogue rebuilds -M/reg on the fly: one n x m buffer less at the peak
            if verbose:
                _print_table(errs, after_stop_test=False)
            if status == _lib.ST_MAXITER:
                warnings.warn(_NOCONV)
        else:  # grad="last_step" with max_iter=1: no device iteration, potentials start at zero (:865-866)
            f = torch.zeros(n, dtype=Md.dtype, device=Md.device)
            g = torch.zeros(m, dtype=Md.dtype, device=Md.device)
            niter = -1
        log = {"niter": niter, "log_u": f, "log_v": g}

        if grad == "last_step":
            # one step with the non-detached inputs (:463-474); both updates start from the device potentials
            at = a0 if isinstance(a0, torch.Tensor) else a_
            bt = b0 if isinstance(b0, torch.Tensor) else b_
            loga, logb = torch.log(at.to(Md.dtype)), torch.log(bt.to(Md.dtype))
            v = logb - torch.logsumexp(-M0 / reg + f[:, None], 0)
            u = loga - torch.logsumexp(-M0 / reg + g[None, :], 1)
            plan = torch.exp(-M0 / reg + u[:, None] + v[None, :])
            potentials = (u, v)
            log.update(niter=max_iter + 1, log_u=u, log_v=v, u=torch.exp(u), v=torch.exp(v))
            # value terms on the device from the detached plan inputs; the differentiable value is rebuilt in
            # torch only when someone can backpropagate through it
            if _any_requires_grad(M0, a0, b0):
                value_linear = torch.sum(Md * plan)
                if rt == "entropy":
                    value = value_linear + reg * torch.sum(plan * torch.log(plan + 1e-16))
                else:
                    value = value_linear + reg * torch.sum(plan * torch.log(plan / (a_[:, None] * b_[None, :]) + 1e-16))
                return OTResult(potentials=potentials, value=value, value_linear=value_linear, plan=plan, log=log)
            potentials = (u.detach().contiguous(), v.detach().contiguous())
            plan, vals, _ = _dense.plan_value("log", potentials[0], potentials[1], M=Md, reg=reg, a=a_, b=b_,
                                              terms=("linear", rt))
        else:
            plan, vals, _ = _dense.plan_value("log", f, g, M=Md, reg=reg, a=a_, b=b_, terms=("linear", rt))
            potentials = (f, g)
        value_linear = vals[0].to(Md.dtype)
        value = (vals[0] + reg * vals[1 if rt == "entropy" else 2]).to(Md.dtype)

        wants_grad = _any_requires_grad(M0, a0, b0)
        if grad == "autodiff" and wants_grad:
            warnings.warn("pot_b200: the CUDA kernels are forward-only; grad='autodiff' attaches the envelope "
                          "gradient to `value` (exact at convergence); plan and potentials are detached")
        if grad in ("envelope", "autodiff") and wants_grad:
            inputs = (M0, a0 if isinstance(a0, torch.Tensor) else a_, b0 if isinstance(b0, torch.Tensor) else b_)
            value = _set_gradients(value, inputs,
                                   (plan, reg * (potentials[0] - potentials[0].mean()),
                                    reg * (potentials[1] - potentials[1].mean())))
        return OTResult(potentials=potentials, value=value, value_linear=value_linear, plan=plan, log=log)

    # unbalanced AND regularised (:522-590)
    if (not isinstance(reg_type, tuple) and reg_type.lower() in ["kl"] and unbalanced_type.lower() == "kl"):
        from .unbalanced import sinkhorn_knopp_unbalanced

        if max_iter is None:
            max_iter = 1000
        if tol is None:
            tol = 1e-9
        plan, log = sinkhorn_knopp_unbalanced(a_, b_, Md, reg=reg, reg_m=unbalanced, method=method,
                                              reg_type=reg_type, c=c, warmstart=potentials_init,
                                              numItermax=max_iter, stopThr=tol, verbose=verbose, log=True)
        return OTResult(potentials=(log["logu"], log["logv"]), value=log["total_cost"], value_linear=log["cost"],
                        plan=plan, log=log)
    if (isinstance(reg_type, tuple) or reg_type.lower() in ["kl", "l2", "entropy"]) and \
            unbalanced_type.lower() in ["kl", "l2", "tv"]:
        raise NotImplementedError("pot_b200: the L-BFGS-B unbalanced solver is outside the accelerated path")
    raise NotImplementedError(
        'Not implemented reg_type="{}" and unbalanced_type="{}"'.format(reg_type, unbalanced_type))  # :583-589


def solve_sample(X_a, X_b, a=None, b=None, metric="sqeuclidean", reg=None, c=None, reg_type="KL", unbalanced=None,
                 unbalanced_type="KL", lazy=False, batch_size=None, method=None, n_threads=1, max_iter=None,
                 plan_init=None, rank=100, scaling=0.95, potentials_init=None, X_init=None, tol=None, verbose=False,
                 grad="autodiff", random_state=None):
    """Drop-in for ot.solve_sample (ot/solvers.py:1354-2060) on the Sinkhorn branches.

    lazy=False: M = dist(X_a, X_b, metric) then `solve` (:1727-1755).
    lazy=True, reg > 0, balanced: online-cost solver, the cost matrix is never stored (:1977-2002); the result carries
    `value_linear`, the potentials (log-scalings, as the reference returns them) and `lazy_plan`.
    """
    from .empirical import dist, empirical_sinkhorn2

    if method is not None:
        raise NotImplementedError(
            f"pot_b200: method={method!r} (1d / gaussian / factored / lowrank / geomloss / nystroem) is outside the "
            "accelerated Sinkhorn path")
    if not lazy:
        # `dist` is differentiable wrt the clouds (pot_b200/_autograd.py::DistFunction), so the gradient `solve`
        # attaches to `value` wrt M reaches X_a / X_b, as in the reference (:1727-1755)
        M = dist(X_a, X_b, metric)
        return solve(M, a, b, reg, c, reg_type, unbalanced, unbalanced_type, method, n_threads, max_iter, plan_init,
                     potentials_init, tol, verbose, grad)
    if reg is None or reg == 0:
        if unbalanced is None:
            raise NotImplementedError("Exact OT solver with lazy=True not implemented")  # :1963-1967
        raise NotImplementedError(
            'Non regularized solver with unbalanced_type="{}" not implemented'.format(unbalanced_type))  # :1969-1975
    if unbalanced is not None:
        raise NotImplementedError(
            'Not implemented unbalanced_type="{}" with regularization'.format(unbalanced_type))  # :2004-2010
    if max_iter is None:
        max_iter = 1000
    if tol is None:
        tol = 1e-9
    if batch_size is None:
        batch_size = 100
    value_linear, log = empirical_sinkhorn2(X_a, X_b, reg, a, b, metric=metric, numIterMax=max_iter, stopThr=tol,
                                            isLazy=True, batchSize=batch_size, verbose=verbose, log=True)
    return OTResult(potentials=(log["u"], log["v"]), value=None, value_linear=value_linear, plan=None,
                    lazy_plan=log["lazy_plan"], log=log, batch_size=batch_size)
