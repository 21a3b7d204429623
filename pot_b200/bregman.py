"""Dense entropic Sinkhorn with POT's signatures, running on hand-written sm_100a kernels.

Mirrors (same names, argument meaning, return values, warnings and errors):

    ot.sinkhorn                 ot/bregman/_sinkhorn.py:21-234   (methods "sinkhorn", "sinkhorn_log"; "sinkhorn_stabilized"
                                                                   and "sinkhorn_epsilon_scaling" via stabilized.py)
    ot.sinkhorn2                ot/bregman/_sinkhorn.py:237-484
    ot.bregman.sinkhorn_knopp   ot/bregman/_sinkhorn.py:487-687
    ot.bregman.sinkhorn_log     ot/bregman/_sinkhorn.py:690-918

Inputs must be torch CUDA tensors (python lists are accepted for the weights, like the reference's
``list_to_array``); anything else raises — there is no CPU path and no other backend.
The iteration runs entirely on the device: the marginal error every 10th iteration, the numerical
guard and the stop decision are evaluated by the finalize kernel, the host only polls a status word.
"""

import warnings

import torch

from . import _autograd, _dense

__all__ = ["sinkhorn", "sinkhorn2", "sinkhorn_knopp", "sinkhorn_log"]

_NOCONV = (
    "Sinkhorn did not converge. You might want to "
    "increase the number of iterations `numItermax` "
    "or the regularization parameter `reg`."
)
POLL_EVERY = 50  # iterations enqueued between two reads of the device status word


def _require_cuda_matrix(M, name="M"):
    if not isinstance(M, torch.Tensor):
        raise TypeError(
            f"pot_b200: `{name}` must be a torch CUDA tensor (got {type(M).__name__}); "
            "this package has no numpy/CPU backend"
        )
    if not M.is_cuda:
        raise ValueError(f"pot_b200: `{name}` must live on a CUDA device (got {M.device}); there is no CPU fallback")
    if M.dtype not in (torch.float32, torch.float64):
        raise TypeError(f"pot_b200: `{name}` must be float32 or float64 (got {M.dtype})")
    return M


def _weights(w, size, like, name):
    """Reference semantics: list -> array typed like M; empty -> uniform (_sinkhorn.py:588-595)."""
    if isinstance(w, (list, tuple)):
        w = torch.as_tensor(w, dtype=like.dtype, device=like.device)
    if not isinstance(w, torch.Tensor):
        raise TypeError(f"pot_b200: `{name}` must be a torch tensor or a list (got {type(w).__name__})")
    if w.numel() == 0 and w.dim() <= 1:
        return torch.full((size,), 1.0 / size, dtype=like.dtype, device=like.device)
    if w.device != like.device:
        raise ValueError(f"pot_b200: `{name}` is on {w.device} but M is on {like.device}")
    if w.shape[0] != size:
        raise ValueError(f"pot_b200: `{name}` has {w.shape[0]} entries, expected {size}")
    return w.to(like.dtype).contiguous()


def _rowmajor(M):
    return M if M.stride(1) == 1 and M.stride(0) >= M.shape[1] else M.contiguous()


def _print_table(errs, check_every=10, after_stop_test=True, stop_thr=None):
    # reference: header every 200 iterations, one row per check (_sinkhorn.py:660-663, :894-897)
    for k, e in enumerate(errs):
        ii = k * check_every
        if after_stop_test and stop_thr is not None and e < stop_thr:
            break
        if ii % 200 == 0:
            print("{:5s}|{:12s}".format("It.", "Err") + "\n" + "-" * 19)
        print("{:5d}|{:8e}|".format(ii, e))


def _solve_dense(mode, a, b, M, reg, numItermax, stopThr, warmstart, variant="auto", Kt=None):
    """Shared driver.  Returns (K_or_Mr, u, v, errs, niter, status).  `Kt`: an already transformed cost to reuse."""
    from . import _lib

    n, m = M.shape
    Mc = _rowmajor(M)
    with torch.cuda.device(M.device):
        if mode == "knopp":
            if Kt is None:
                Kt = _dense.cost_transform(Mc, reg, "kexp")
            if warmstart is None:
                u0 = torch.full((n,), 1.0 / n, dtype=M.dtype, device=M.device)
                v0 = torch.full((m,), 1.0 / m, dtype=M.dtype, device=M.device)
            else:
                u0 = torch.exp(warmstart[0].to(M.dtype)).contiguous()
                v0 = torch.exp(warmstart[1].to(M.dtype)).contiguous()
            av, logb = a, None
        else:
            if Kt is None:
                Kt = _dense.cost_transform(Mc, reg, "negscale")
            if warmstart is None:
                u0 = torch.zeros(n, dtype=M.dtype, device=M.device)
                v0 = torch.zeros(m, dtype=M.dtype, device=M.device)
            else:
                u0 = warmstart[0].to(M.dtype).contiguous()
                v0 = warmstart[1].to(M.dtype).contiguous()
            av, logb = torch.log(a), torch.log(b)
        solver = _dense.DenseSolver(mode, Kt, av, b, logb, max_iter=numItermax, check_every=10,
                                    stop_thr=stopThr, variant=variant)
        status, niter, _ = solver.solve(u0, v0, poll_every=POLL_EVERY)
        u, v, errs = solver.export()
        solver.close()
    return Kt, u, v, errs, niter, status


def sinkhorn_knopp(a, b, M, reg, numItermax=1000, stopThr=1e-9, verbose=False, log=False, warn=True,
                   warmstart=None, **kwargs):
    """Scaling-domain Sinkhorn-Knopp; drop-in for ot.bregman.sinkhorn_knopp (_sinkhorn.py:487-687).

    Returns the plan (same dtype/device as M) for 1-D ``b``; per-histogram losses for 2-D ``b``;
    with ``log=True`` also the dict {"err", "niter", "u", "v"}.
    """
    from . import _lib

    if not kwargs.get("b2ot_no_grad_warning"):
        _autograd.warn_detached("sinkhorn_knopp", M, a, b)
    M = _require_cuda_matrix(M).detach()
    a = _weights(a, M.shape[0], M, "a").detach()
    b = _weights(b, M.shape[1], M, "b").detach()
    variant = kwargs.get("b2ot_variant", "auto")
    if b.dim() > 1:
        return _multi_hist(sinkhorn_knopp, a, b, M, reg, numItermax, stopThr, verbose, log, warn, warmstart, kwargs)
    K, u, v, errs, niter, status = _solve_dense("knopp", a, b, M, reg, int(numItermax), float(stopThr), warmstart,
                                                variant)
    if status == _lib.ST_NUMERR:
        warnings.warn("Warning: numerical errors at iteration %d" % niter)  # not gated by `warn` (:642)
    elif status == _lib.ST_MAXITER and warn:
        warnings.warn(_NOCONV)
    if verbose:
        _print_table(errs, after_stop_test=True, stop_thr=stopThr)
    G = _dense.plan("knopp", K, u, v)
    if log:
        return G, {"err": [torch.tensor(e, dtype=M.dtype, device=M.device) for e in errs], "niter": niter,
                   "u": u, "v": v}
    return G


def sinkhorn_log(a, b, M, reg, numItermax=1000, stopThr=1e-9, verbose=False, log=False, warn=True,
                 warmstart=None, **kwargs):
    """Log-domain Sinkhorn; drop-in for ot.bregman.sinkhorn_log (_sinkhorn.py:690-918).

    ``warmstart`` is (log u, log v) used as is (:868).  The log dict carries "log_u"/"log_v" and
    their exponentials "u"/"v" (:909-913).
    """
    from . import _lib

    if not kwargs.get("b2ot_no_grad_warning"):
        _autograd.warn_detached("sinkhorn_log", M, a, b)
    M = _require_cuda_matrix(M).detach()
    a = _weights(a, M.shape[0], M, "a").detach()
    b = _weights(b, M.shape[1], M, "b").detach()
    variant = kwargs.get("b2ot_variant", "auto")
    if b.dim() > 1:
        return _multi_hist(sinkhorn_log, a, b, M, reg, numItermax, stopThr, verbose, log, warn, warmstart, kwargs)
    Mr, u, v, errs, niter, status = _solve_dense("log", a, b, M, reg, int(numItermax), float(stopThr), warmstart,
                                                 variant)
    if verbose:
        _print_table(errs, after_stop_test=False)
    if status == _lib.ST_MAXITER and warn:
        warnings.warn(_NOCONV)
    G = _dense.plan("log", Mr, u, v)
    if log:
        return G, {"err": [torch.tensor(e, dtype=M.dtype, device=M.device) for e in errs], "niter": niter,
                   "log_u": u, "log_v": v, "u": torch.exp(u), "v": torch.exp(v)}
    return G


def _multi_hist(fn, a, b, M, reg, numItermax, stopThr, verbose, log, warn, warmstart, kwargs, attach=None):
    """2-D ``b``: losses only (_sinkhorn.py:676-681 and :819-854).

    sinkhorn_knopp keeps all histograms on ONE Gibbs kernel with a joint guard and a joint error / stop test
    (:626-663): `_dense.multi_hist_solve` (two passes over K per iteration whatever the number of histograms,
    decisions on the device).  sinkhorn_log solves one histogram after the other, like the reference (:819-854).
    Each loss <M, plan> is reduced straight from (K, u, v) by the loss kernel, no plan is materialised.
    """
    from . import _lib

    k_h = b.shape[1]
    is_log = fn is sinkhorn_log
    Mc = _rowmajor(M)
    if not is_log:
        n, m = Mc.shape
        dt, dev = Mc.dtype, Mc.device
        with torch.cuda.device(dev):
            Kt = _dense.cost_transform(Mc, reg, "kexp")
            if warmstart is None:
                u0 = torch.full((n, k_h), 1.0 / n, dtype=dt, device=dev)  # :610-613
                v0 = torch.full((m, k_h), 1.0 / m, dtype=dt, device=dev)
            else:
                u0 = torch.exp(warmstart[0].detach().to(dt)).reshape(n, -1).expand(n, k_h)
                v0 = torch.exp(warmstart[1].detach().to(dt)).reshape(m, -1).expand(m, k_h)
            u, v, errs, niter, status = _dense.multi_hist_solve("knopp", Kt, a, b.contiguous(), u0, v0,
                                                                int(numItermax), float(stopThr), poll_every=POLL_EVERY)
            if status == _lib.ST_NUMERR:
                warnings.warn("Warning: numerical errors at iteration %d" % niter)
            elif status == _lib.ST_MAXITER and warn:
                warnings.warn(_NOCONV)
            if verbose:
                _print_table(errs, after_stop_test=True, stop_thr=stopThr)
            losses = []
            for k in range(k_h):
                uk, vk = u[:, k].contiguous(), v[:, k].contiguous()
                loss_k = _dense.loss("knopp", Kt, Mc, uk, vk).to(M.dtype)
                losses.append(loss_k if attach is None else attach(loss_k, Kt, uk, vk, k))
        res = torch.stack(losses)
        if not log:
            return res
        return res, {"err": [torch.tensor(e, dtype=dt, device=dev) for e in errs], "niter": niter, "u": u, "v": v}

    ws = warmstart if warmstart is not None else [None] * k_h
    variant = kwargs.get("b2ot_variant", "auto")
    Kt = None
    losses, us, vs = [], [], []
    for k in range(k_h):
        Kt, u, v, errs, niter, status = _solve_dense("log", a, b[:, k].contiguous(), M, reg, int(numItermax),
                                                     float(stopThr), ws[k], variant, Kt=Kt)
        if status == _lib.ST_MAXITER and warn:
            warnings.warn(_NOCONV)
        if verbose:
            _print_table(errs, after_stop_test=False, stop_thr=None)
        loss_k = _dense.loss("log", Kt, Mc, u, v).to(M.dtype)
        losses.append(loss_k if attach is None else attach(loss_k, Kt, u, v, k))
        us.append(u)
        vs.append(v)
    res = torch.stack(losses)
    if not log:
        return res
    out = {"log_u": torch.stack(us, 1), "log_v": torch.stack(vs, 1)}
    out["u"], out["v"] = torch.exp(out["log_u"]), torch.exp(out["log_v"])
    return res, out


def sinkhorn(a, b, M, reg, method="sinkhorn", numItermax=1000, stopThr=1e-9, verbose=False, log=False, warn=True,
             warmstart=None, **kwargs):
    """Drop-in for ot.sinkhorn (_sinkhorn.py:21-234); methods "sinkhorn" and "sinkhorn_log"."""
    name = method.lower()
    if name == "sinkhorn":
        fn = sinkhorn_knopp
    elif name == "sinkhorn_log":
        fn = sinkhorn_log
    elif name in ("sinkhorn_stabilized", "sinkhorn_epsilon_scaling"):
        from . import stabilized

        fn = stabilized.sinkhorn_stabilized if name == "sinkhorn_stabilized" else stabilized.sinkhorn_epsilon_scaling
    elif name == "greenkhorn":
        raise NotImplementedError(f"pot_b200: method '{method}' is outside the accelerated hot path")
    else:
        raise ValueError("Unknown method '%s'." % method)
    return fn(a, b, M, reg, numItermax=numItermax, stopThr=stopThr, verbose=verbose, log=log, warn=warn,
              warmstart=warmstart, **kwargs)


def sinkhorn2(a, b, M, reg, method="sinkhorn", numItermax=1000, stopThr=1e-9, verbose=False, log=False,
              warn=False, warmstart=None, **kwargs):
    """Drop-in for ot.sinkhorn2 (_sinkhorn.py:237-484): the loss <M, plan> (1-D b), losses per histogram (2-D b).

    The loss is reduced on the device straight from (K, u, v): the plan is never materialised -- unless M, a or b
    require grad: then the plan is formed once and the loss carries the envelope gradient (plan, reg (f - mean f),
    reg (g - mean g)) of ot/solvers.py:485-494 (pot_b200/_autograd.py; a UserWarning says that this replaces the
    reference's unrolled autodiff).
    """
    from . import _lib

    M0, a0, b0 = M, a, b
    wants_grad = _autograd.any_requires_grad(M0, a0, b0)
    M = _require_cuda_matrix(M).detach()
    a = _weights(a, M.shape[0], M, "a").detach()
    b = _weights(b, M.shape[1], M, "b").detach()
    name = method.lower()
    if name == "sinkhorn_stabilized":  # :420-432: loss of the stabilized plan
        from . import stabilized

        _autograd.warn_detached("sinkhorn2(method='sinkhorn_stabilized')", M0, a0, b0)
        return stabilized.sinkhorn_stabilized(a, b, M, reg, numItermax=numItermax, stopThr=stopThr,
                                              warmstart=warmstart, verbose=verbose, log=log, warn=warn,
                                              b2ot_loss=True, **kwargs)
    if name not in ("sinkhorn", "sinkhorn_log"):
        raise ValueError("Unknown method '%s'." % method)
    mode = "knopp" if name == "sinkhorn" else "log"
    if wants_grad:
        warnings.warn(_autograd.ENVELOPE_NOTE)
    ain = a0 if isinstance(a0, torch.Tensor) and a0.numel() else a
    bin_ = b0 if isinstance(b0, torch.Tensor) and b0.numel() else b

    def attach(value, Kt, u, v, b_in):
        if not wants_grad:
            return value
        lu, lv = (torch.log(u), torch.log(v)) if mode == "knopp" else (u, v)
        return _autograd.set_gradients(value, (M0, ain, b_in),
                                       _autograd.envelope_grads(_dense.plan(mode, Kt, u, v), lu, lv, reg))

    if b.dim() > 1:
        if not wants_grad:
            return sinkhorn(a, b, M, reg, method=method, numItermax=numItermax, stopThr=stopThr, verbose=verbose,
                            log=log, warn=warn, warmstart=warmstart, b2ot_no_grad_warning=True, **kwargs)
        return _multi_hist(sinkhorn_knopp if mode == "knopp" else sinkhorn_log, a, b, M, reg, numItermax, stopThr,
                           verbose, log, warn, warmstart, kwargs,
                           attach=lambda value, Kt, u, v, k: attach(value, Kt, u, v, bin_[:, k]))
    Kt, u, v, errs, niter, status = _solve_dense(mode, a, b, M, reg, int(numItermax), float(stopThr), warmstart,
                                                 kwargs.get("b2ot_variant", "auto"))
    if status == _lib.ST_NUMERR:
        warnings.warn("Warning: numerical errors at iteration %d" % niter)
    elif status == _lib.ST_MAXITER and warn:
        warnings.warn(_NOCONV)
    value = attach(_dense.loss(mode, Kt, _rowmajor(M), u, v).to(M.dtype), Kt, u, v, bin_)
    if not log:
        return value
    errl = [torch.tensor(e, dtype=M.dtype, device=M.device) for e in errs]
    if mode == "knopp":
        return value, {"err": errl, "niter": niter, "u": u, "v": v}
    return value, {"err": errl, "niter": niter, "log_u": u, "log_v": v, "u": torch.exp(u), "v": torch.exp(v)}
