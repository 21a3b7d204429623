"""Batched entropic OT with POT's ot.batch signatures, every problem solved on chip by libb2ot.so.

Mirrors (same names, argument meaning, return values and errors):

    ot.batch.bregman_projection_batch       ot/batch/_utils.py:46-162
    ot.batch.bregman_log_projection_batch   ot/batch/_utils.py:165-290
    ot.batch.solve_batch                    ot/batch/_linear.py:236-357
    ot.batch.loss_linear_batch / entropy_batch (as the fused value epilogue of solve_batch)

One thread-block cluster keeps one n x m problem in shared memory for the whole solve; the cost matrix is
read from HBM once (the Gibbs transform is applied in shared memory) and the plan is written once.  The
reference's stop test (max over the whole batch of the L2 row-marginal violation, every 10th iteration) is
evaluated on the device.  Inputs must be torch CUDA tensors; there is no CPU path.
"""

import ctypes as C

import torch

from . import _dense, _lib
from .utils import OTResult

__all__ = ["bregman_projection_batch", "bregman_log_projection_batch", "solve_batch", "batch_config"]


def _check3(K, name):
    if not isinstance(K, torch.Tensor):
        raise TypeError(f"pot_b200: `{name}` must be a torch CUDA tensor (got {type(K).__name__}); "
                        "this package has no numpy/CPU backend")
    if not K.is_cuda:
        raise ValueError(f"pot_b200: `{name}` must live on a CUDA device (got {K.device}); there is no CPU fallback")
    if K.dim() != 3:
        raise ValueError(f"pot_b200: `{name}` must have shape (B, n, m), got {tuple(K.shape)}")
    if K.dtype not in (torch.float32, torch.float64):
        raise TypeError(f"pot_b200: `{name}` must be float32 or float64 (got {K.dtype})")
    return K


def _weights(w, B, size, like):
    if w is None:
        return torch.full((B, size), 1.0 / size, dtype=like.dtype, device=like.device)
    if w.shape != (B, size):
        raise ValueError(f"pot_b200: weights have shape {tuple(w.shape)}, expected {(B, size)}")
    return w.detach().to(device=like.device, dtype=like.dtype).contiguous()


def batch_config(mode, dtype, n, m):
    """(cluster_size, shared-memory bytes per CTA, columns per lane) the on-chip solver uses for n x m problems."""
    lib = _lib.load()
    cs, sm, cpl = C.c_int(), C.c_int(), C.c_int()
    _lib.check(lib.b2ot_batch_config(_lib.MODE_KNOPP if mode == "knopp" else _lib.MODE_LOG, _dense.dtype_code(dtype),
                                     int(n), int(m), C.byref(cs), C.byref(sm), C.byref(cpl)), "b2ot_batch_config")
    return cs.value, sm.value, cpl.value


def _run(mode, X, transform, reg, Mval, a, b, max_iter, tol, want_plan, want_logT, reg_type, flags=0,
         sync_n_iters=True):
    """One call of b2ot_batch_solve.  Returns dict(T, log_T, pa, pb, value_linear, value_reg, n_iters)."""
    lib = _lib.load()
    B, n, m = X.shape
    dt, dev = X.dtype, X.device
    Xc = X.detach().contiguous()
    Mv = Mval.detach().contiguous() if Mval is not None else None
    pa = torch.empty((B, n), dtype=dt, device=dev)
    pb = torch.empty((B, m), dtype=dt, device=dev)
    T = torch.empty((B, n, m), dtype=dt, device=dev) if want_plan else None
    logT = torch.empty((B, n, m), dtype=dt, device=dev) if want_logT else None
    vl = torch.empty(B, dtype=torch.float64, device=dev) if Mv is not None else None
    vr = torch.empty(B, dtype=torch.float64, device=dev) if reg_type else None
    nbytes = lib.b2ot_batch_workspace_bytes(B)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    nit = C.c_int(-1)
    ptr = lambda t: t.data_ptr() if t is not None else None  # noqa: E731
    with torch.cuda.device(dev):
        _lib.check(lib.b2ot_batch_solve(
            _lib.MODE_KNOPP if mode == "knopp" else _lib.MODE_LOG, _dense.dtype_code(dt), B, n, m, Xc.data_ptr(),
            int(transform), float(reg), ptr(Mv), a.data_ptr(), b.data_ptr(), pa.data_ptr(), pb.data_ptr(), ptr(T),
            ptr(logT), ptr(vl), ptr(vr), int(reg_type), int(max_iter), float(tol), ws.data_ptr(), nbytes,
            C.byref(nit) if sync_n_iters else None, int(flags), _dense.stream_ptr(dev)), "b2ot_batch_solve")
    return {"T": T, "log_T": logT, "pa": pa, "pb": pb, "value_linear": vl, "value_reg": vr,
            "n_iters": nit.value if sync_n_iters else int(max_iter) - 1}


def bregman_projection_batch(K, a=None, b=None, nx=None, max_iter=10000, tol=1e-5, grad="detach", **kwargs):
    """Scaling-domain Bregman projection of a batch of affinity matrices (ot/batch/_utils.py:46-162).

    Returns {"T", "n_iters", "potentials": (f, g)} with f, g the scalings.  `nx` is accepted and ignored
    (there is one backend).  grad="last_step" re-applies one differentiable torch step after the solve
    (:151-155); other modes return detached tensors.
    """
    K = _check3(K, "K")
    B, n, m = K.shape
    a, b = _weights(a, B, n, K), _weights(b, B, m, K)
    out = _run("knopp", K, 0, 1.0, None, a, b, max_iter, tol, True, False, 0, kwargs.get("b2ot_flags", 0))
    f, g, T = out["pa"], out["pb"], out["T"]
    if grad == "last_step" and K.requires_grad:
        f = a / torch.sum(K * g[:, None, :], dim=2)
        g = b / torch.sum(K * f[:, :, None], dim=1)
        T = K * f[:, :, None] * g[:, None, :]
    return {"T": T, "n_iters": out["n_iters"], "potentials": (f, g)}


def bregman_log_projection_batch(K, a=None, b=None, nx=None, max_iter=10000, tol=1e-5, grad="detach", **kwargs):
    """Log-domain Bregman projection on K = -M/reg (ot/batch/_utils.py:165-290).

    Returns {"T", "log_T", "n_iters", "potentials": (u, v)} with u, v the log-scalings.
    """
    K = _check3(K, "K")
    B, n, m = K.shape
    a, b = _weights(a, B, n, K), _weights(b, B, m, K)
    out = _run("log", K, 0, 1.0, None, a, b, max_iter, tol, True, True, 0, kwargs.get("b2ot_flags", 0))
    u, v, T, logT = out["pa"], out["pb"], out["T"], out["log_T"]
    if grad == "last_step" and K.requires_grad:
        loga, logb = torch.log(a), torch.log(b)
        u = loga - torch.logsumexp(K + v[:, None, :], dim=2)
        v = logb - torch.logsumexp(K + u[:, :, None], dim=1)
        logT = K + u[:, :, None] + v[:, None, :]
        T = torch.exp(logT)
    return {"T": T, "log_T": logT, "n_iters": out["n_iters"], "potentials": (u, v)}


def solve_batch(M, reg, a=None, b=None, max_iter=1000, tol=1e-5, solver="log_sinkhorn", reg_type="entropy",
                grad="envelope", **kwargs):
    """Batched ot.solve (ot/batch/_linear.py:236-357): B entropic problems in one call, returns an OTResult with
    (B,) `value` / `value_linear`, `potentials`, (B, n, m) `plan` and log {"n_iter"}.

    The Gibbs transform, the iterations, the plan and both objective terms are produced by one kernel per
    problem; `grad="envelope"` keeps the reference's envelope gradient wrt M by forming <M, T> in torch with
    T detached when M requires grad.
    """
    M = _check3(M, "M")
    B, n, m = M.shape
    a, b = _weights(a, B, n, M), _weights(b, B, m, M)
    if solver == "log_sinkhorn":
        mode, transform = "log", 2
    elif solver == "sinkhorn":
        mode, transform = "knopp", 1
    else:
        raise ValueError(f"Unknown solver: {solver}")
    rt = {"entropy": 1, "kl": 2}.get(reg_type.lower(), 0)
    out = _run(mode, M, transform, reg, M, a, b, max_iter, tol, True, False, rt, kwargs.get("b2ot_flags", 0))
    T = out["T"]
    if grad == "envelope" and M.requires_grad:
        value_linear = torch.sum(M * T, dim=(1, 2))
    else:
        value_linear = out["value_linear"].to(M.dtype)
    value = None
    if rt:
        value = value_linear + reg * out["value_reg"].to(M.dtype)
    return OTResult(value=value, value_linear=value_linear, potentials=(out["pa"], out["pb"]), plan=T,
                    log={"n_iter": out["n_iters"]})
