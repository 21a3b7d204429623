"""Batched entropic OT with POT's ot.batch signatures, every problem solved on chip by libb2ot.so.

Mirrors (same names, argument meaning, return values and errors):

    ot.batch.bregman_projection_batch       ot/batch/_utils.py:46-162
    ot.batch.bregman_log_projection_batch   ot/batch/_utils.py:165-290
    ot.batch.solve_batch                    ot/batch/_linear.py:236-357
    ot.batch.solve_sample_batch             ot/batch/_linear.py:360-434
    ot.batch.dist_batch                     ot/batch/_linear.py:178-233  (one launch for the whole batch, no clamp)
    ot.batch.loss_linear_batch              ot/batch/_linear.py:123-147
    ot.batch.loss_linear_samples_batch      ot/batch/_linear.py:150-175
    ot.batch.entropy_batch                  ot/batch/_utils.py:4-22
    (loss_linear_batch / entropy_batch are also fused into the value epilogue of solve_batch)

One thread-block cluster keeps one n x m problem in shared memory for the whole solve; the cost matrix is
read from HBM once (the Gibbs transform is applied in shared memory) and the plan is written once.  The
reference's stop test (max over the whole batch of the L2 row-marginal violation, every 10th iteration) is
evaluated on the device.  Inputs must be torch CUDA tensors; there is no CPU path.
"""

import contextlib
import ctypes as C
import warnings

import torch

from . import _autograd, _dense, _lib
from .utils import OTResult

__all__ = ["bregman_projection_batch", "bregman_log_projection_batch", "solve_batch", "solve_sample_batch",
           "dist_batch", "loss_linear_batch", "loss_linear_samples_batch", "entropy_batch", "batch_config"]

_FLAGS = 0  # test hook (B2OT_BATCH_FORCE_EXACT); the public signatures are the reference's, without extra keywords


@contextlib.contextmanager
def _solver_flags(flags):
    """`with batch._solver_flags(_lib.BATCH_FORCE_EXACT):` -- flags of b2ot_batch_solve for the calls inside."""
    global _FLAGS
    old, _FLAGS = _FLAGS, int(flags)
    try:
        yield
    finally:
        _FLAGS = old


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


def bregman_projection_batch(K, a=None, b=None, nx=None, max_iter=10000, tol=1e-5, grad="detach"):
    """Scaling-domain Bregman projection of a batch of affinity matrices (ot/batch/_utils.py:46-162).

    Returns {"T", "n_iters", "potentials": (f, g)} with f, g the scalings.  `nx` is accepted and ignored
    (there is one backend).  grad="last_step" re-applies one differentiable torch step after the solve
    (:151-155); other modes return detached tensors.
    """
    K = _check3(K, "K")
    B, n, m = K.shape
    a, b = _weights(a, B, n, K), _weights(b, B, m, K)
    out = _run("knopp", K, 0, 1.0, None, a, b, max_iter, tol, True, False, 0, _FLAGS)
    f, g, T = out["pa"], out["pb"], out["T"]
    if grad == "last_step" and K.requires_grad:
        f = a / torch.sum(K * g[:, None, :], dim=2)
        g = b / torch.sum(K * f[:, :, None], dim=1)
        T = K.detach() * f[:, :, None] * g[:, None, :]  # K detached, K_ in the step (:151-157)
    return {"T": T, "n_iters": out["n_iters"], "potentials": (f, g)}


def bregman_log_projection_batch(K, a=None, b=None, nx=None, max_iter=10000, tol=1e-5, grad="detach"):
    """Log-domain Bregman projection on K = -M/reg (ot/batch/_utils.py:165-290).

    Returns {"T", "log_T", "n_iters", "potentials": (u, v)} with u, v the log-scalings.
    """
    K = _check3(K, "K")
    B, n, m = K.shape
    a, b = _weights(a, B, n, K), _weights(b, B, m, K)
    out = _run("log", K, 0, 1.0, None, a, b, max_iter, tol, True, True, 0, _FLAGS)
    u, v, T, logT = out["pa"], out["pb"], out["T"], out["log_T"]
    if grad == "last_step" and K.requires_grad:
        loga, logb = torch.log(a), torch.log(b)
        u = loga - torch.logsumexp(K + v[:, None, :], dim=2)
        v = logb - torch.logsumexp(K + u[:, :, None], dim=1)
        logT = K.detach() + u[:, :, None] + v[:, None, :]  # :279-285
        T = torch.exp(logT)
    return {"T": T, "log_T": logT, "n_iters": out["n_iters"], "potentials": (u, v)}


def solve_batch(M, reg, a=None, b=None, max_iter=1000, tol=1e-5, solver="log_sinkhorn", reg_type="entropy",
                grad="envelope"):
    """Batched ot.solve (ot/batch/_linear.py:236-357): B entropic problems in one call, returns an OTResult with
    (B,) `value` / `value_linear`, `potentials`, (B, n, m) `plan` and log {"n_iter"}.

    The Gibbs transform, the iterations, the plan and both objective terms are produced by one kernel per
    problem.  Gradient modes when M requires grad (:333-337):
      "detach" / None : everything detached;
      "envelope"      : plan detached, value_linear = <M, T> differentiable wrt M (formed in torch);
      "last_step"     : max_iter iterations on the device, then the reference's one differentiable step
                        (ot/batch/_utils.py:151-155, :279-283): plan and values carry its graph;
      "autodiff"      : the kernels are forward-only; the last-step graph is used instead and a UserWarning says so.
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
    wants_grad = _autograd.any_requires_grad(M)
    if wants_grad and grad in ("autodiff", "last_step"):
        if grad == "autodiff":
            warnings.warn("pot_b200.solve_batch: the CUDA kernels are forward-only; grad='autodiff' differentiates "
                          "through the last Sinkhorn step only (the reference's grad='last_step')")
        out = _run(mode, M, transform, reg, None, a, b, max_iter, tol, False, False, 0, _FLAGS)
        pa, pb = out["pa"], out["pb"]
        if mode == "log":  # ot/batch/_utils.py:279-286
            K = -M / reg
            u = torch.log(a) - torch.logsumexp(K + pb[:, None, :], dim=2)
            v = torch.log(b) - torch.logsumexp(K + u[:, :, None], dim=1)
            T = torch.exp(K.detach() + u[:, :, None] + v[:, None, :])
        else:  # :151-157
            K = torch.exp(-M / reg)
            u = a / torch.sum(K * pb[:, None, :], dim=2)
            v = b / torch.sum(K * u[:, :, None], dim=1)
            T = K.detach() * u[:, :, None] * v[:, None, :]
        value_linear = torch.sum(M * T, dim=(1, 2))
        value = None
        if rt == 1:
            value = value_linear + reg * torch.sum(T * torch.log(T + 1e-16), dim=(1, 2))
        elif rt == 2:
            value = value_linear + reg * torch.sum(T * torch.log(T / (a[:, :, None] * b[:, None, :]) + 1e-16),
                                                   dim=(1, 2))
        return OTResult(value=value, value_linear=value_linear, potentials=(u, v), plan=T,
                        log={"n_iter": out["n_iters"]})
    out = _run(mode, M, transform, reg, M, a, b, max_iter, tol, True, False, rt, _FLAGS)
    T = out["T"]
    if grad == "envelope" and wants_grad:
        value_linear = loss_linear_batch(M, T)
    else:
        value_linear = out["value_linear"].to(M.dtype)
    value = None
    if rt:
        value = value_linear + reg * out["value_reg"].to(M.dtype)
    return OTResult(value=value, value_linear=value_linear, potentials=(out["pa"], out["pb"]), plan=T,
                    log={"n_iter": out["n_iters"]})


def solve_sample_batch(X_a, X_b, reg, a=None, b=None, metric="sqeuclidean", p=2, max_iter=1000, tol=1e-5,
                       solver="log_sinkhorn", reg_type="entropy", grad="envelope"):
    """Drop-in for ot.batch.solve_sample_batch (ot/batch/_linear.py:360-434): dist_batch, then solve_batch."""
    M = dist_batch(X_a, X_b, metric=metric, p=p)
    return solve_batch(M, reg, a=a, b=b, max_iter=max_iter, tol=tol, solver=solver, reg_type=reg_type, grad=grad)


# ------------------------------------------------------------------------------------------------
# batched cost matrices and per-problem reductions (csrc/batchops_kernels.cuh)
# ------------------------------------------------------------------------------------------------
def _dist_batch_kernel(X1, X2, metric, p):
    lib = _lib.load()
    X1c, X2c = X1.detach().contiguous(), X2.detach().contiguous()
    B, n, d = X1c.shape
    m = X2c.shape[1]
    out = torch.empty((B, n, m), dtype=X1c.dtype, device=X1c.device)
    with torch.cuda.device(X1c.device):
        _lib.check(lib.b2ot_batch_dist(_dense.dtype_code(X1c.dtype), X1c.data_ptr(), X2c.data_ptr(), B, n, m, d,
                                       _lib.BMETRIC[metric], float(p), None, 1e-10, out.data_ptr(),
                                       _dense.stream_ptr(X1c.device)), "b2ot_batch_dist")
    return out


class _DistBatchFunction(torch.autograd.Function):
    """dist_batch differentiable wrt the samples (sqeuclidean / euclidean), as the reference's torch backend is:
    with G = dL/dM per problem, dL/dX = 2 (rowsum(G) X - G Y), dL/dY = 2 (colsum(G) Y - G^T X); euclidean uses
    G / (2 M) (zero where M = 0)."""

    @staticmethod
    def forward(ctx, X1, X2, metric):
        M = _dist_batch_kernel(X1, X2, metric, 2)
        ctx.metric = metric
        ctx.save_for_backward(X1.detach(), X2.detach(), M if metric == "euclidean" else None)
        return M

    @staticmethod
    def backward(ctx, G):
        X, Y, M = ctx.saved_tensors
        if ctx.metric == "euclidean":
            G = torch.where(M > 0, G / (2 * M), torch.zeros_like(G))
        gX = 2 * (G.sum(2, keepdim=True) * X - torch.bmm(G, Y)) if ctx.needs_input_grad[0] else None
        gY = 2 * (G.sum(1)[:, :, None] * Y - torch.bmm(G.transpose(1, 2), X)) if ctx.needs_input_grad[1] else None
        return gX, gY, None


def dist_batch(X1, X2=None, metric="sqeuclidean", p=2, nx=None):
    """Drop-in for ot.batch.dist_batch (ot/batch/_linear.py:178-233): (B, n, d), (B, m, d) -> (B, n, m) in one launch.

    metric: "sqeuclidean", "euclidean" (no clamp at 0, unlike ot.dist -- dist_euclidean_batch :79-85), "minkowski"
    (order p), "kl"; anything else raises ValueError("Unknown metric: ...") like the reference.  `nx` is accepted and
    ignored.  Differentiable wrt the samples for the two euclidean metrics.
    """
    X1 = _check3(X1, "X1")
    X2 = X1 if X2 is None else _check3(X2, "X2")
    if X2.dtype != X1.dtype or X2.device != X1.device or X2.shape[0] != X1.shape[0] or X2.shape[2] != X1.shape[2]:
        raise ValueError("pot_b200: X1 and X2 must share dtype, device, batch size and dimension")
    metric = metric.lower()
    if metric not in _lib.BMETRIC:
        raise ValueError(f"Unknown metric: {metric}")
    if _autograd.any_requires_grad(X1, X2):
        if metric not in ("sqeuclidean", "euclidean"):
            raise NotImplementedError(f"pot_b200: dist_batch(metric={metric!r}) is forward-only; detach the samples")
        return _DistBatchFunction.apply(X1, X2, metric)
    return _dist_batch_kernel(X1, X2, metric, p)


def _reduce(op, A, Bm, eps, sign):
    lib = _lib.load()
    B = A.shape[0]
    per = A.shape[1] * A.shape[2]
    out = torch.empty(B, dtype=A.dtype, device=A.device)
    scratch = torch.empty(lib.b2ot_batch_reduce_scratch_bytes(B, per), dtype=torch.uint8, device=A.device)
    with torch.cuda.device(A.device):
        _lib.check(lib.b2ot_batch_reduce(_dense.dtype_code(A.dtype), op, A.data_ptr(),
                                         Bm.data_ptr() if Bm is not None else None, B, per, float(eps), float(sign),
                                         out.data_ptr(), scratch.data_ptr(), _dense.stream_ptr(A.device)),
                   "b2ot_batch_reduce")
    return out


class _LossLinearFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, M, T):
        Mc, Tc = M.detach().contiguous(), T.detach().contiguous()
        ctx.save_for_backward(Mc, Tc)
        return _reduce(_lib.BREDUCE_DOT, Mc, Tc, 0.0, 1.0)

    @staticmethod
    def backward(ctx, g):
        M, T = ctx.saved_tensors
        g = g[:, None, None]
        return (g * T if ctx.needs_input_grad[0] else None), (g * M if ctx.needs_input_grad[1] else None)


class _EntropyFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, T, eps):
        Tc = T.detach().contiguous()
        ctx.eps = eps
        ctx.save_for_backward(Tc)
        return _reduce(_lib.BREDUCE_ENTROPY, Tc, None, eps, -1.0)

    @staticmethod
    def backward(ctx, g):
        (T,) = ctx.saved_tensors
        return -g[:, None, None] * (torch.log(T + ctx.eps) + T / (T + ctx.eps)), None


def _same3(A, Bm, na, nb):
    A, Bm = _check3(A, na), _check3(Bm, nb)
    if A.shape != Bm.shape or A.dtype != Bm.dtype or A.device != Bm.device:
        raise ValueError(f"pot_b200: `{na}` and `{nb}` must share shape, dtype and device")
    return A, Bm


def loss_linear_batch(M, T, nx=None):
    """Drop-in for ot.batch.loss_linear_batch (ot/batch/_linear.py:123-147): sum_ij M_bij T_bij per problem, one read
    of both operands, fp64 accumulation in a fixed order."""
    M, T = _same3(M, T, "M", "T")
    if _autograd.any_requires_grad(M, T):
        return _LossLinearFunction.apply(M, T)
    return _reduce(_lib.BREDUCE_DOT, M.contiguous(), T.contiguous(), 0.0, 1.0)


def loss_linear_samples_batch(X, Y, T, metric="l2"):
    """Drop-in for ot.batch.loss_linear_samples_batch (ot/batch/_linear.py:150-175): dist_batch then loss_linear_batch.
    (The reference's default metric "l2" is not a dist_batch metric and raises there too.)"""
    return loss_linear_batch(dist_batch(X, Y, metric=metric), T)


def entropy_batch(T, nx=None, eps=1e-16):
    """Drop-in for ot.batch.entropy_batch (ot/batch/_utils.py:4-22): -sum_ij T log(T + eps) per problem."""
    T = _check3(T, "T")
    if _autograd.any_requires_grad(T):
        return _EntropyFunction.apply(T, float(eps))
    return _reduce(_lib.BREDUCE_ENTROPY, T.contiguous(), None, eps, -1.0)
