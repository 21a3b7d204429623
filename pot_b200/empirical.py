"""Empirical (point-cloud) Sinkhorn with POT's signatures; the cost is rebuilt on chip, never stored.

Mirrors (same names, argument meaning, return values, warnings and errors):

    ot.dist (sqeuclidean / euclidean)          ot/utils.py:400-526, :270-307
    ot.bregman.empirical_sinkhorn              ot/bregman/_empirical.py:74-302
    ot.bregman.empirical_sinkhorn2             ot/bregman/_empirical.py:305-517
    ot.bregman.get_sinkhorn_lazytensor         ot/bregman/_empirical.py:26-71
    ot.bregman.empirical_sinkhorn_divergence   ot/bregman/_empirical.py:520-763

`isLazy=True` runs the online-cost log-domain solver of libb2ot.so: every half-iteration is one sweep that
forms the cost tile from the point clouds in registers (-2 x.y^T + |x|^2 + |y|^2, clamped at 0), adds the
potential, and reduces it with an online log-sum-exp; `batchSize` of the reference (its host-side block loop)
has no effect on the result and is only validated.  `isLazy=False` materialises M with the dist kernel and
calls the dense solver, as the reference does.  Inputs must be torch CUDA tensors; there is no CPU path.
"""

import ctypes as C
import warnings

import torch

from . import _autograd, _dense, _lib
from .bregman import _NOCONV, sinkhorn, sinkhorn2
from .utils import LazyTensor

__all__ = ["dist", "empirical_sinkhorn", "empirical_sinkhorn2", "empirical_sinkhorn_divergence",
           "get_sinkhorn_lazytensor", "EmpiricalSolver"]

POLL_EVERY = 50
_METRICS = {"sqeuclidean": _lib.METRIC_SQEUCLIDEAN, "euclidean": _lib.METRIC_EUCLIDEAN}


def _metric_code(metric):
    try:
        return _METRICS[metric]
    except (KeyError, TypeError):
        raise NotImplementedError(
            f"pot_b200: metric {metric!r} is outside the accelerated hot path (supported: {sorted(_METRICS)})") from None


def _points(X, name):
    if isinstance(X, (list, tuple)):
        raise TypeError(f"pot_b200: `{name}` must be a torch CUDA tensor; this package has no numpy/CPU backend")
    if not isinstance(X, torch.Tensor):
        raise TypeError(f"pot_b200: `{name}` must be a torch CUDA tensor (got {type(X).__name__}); "
                        "this package has no numpy/CPU backend")
    if not X.is_cuda:
        raise ValueError(f"pot_b200: `{name}` must live on a CUDA device (got {X.device}); there is no CPU fallback")
    if X.dtype not in (torch.float32, torch.float64):
        raise TypeError(f"pot_b200: `{name}` must be float32 or float64 (got {X.dtype})")
    assert X.dim() == 2, f"{name} must be a 2d array, got {X.dim()}d array instead"  # ot/utils.py:354-355
    X = X.detach()
    return X if X.stride(1) == 1 and X.stride(0) >= X.shape[1] else X.contiguous()


def dist(x1, x2=None, metric="sqeuclidean", p=2, w=None, backend="auto", nx=None, use_tensor=False):
    """Pairwise cost matrix; drop-in for ot.dist (ot/utils.py:400-526, same signature) with metric "sqeuclidean"
    (default) or "euclidean".

    C = |x|^2 + |y|^2 - 2 x.y^T clamped at 0 (ot/utils.py:292-299); with x2 None (or x2 is x1) the diagonal is zeroed
    like the reference does when X is Y (:304-305).  `p`, `w` and `use_tensor` do not enter these two metrics in the
    reference either (:465-468); `nx` is accepted and ignored (there is one backend here); `backend="scipy"` (the
    host cdist path, :457-464) and every other metric raise NotImplementedError: they are outside the hot path.
    """
    if backend == "scipy":
        raise NotImplementedError("pot_b200: backend='scipy' (host cdist) is outside the accelerated hot path")
    same = x2 is None or x2 is x1
    if _autograd.any_requires_grad(x1, x2):
        _metric_code(metric)
        _points(x1, "x1")  # type / device checks before entering autograd
        if not same:
            _points(x2, "x2")
        return _autograd.DistFunction.apply(x1, x1 if same else x2, metric, same, _dist_kernel)
    return _dist_kernel(x1, None if same else x2, metric)


def _dist_kernel(x1, x2, metric):
    """One launch of the dist kernel on detached inputs; x2 None = the X-is-Y case (diagonal zeroed)."""
    lib = _lib.load()
    X = _points(x1, "x1")
    same = x2 is None
    Y = X if same else _points(x2, "x2")
    if Y.dtype != X.dtype or Y.device != X.device or Y.shape[1] != X.shape[1]:
        raise ValueError("pot_b200: x1 and x2 must share dtype, device and dimension")
    code = _metric_code(metric)
    n, m, d = X.shape[0], Y.shape[0], X.shape[1]
    out = torch.empty((n, m), dtype=X.dtype, device=X.device)
    scratch = torch.empty(lib.b2ot_dist_scratch_bytes(n, m, _dense.dtype_code(X.dtype)), dtype=torch.uint8,
                          device=X.device)
    with torch.cuda.device(X.device):
        _lib.check(lib.b2ot_dist(_dense.dtype_code(X.dtype), X.data_ptr(), Y.data_ptr(), n, m, d, X.stride(0),
                                 Y.stride(0), code, out.data_ptr(), out.stride(0), scratch.data_ptr(),
                                 _dense.stream_ptr(X.device)), "b2ot_dist")
    if same:
        out.fill_diagonal_(0.0)
    return out


def _plan_block(X_a, X_b, f, g, reg, metric):
    lib = _lib.load()
    n, m, d = X_a.shape[0], X_b.shape[0], X_a.shape[1]
    out = torch.empty((n, m), dtype=X_a.dtype, device=X_a.device)
    if n == 0 or m == 0:
        return out
    code = _dense.dtype_code(X_a.dtype)
    scratch = torch.empty(lib.b2ot_dist_scratch_bytes(n, m, code), dtype=torch.uint8, device=X_a.device)
    f, g = f.contiguous(), g.contiguous()
    with torch.cuda.device(X_a.device):
        _lib.check(lib.b2ot_emp_plan_block(code, X_a.data_ptr(), X_b.data_ptr(), f.data_ptr(), g.data_ptr(), n, m, d,
                                           X_a.stride(0), X_b.stride(0), float(reg), _metric_code(metric),
                                           out.data_ptr(), out.stride(0), scratch.data_ptr(),
                                           _dense.stream_ptr(X_a.device)), "b2ot_emp_plan_block")
    return out


def get_sinkhorn_lazytensor(X_a, X_b, f, g, metric="sqeuclidean", reg=1e-1, nx=None):
    """LazyTensor T[i, j] = exp(f_i + g_j - C_ij / reg) (ot/bregman/_empirical.py:26-71); each requested block is
    produced by one kernel from the point clouds (no cost block is stored)."""
    X_a, X_b = _points(X_a, "X_a"), _points(X_b, "X_b")
    shape = (X_a.shape[0], X_b.shape[0])

    def func(i, j, X_a, X_b, f, g, metric, reg):
        if isinstance(i, int):
            i = slice(i, i + 1)
        if isinstance(j, int):
            j = slice(j, j + 1)
        xa, xb, fi, gj = X_a[i], X_b[j], f[i], g[j]
        if xa.stride(-1) != 1 or xb.stride(-1) != 1:  # strided slices
            xa, xb = xa.contiguous(), xb.contiguous()
        return _plan_block(xa, xb, fi, gj, reg, metric)

    return LazyTensor(shape, func, X_a=X_a, X_b=X_b, f=f, g=g, metric=metric, reg=reg)


class EmpiricalSolver:
    """One online-cost solve (or one block of source points of a sharded solve) on the current device."""

    def __init__(self, X_s, X_t, a, b, reg, metric="sqeuclidean", max_iter=10000, stop_thr=1e-9, variant="auto",
                 group=None, exchange="peer", operand="f16"):
        self.lib = _lib.load()
        self.X_s, self.X_t = X_s, X_t
        self.n, self.m, self.d = X_s.shape[0], X_t.shape[0], X_s.shape[1]
        self.dtype, self.device = X_s.dtype, X_s.device
        self.a = a.contiguous()
        self.loga, self.logb = torch.log(a).contiguous(), torch.log(b).contiguous()
        self.max_iter = int(max_iter)
        code = _dense.dtype_code(self.dtype)
        var = {"auto": _lib.EMP_AUTO, "simt": _lib.EMP_SIMT, "tc": _lib.EMP_TC}[variant]
        nbytes = self.lib.b2ot_emp_workspace_bytes(self.n, self.m, self.d, code, _metric_code(metric), var,
                                                   self.max_iter)
        self.ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_emp_create(
                C.byref(h), code, self.n, self.m, self.d, X_s.data_ptr(), X_s.stride(0), X_t.data_ptr(),
                X_t.stride(0), self.a.data_ptr(), self.loga.data_ptr(), self.logb.data_ptr(), float(reg),
                _metric_code(metric), self.ws.data_ptr(), nbytes, self.max_iter, float(stop_thr), var,
                _dense.stream_ptr(self.device)), "b2ot_emp_create")
        self.h = h
        self._cb = None
        self.comm = None
        try:
            op = {"f16": _lib.EMP_OPERAND_F16, "tf32": _lib.EMP_OPERAND_TF32}[operand]
        except KeyError:
            raise ValueError(f"unknown operand split {operand!r} (f16 | tf32)") from None
        _lib.check(self.lib.b2ot_emp_set_operand(self.h, op), "b2ot_emp_set_operand")
        self.operand = operand
        if group is not None:
            if exchange in ("peer", "peer_ll"):
                from .comm import get_comm

                self.comm = get_comm(group, self.device, (self.m + 1) * 8)
                _lib.check(self.lib.b2ot_emp_set_comm(self.h, self.comm.h), "b2ot_emp_set_comm")
            elif exchange == "nccl":
                self._install_allreduce(group)
            else:
                raise ValueError(f"unknown exchange {exchange!r} (peer | nccl)")

    def _install_allreduce(self, group):
        import torch.distributed as dist_

        ws, base = self.ws, self.ws.data_ptr()
        tdt = {_lib.F32: torch.float32, _lib.F64: torch.float64}

        def hook(buf, count, dtype_code, op, _stream, _ctx):
            try:
                dt = tdt[dtype_code]
                off = buf - base
                t = ws[off:off + count * (4 if dt == torch.float32 else 8)].view(dt)
                dist_.all_reduce(t, op=dist_.ReduceOp.MAX if op == _lib.OP_MAX else dist_.ReduceOp.SUM, group=group)
                return 0
            except Exception:
                import traceback
                traceback.print_exc()
                return 1

        self._cb = _lib.ALLREDUCE_FN(hook)
        _lib.check(self.lib.b2ot_emp_set_allreduce(self.h, self._cb, None), "b2ot_emp_set_allreduce")

    def info(self):
        v, a, b = C.c_int(), C.c_int(), C.c_int()
        _lib.check(self.lib.b2ot_emp_info(self.h, C.byref(v), C.byref(a), C.byref(b)), "b2ot_emp_info")
        return {"variant": {_lib.EMP_SIMT: "simt", _lib.EMP_TC: "tc"}[v.value], "nsplit_rows": a.value,
                "nsplit_cols": b.value}

    def init(self, f0, g0):
        self._f0, self._g0 = f0, g0
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_emp_init(self.h, f0.data_ptr(), g0.data_ptr()), "b2ot_emp_init")

    def iterate(self, k):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_emp_iterate(self.h, int(k)), "b2ot_emp_iterate")

    def profile(self, k):
        """k iterations with CUDA events around each sweep launch -> (one-sweep ms avg, iteration ms avg)."""
        a, b = C.c_float(), C.c_float()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_emp_profile(self.h, int(k), C.byref(a), C.byref(b)), "b2ot_emp_profile")
        return a.value, b.value

    def poll(self):
        st, it = C.c_int(), C.c_int()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_emp_poll(self.h, C.byref(st), C.byref(it)), "b2ot_emp_poll")
        return st.value, it.value

    def solve(self, f0, g0, poll_every=POLL_EVERY):
        self._f0, self._g0 = f0, g0
        st, ni, ne = C.c_int(), C.c_int(), C.c_int()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_emp_solve(self.h, f0.data_ptr(), g0.data_ptr(), int(poll_every), C.byref(st),
                                               C.byref(ni), C.byref(ne)), "b2ot_emp_solve")
        if st.value == _lib.ST_COMM_TIMEOUT:
            raise _lib.B2otError("sharded solve: a peer rank never published its exchange flag (timeout)")
        return st.value, ni.value, ne.value

    def export(self):
        f = torch.empty(self.n, dtype=self.dtype, device=self.device)
        g = torch.empty(self.m, dtype=self.dtype, device=self.device)
        cap = self.max_iter // 10 + 2
        errs = (C.c_double * cap)()
        ne = C.c_int()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_emp_export(self.h, f.data_ptr(), g.data_ptr(), errs, cap, C.byref(ne)),
                       "b2ot_emp_export")
        return f, g, [errs[i] for i in range(ne.value)]

    def close(self):
        if getattr(self, "h", None):
            self.lib.b2ot_emp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _unif(n, like):
    return torch.full((n,), 1.0 / n, dtype=like.dtype, device=like.device)


def _lazy_solve(X_s, X_t, reg, a, b, metric, numIterMax, stopThr, batchSize, verbose, warn, warmstart, variant="auto",
                operand="f16"):
    if isinstance(batchSize, int):
        pass
    elif isinstance(batchSize, tuple) and len(batchSize) == 2:
        pass
    else:
        raise ValueError("Batch size must be in integer or a tuple of two integers")
    ns, nt = X_s.shape[0], X_t.shape[0]
    if warmstart is None:
        f0 = torch.zeros(ns, dtype=X_s.dtype, device=X_s.device)
        g0 = torch.zeros(nt, dtype=X_s.dtype, device=X_s.device)
    else:
        f0, g0 = (w.detach().to(X_s.dtype).contiguous() for w in warmstart)
    solver = EmpiricalSolver(X_s, X_t, a, b, reg, metric, max_iter=numIterMax, stop_thr=stopThr, variant=variant,
                             operand=operand)
    status, niter, _ = solver.solve(f0, g0)
    f, g, errs = solver.export()
    solver.close()
    if verbose:
        for k, e in enumerate(errs):
            it1 = (k + 1) * 10
            if it1 % 100 == 0:
                print("Error in marginal at iteration {} = {}".format(it1, e))
    if status == _lib.ST_MAXITER and warn:
        warnings.warn(_NOCONV)
    return f, g, errs, niter


def _prep(X_s, X_t, a, b):
    # identity is decided BEFORE detaching: the reference's `dist` zeroes the diagonal iff `X is Y` (ot/utils.py:304),
    # and list_to_array keeps the caller's objects, so empirical_sinkhorn(X, X, isLazy=False) takes that branch
    same = X_t is X_s
    X_s = _points(X_s, "X_s")
    X_t = X_s if same else _points(X_t, "X_t")
    if X_t.dtype != X_s.dtype or X_t.device != X_s.device or X_t.shape[1] != X_s.shape[1]:
        raise ValueError("pot_b200: X_s and X_t must share dtype, device and dimension")
    a = _unif(X_s.shape[0], X_s) if a is None else a.detach().to(X_s.dtype).contiguous()
    b = _unif(X_t.shape[0], X_s) if b is None else b.detach().to(X_s.dtype).contiguous()
    return X_s, X_t, a, b


def empirical_sinkhorn(X_s, X_t, reg, a=None, b=None, metric="sqeuclidean", numIterMax=10000, stopThr=1e-9,
                       isLazy=False, batchSize=100, verbose=False, log=False, warn=True, warmstart=None, **kwargs):
    """Drop-in for ot.bregman.empirical_sinkhorn (ot/bregman/_empirical.py:74-302).

    isLazy=True : returns the log-scalings (f, g) [+ dict_log {"err", "u": f, "v": g, "niter", "lazy_plan"}];
    isLazy=False: M = dist(X_s, X_t), then sinkhorn(a, b, M, reg, ...) -> plan [+ log].
    """
    _autograd.warn_detached("empirical_sinkhorn", X_s, X_t, a, b)
    X_s, X_t, a, b = _prep(X_s, X_t, a, b)
    if isLazy:
        f, g, errs, niter = _lazy_solve(X_s, X_t, reg, a, b, metric, int(numIterMax), float(stopThr), batchSize,
                                        verbose, warn, warmstart, kwargs.get("b2ot_variant", "auto"),
                                        kwargs.get("b2ot_operand", "f16"))
        if log:
            dict_log = {"err": [torch.tensor(e, dtype=X_s.dtype, device=X_s.device) for e in errs], "u": f, "v": g,
                        "niter": niter, "lazy_plan": get_sinkhorn_lazytensor(X_s, X_t, f, g, metric, reg)}
            return f, g, dict_log
        return f, g
    M = dist(X_s, X_t, metric=metric)
    # the reference forwards everything but `warn` here (:276-301)
    return sinkhorn(a, b, M, reg, numItermax=numIterMax, stopThr=stopThr, verbose=verbose, log=log,
                    warmstart=warmstart, **kwargs)


def empirical_sinkhorn2(X_s, X_t, reg, a=None, b=None, metric="sqeuclidean", numIterMax=10000, stopThr=1e-9,
                        isLazy=False, batchSize=100, verbose=False, log=False, warn=True, warmstart=None, **kwargs):
    """Drop-in for ot.bregman.empirical_sinkhorn2 (ot/bregman/_empirical.py:305-517): the loss <C, plan>.

    Lazy mode reduces sum_ij C_ij exp(f_i + g_j - C_ij/reg) in one sweep from the point clouds (:466-483).

    When X_s, X_t, a or b require grad the loss carries the envelope gradient (pot_b200/_autograd.py): the lazy path
    rebuilds the plan block by block in backward; the dense path goes through the differentiable `dist` and
    `sinkhorn2`.
    """
    X_s0, X_t0, a0, b0 = X_s, X_t, a, b
    wants_grad = _autograd.any_requires_grad(X_s0, X_t0, a0, b0)
    X_s, X_t, a, b = _prep(X_s, X_t, a, b)
    if isLazy:
        lib = _lib.load()
        f, g, errs, niter = _lazy_solve(X_s, X_t, reg, a, b, metric, int(numIterMax), float(stopThr), batchSize,
                                        verbose, warn, warmstart, kwargs.get("b2ot_variant", "auto"),
                                        kwargs.get("b2ot_operand", "f16"))
        ns, nt, d = X_s.shape[0], X_t.shape[0], X_s.shape[1]
        code = _dense.dtype_code(X_s.dtype)
        out = torch.empty((), dtype=torch.float64, device=X_s.device)
        scratch = torch.empty(lib.b2ot_emp_loss_scratch_bytes(ns, nt, code), dtype=torch.uint8, device=X_s.device)
        with torch.cuda.device(X_s.device):
            _lib.check(lib.b2ot_emp_loss(code, X_s.data_ptr(), X_t.data_ptr(), f.data_ptr(), g.data_ptr(), ns, nt, d,
                                         X_s.stride(0), X_t.stride(0), float(reg), _metric_code(metric),
                                         out.data_ptr(), scratch.data_ptr(), _dense.stream_ptr(X_s.device)),
                       "b2ot_emp_loss")
        loss = out.to(X_s.dtype)
        if wants_grad:
            warnings.warn(_autograd.ENVELOPE_NOTE)
            block = batchSize[0] if isinstance(batchSize, tuple) else batchSize
            loss = _autograd.LazyLossFunction.apply(
                loss, X_s0, X_t0, a0 if isinstance(a0, torch.Tensor) else a, b0 if isinstance(b0, torch.Tensor) else b,
                f, g, float(reg), metric, _plan_block, max(int(block), 1024))
        if log:
            dict_log = {"err": [torch.tensor(e, dtype=X_s.dtype, device=X_s.device) for e in errs], "u": f, "v": g,
                        "niter": niter, "lazy_plan": get_sinkhorn_lazytensor(X_s, X_t, f, g, metric, reg)}
            return loss, dict_log
        return loss
    if wants_grad:
        M = dist(X_s0, X_t0, metric=metric)  # differentiable wrt the clouds; keeps the X-is-Y identity
        a, b = (a0 if isinstance(a0, torch.Tensor) else a), (b0 if isinstance(b0, torch.Tensor) else b)
    else:
        M = dist(X_s, X_t, metric=metric)
    return sinkhorn2(a, b, M, reg, numItermax=numIterMax, stopThr=stopThr, verbose=verbose, log=log, warn=warn,
                     warmstart=warmstart, **kwargs)


def empirical_sinkhorn_divergence(X_s, X_t, reg, a=None, b=None, metric="sqeuclidean", numIterMax=10000,
                                  stopThr=1e-9, verbose=False, log=False, warn=True, warmstart=None, **kwargs):
    """Drop-in for ot.bregman.empirical_sinkhorn_divergence (ot/bregman/_empirical.py:520-763).

    S = max(0, W(X_s, X_t) - (W(X_s, X_s) + W(X_t, X_t)) / 2), three `empirical_sinkhorn2` solves; `isLazy=True`
    (through **kwargs, as in the reference) keeps all three on the online-cost kernels.
    """
    if warmstart is None:
        warmstart_a, warmstart_b = None, None
    else:
        u, v = warmstart
        warmstart_a, warmstart_b = (u, u), (v, v)  # :646-648
    common = dict(metric=metric, numIterMax=numIterMax, stopThr=stopThr, verbose=verbose, log=log, warn=warn)
    r_ab = empirical_sinkhorn2(X_s, X_t, reg, a, b, warmstart=warmstart, **common, **kwargs)
    r_a = empirical_sinkhorn2(X_s, X_s, reg, a, a, warmstart=warmstart_a, **common, **kwargs)
    r_b = empirical_sinkhorn2(X_t, X_t, reg, b, b, warmstart=warmstart_b, **common, **kwargs)
    if log:
        (l_ab, log_ab), (l_a, log_a), (l_b, log_b) = r_ab, r_a, r_b
    else:
        l_ab, l_a, l_b = r_ab, r_a, r_b
    div = torch.clamp(l_ab - 0.5 * (l_a + l_b), min=0)
    if log:
        return div, {"sinkhorn_loss_ab": l_ab, "sinkhorn_loss_a": l_a, "sinkhorn_loss_b": l_b,
                     "log_sinkhorn_ab": log_ab, "log_sinkhorn_a": log_a, "log_sinkhorn_b": log_b}
    return div
