"""CPU oracle for the entropic Sinkhorn hot path (TEST INFRASTRUCTURE ONLY).

This module is a numpy restatement of the reference algorithms of PythonOT/POT
for the path named by BASELINE.json.  It is the *checker*: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it.  Nothing under ``pot_b200/`` imports it
and the product path never falls back to it.

Parity status: PINNED.  ``tests/golden/make_golden.py`` ran the unmodified
reference (imported from /root/reference in the build container) on seeded
inputs and committed its outputs under ``tests/golden/*.npz``;
``tests/test_oracle_golden.py`` checks every function below against those
vectors and against the reference's doctest known-answer values.

Each function cites the reference file:line (relative to /root/reference) that
it restates.  The arithmetic (operation order, dtype, where temporaries are
rounded) follows the reference's numpy backend so that fp32 runs round the same
way: np.dot for the GEMVs, scipy.special.logsumexp for the log domain,
np.einsum for the marginal check.
"""

import warnings

import numpy as np
from scipy.special import logsumexp as _lse

__all__ = [
    "sqeuclidean",
    "sinkhorn_knopp",
    "sinkhorn_log",
    "sinkhorn",
    "sinkhorn2",
    "empirical_sinkhorn_lazy",
    "lazy_plan_block",
    "bregman_projection_batch",
    "bregman_log_projection_batch",
    "solve_batch",
    "dist_sqeuclidean_batch",
]

_NOCONV = (
    "Sinkhorn did not converge. You might want to increase the number of "
    "iterations `numItermax` or the regularization parameter `reg`."
)


def sqeuclidean(X, Y):
    """Squared euclidean cost, clamped at 0 (ot/utils.py:292-299, via dist :465-466)."""
    xx = np.einsum("ij,ij->i", X, X)
    yy = np.einsum("ij,ij->i", Y, Y)
    C = -2 * np.dot(X, Y.T)
    C += xx[:, None]
    C += yy[None, :]
    return np.maximum(C, 0)


def _as_arrays(a, b, M):
    """list -> array typed like the first array argument (ot/utils.py:59-85)."""
    M = np.asarray(M)
    a = np.asarray(a, dtype=M.dtype) if isinstance(a, list) else np.asarray(a)
    b = np.asarray(b, dtype=M.dtype) if isinstance(b, list) else np.asarray(b)
    if a.shape[0] == 0:  # ot/bregman/_sinkhorn.py:592-595
        a = np.full((M.shape[0],), 1.0 / M.shape[0], dtype=M.dtype)
    if b.shape[0] == 0:
        b = np.full((M.shape[1],), 1.0 / M.shape[1], dtype=M.dtype)
    return a, b, M


def knopp_loop(K, Kp, b, u, v, numItermax, stopThr, hist):
    """The hot loop of sinkhorn_knopp (ot/bregman/_sinkhorn.py:626-663), split out so bench.py can time
    the iterations without the one-off K / Kp / plan passes.  Returns (u, v, last index, stopped)."""
    multi = b.ndim > 1
    ii = 0
    for ii in range(numItermax):
        u_old, v_old = u, v
        KtU = np.dot(K.T, u)
        v = b / KtU
        u = 1.0 / np.dot(Kp, v)
        if (np.any(KtU == 0) or np.any(np.isnan(u)) or np.any(np.isnan(v))
                or np.any(np.isinf(u)) or np.any(np.isinf(v))):
            warnings.warn("Warning: numerical errors at iteration %d" % ii)
            return u_old, v_old, ii, True
        if ii % 10 == 0:
            if multi:
                marg = np.einsum("ik,ij,jk->jk", u, K, v)
            else:
                marg = np.einsum("i,ij,j->j", u, K, v)
            err = np.linalg.norm(marg - b)
            hist["err"].append(err)
            if err < stopThr:
                return u, v, ii, True
    return u, v, ii, False


def sinkhorn_knopp(a, b, M, reg, numItermax=1000, stopThr=1e-9, log=False,
                   warn=True, warmstart=None):
    """Scaling-domain Sinkhorn-Knopp (ot/bregman/_sinkhorn.py:588-687).

    v-then-u update order (:629-631), per-iteration numerical guard with revert
    (:633-645), marginal check every 10th iteration on the column marginal with
    a strict ``<`` stop test (:646-659), ``niter`` = last loop index (:672).
    1-D ``b`` returns the plan, 2-D ``b`` returns per-histogram losses (:676-681).
    """
    a, b, M = _as_arrays(a, b, M)
    n, m = a.shape[0], b.shape[0]
    multi = b.ndim > 1
    hist = {"err": []}
    if warmstart is None:
        shape_u = (n, b.shape[1]) if multi else (n,)
        shape_v = (m, b.shape[1]) if multi else (m,)
        u = np.ones(shape_u, dtype=M.dtype) / n
        v = np.ones(shape_v, dtype=M.dtype) / m
    else:
        u, v = np.exp(warmstart[0]), np.exp(warmstart[1])
    K = np.exp(M / (-reg))
    Kp = (1 / a).reshape(-1, 1) * K
    u, v, ii, converged_or_broke = knopp_loop(K, Kp, b, u, v, numItermax, stopThr, hist)
    if not converged_or_broke and warn:
        warnings.warn(_NOCONV)
    hist["niter"] = ii
    hist["u"] = u
    hist["v"] = v
    if multi:
        res = np.einsum("ik,ij,jk,ij->k", u, K, v, M)
    else:
        res = u.reshape((-1, 1)) * K * v.reshape((1, -1))
    return (res, hist) if log else res


def sinkhorn_log(a, b, M, reg, numItermax=1000, stopThr=1e-9, log=False,
                 warn=True, warmstart=None):
    """Log-domain Sinkhorn (ot/bregman/_sinkhorn.py:797-918).

    ``Mr = -M/reg`` (:860); v then u by logsumexp (:881-882); check every 10th
    iteration on the column marginal of exp(Mr+u+v) (:884-899); no guard.
    2-D ``b`` is a loop of single solves returning losses (:819-854).
    """
    a, b, M = _as_arrays(a, b, M)
    if b.ndim > 1:
        k_h = b.shape[1]
        ws = warmstart if warmstart is not None else [None] * k_h
        if k_h == 1 and warmstart is None:
            ws = [None]
        losses, lu, lv = [], [], []
        for k in range(k_h):
            G, lg = sinkhorn_log(a, b[:, k], M, reg, numItermax, stopThr, True,
                                 warn, ws[k])
            losses.append(np.sum(M * G))
            lu.append(lg["log_u"])
            lv.append(lg["log_v"])
        res = np.stack(losses)
        if log:
            out = {"log_u": np.stack(lu, 1), "log_v": np.stack(lv, 1)}
            out["u"] = np.exp(out["log_u"])
            out["v"] = np.exp(out["log_v"])
            return res, out
        return res
    n, m = a.shape[0], b.shape[0]
    hist = {"err": []}
    Mr = -M / reg
    if warmstart is None:
        u = np.zeros(n, dtype=M.dtype)
        v = np.zeros(m, dtype=M.dtype)
    else:
        u, v = warmstart
    loga, logb = np.log(a), np.log(b)
    ii = 0
    stopped = False
    for ii in range(numItermax):
        v = logb - _lse(Mr + u[:, None], 0)
        u = loga - _lse(Mr + v[None, :], 1)
        if ii % 10 == 0:
            marg = np.sum(np.exp(Mr + u[:, None] + v[None, :]), 0)
            err = np.linalg.norm(marg - b)
            hist["err"].append(err)
            if err < stopThr:
                stopped = True
                break
    if not stopped and warn:
        warnings.warn(_NOCONV)
    G = np.exp(Mr + u[:, None] + v[None, :])
    if log:
        hist.update(niter=ii, log_u=u, log_v=v, u=np.exp(u), v=np.exp(v))
        return G, hist
    return G


def sinkhorn(a, b, M, reg, method="sinkhorn", **kw):
    """Dispatcher (ot/bregman/_sinkhorn.py:164-234), in-scope methods only."""
    name = method.lower()
    if name == "sinkhorn":
        return sinkhorn_knopp(a, b, M, reg, **kw)
    if name == "sinkhorn_log":
        return sinkhorn_log(a, b, M, reg, **kw)
    raise ValueError("Unknown method '%s'." % method)


def sinkhorn2(a, b, M, reg, method="sinkhorn", log=False, warn=False, **kw):
    """Loss <M, plan> for 1-D b (ot/bregman/_sinkhorn.py:393-438), passthrough otherwise."""
    a_, b_, M_ = _as_arrays(a, b, M)
    res = sinkhorn(a_, b_, M_, reg, method=method, log=log, warn=warn, **kw)
    if b_.ndim > 1:
        return res
    if log:
        return np.sum(M_ * res[0]), res[1]
    return np.sum(M_ * res)


def empirical_sinkhorn_lazy(X_s, X_t, reg, a=None, b=None, numIterMax=10000,
                            stopThr=1e-9, batchSize=100, log=False, warn=True,
                            warmstart=None):
    """Online-cost log-domain Sinkhorn (ot/bregman/_empirical.py:193-271).

    f then g (:225, :233); cost blocks rebuilt with the clamped squared
    euclidean distance; L1 error on the ROW marginal every 10th iteration,
    counted from 1, with a ``<=`` stop test (:235-256).
    """
    X_s, X_t = np.asarray(X_s), np.asarray(X_t)
    ns, nt = X_s.shape[0], X_t.shape[0]
    if a is None:
        a = (np.ones((ns,)) / ns).astype(X_s.dtype)
    if b is None:
        b = (np.ones((nt,)) / nt).astype(X_s.dtype)
    if isinstance(batchSize, int):
        bs = bt = batchSize
    elif isinstance(batchSize, tuple) and len(batchSize) == 2:
        bs, bt = batchSize
    else:
        raise ValueError("Batch size must be in integer or a tuple of two integers")
    log_a, log_b = np.log(a), np.log(b)
    if warmstart is None:
        f, g = np.zeros((ns,), dtype=a.dtype), np.zeros((nt,), dtype=a.dtype)
    else:
        f, g = warmstart
    hist = {"err": []}
    it = 0
    stopped = False
    for it in range(numIterMax):
        parts = []
        for i in range(0, ns, bs):
            C = sqeuclidean(X_s[i:i + bs, :], X_t).astype(a.dtype, copy=False)
            parts.append(_lse(g[None, :] - C / reg, axis=1))
        f = log_a - np.concatenate(parts, axis=0)
        parts = []
        for j in range(0, nt, bt):
            C = sqeuclidean(X_s, X_t[j:j + bt, :]).astype(a.dtype, copy=False)
            parts.append(_lse(f[:, None] - C / reg, axis=0))
        g = log_b - np.concatenate(parts, axis=0)
        if (it + 1) % 10 == 0:
            parts = []
            for i in range(0, ns, bs):
                C = sqeuclidean(X_s[i:i + bs, :], X_t).astype(a.dtype, copy=False)
                parts.append(np.sum(np.exp(f[i:i + bs, None] + g[None, :] - C / reg), axis=1))
            err = np.sum(np.abs(np.concatenate(parts, axis=0) - a))
            hist["err"].append(err)
            if err <= stopThr:
                stopped = True
                break
    if not stopped and warn:
        warnings.warn(_NOCONV)
    if log:
        hist.update(u=f, v=g, niter=it)
        return f, g, hist
    return f, g


def lazy_plan_block(X_s, X_t, f, g, reg, i=slice(None), j=slice(None)):
    """Block of exp(f_i + g_j - C_ij/reg) (ot/bregman/_empirical.py:61-67)."""
    if isinstance(i, int):
        i = slice(i, i + 1)
    if isinstance(j, int):
        j = slice(j, j + 1)
    C = sqeuclidean(X_s[i], X_t[j])
    return np.exp(f[i, None] + g[None, j] - C / reg)


def _norm_batch(x):
    return np.sum(x ** 2, axis=1) ** (1 / 2)  # ot/batch/_utils.py:25-29


def bregman_projection_batch(K, a=None, b=None, max_iter=10000, tol=1e-5):
    """Batched scaling-domain projection (ot/batch/_utils.py:123-162).

    f then g (:142-143); every 10th iteration the max over the batch of the L2
    row-marginal violation is tested with ``<`` (:144-149).
    NOTE the reference creates f, g without ``type_as`` (:138-139): float64 ones
    on numpy, so fp32 K is promoted; reproduced here.
    """
    B, n, m = K.shape
    if a is None:
        a = np.ones((B, n)) / n
    if b is None:
        b = np.ones((B, m)) / m
    f = np.ones((B, n))
    g = np.ones((B, m))
    it = 0
    for it in range(max_iter):
        f = a / np.sum(K * g[:, None, :], axis=2)
        g = b / np.sum(K * f[:, :, None], axis=1)
        if it % 10 == 0:
            T = K * f[:, :, None] * g[:, None, :]
            err = np.max(_norm_batch(np.sum(T, axis=2) - a))
            if err < tol:
                break
    T = K * f[:, :, None] * g[:, None, :]
    return {"T": T, "n_iters": it, "potentials": (f, g)}


def bregman_log_projection_batch(K, a=None, b=None, max_iter=10000, tol=1e-5):
    """Batched log-domain projection on K = -M/reg (ot/batch/_utils.py:247-290)."""
    B, n, m = K.shape
    if a is None:
        a = np.ones((B, n)) / n
    if b is None:
        b = np.ones((B, m)) / m
    u = np.zeros((B, n), dtype=K.dtype)
    v = np.zeros((B, m), dtype=K.dtype)
    loga, logb = np.log(a), np.log(b)
    it = 0
    for it in range(max_iter):
        u = loga - _lse(K + v[:, None, :], axis=2)
        v = logb - _lse(K + u[:, :, None], axis=1)
        if it % 10 == 0:
            T = np.exp(K + u[:, :, None] + v[:, None, :])
            err = np.max(_norm_batch(np.sum(T, axis=2) - a))
            if err < tol:
                break
    log_T = K + u[:, :, None] + v[:, None, :]
    return {"T": np.exp(log_T), "log_T": log_T, "n_iters": it, "potentials": (u, v)}


def dist_sqeuclidean_batch(X, Y):
    """Batched squared euclidean cost, NOT clamped (ot/batch/_linear.py:79-85)."""
    XX = np.sum(X ** 2, axis=-1, keepdims=True)
    YY = np.transpose(np.sum(Y ** 2, axis=-1, keepdims=True), (0, 2, 1))
    return XX + YY - 2 * np.einsum("bid,bjd->bij", X, Y)


def solve_batch(M, reg, a=None, b=None, max_iter=1000, tol=1e-5,
                solver="log_sinkhorn", reg_type="entropy"):
    """Batched solve + objective values (ot/batch/_linear.py:308-357).

    Returns a dict with the fields of the reference's OTResult: value,
    value_linear, potentials, plan, n_iter.
    """
    B, n, m = M.shape
    if a is None:
        a = np.ones((B, n), dtype=M.dtype) / n
    if b is None:
        b = np.ones((B, m), dtype=M.dtype) / m
    if solver == "log_sinkhorn":
        out = bregman_log_projection_batch(-M / reg, a, b, max_iter, tol)
    elif solver == "sinkhorn":
        out = bregman_projection_batch(np.exp(-M / reg), a, b, max_iter, tol)
    else:
        raise ValueError(f"Unknown solver: {solver}")
    T = out["T"]
    value_linear = np.sum(M * T, axis=(1, 2))
    if reg_type.lower() == "entropy":
        value = value_linear + reg * np.sum(T * np.log(T + 1e-16), axis=(1, 2))
    elif reg_type.lower() == "kl":
        ref = np.einsum("bi,bj->bij", a, b)
        value = value_linear + reg * np.sum(T * np.log(T / ref + 1e-16), axis=(1, 2))
    else:
        value = None
    return {"value": value, "value_linear": value_linear, "plan": T,
            "potentials": out["potentials"], "n_iter": out["n_iters"]}


def dist_batch(X1, X2=None, metric="sqeuclidean", p=2):
    """Batched cost matrices (ot/batch/_linear.py:178-233): sqeuclidean / euclidean without clamp (:79-85),
    minkowski = dist_lp_batch with q = 1 (:46-51), kl = dist_kl_batch with logits_X False, eps 1e-10 (:112-120)."""
    X2 = X1 if X2 is None else X2
    metric = metric.lower()
    if metric == "sqeuclidean":
        return dist_sqeuclidean_batch(X1, X2)
    if metric == "euclidean":
        return np.sqrt(dist_sqeuclidean_batch(X1, X2))
    if metric == "minkowski":
        Mp = (np.abs(X1[:, :, None] - X2[:, None, :]) ** p).sum(axis=-1)
        return Mp ** (1 / p) if p != 1 else Mp
    if metric == "kl":
        eps = 1e-10
        entr_y = np.sum(X2 * np.log(X2 + eps), axis=-1)
        return entr_y[:, None, :] - np.sum(X2[:, None, :] * np.log(X1 + eps)[:, :, None], axis=-1)
    raise ValueError(f"Unknown metric: {metric}")


def loss_linear_batch(M, T):
    """sum_ij M_bij T_bij (ot/batch/_linear.py:147)."""
    return np.sum(M * T, axis=(1, 2))


def entropy_batch(T, eps=1e-16):
    """-sum_ij T log(T + eps) (ot/batch/_utils.py:22)."""
    return -np.sum(T * np.log(T + eps), axis=(1, 2))
