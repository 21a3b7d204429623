"""CPU oracle for the callers right after the Sinkhorn hot path (TEST INFRASTRUCTURE ONLY).

numpy restatement of the reference's `ot.solve` / `ot.solve_sample` Sinkhorn branches, the lazy
`empirical_sinkhorn2` loss, `empirical_sinkhorn_divergence`, `sinkhorn_knopp_unbalanced`,
`sinkhorn_stabilized`, `sinkhorn_epsilon_scaling`, `entropic_gromov_wasserstein` and `solve_gromov_batch`
(SURVEY.md section 8f, ranks 1-4).  Only ``tests/`` (the parity tests and the CPU legs of the measurement helpers
``tests/perf_*.py`` / ``tests/diag_*.py``) may import it; nothing under ``pot_b200/`` does.

Parity status: PINNED.  ``tests/golden/make_golden_solvers.py`` ran the unmodified reference (imported from
/root/reference in the build container) on seeded inputs and committed its outputs as
``tests/golden/golden_solvers.npz``; ``make_golden_stabilized.py``, ``make_golden_gromov.py`` and
``make_golden_gromov_batch.py`` did the same for ``golden_stabilized.npz``, ``golden_gromov.npz`` and
``golden_gromov_batch.npz``.  ``tests/test_oracle_solvers.py``, ``test_oracle_stabilized.py``, ``test_oracle_gromov.py`` and
``test_oracle_gromov_batch.py`` check every function below against them.

Each function cites the reference file:line (relative to /root/reference) it restates.
"""

import warnings

import numpy as np

from . import sinkhorn_oracle as so

__all__ = ["solve", "solve_sample_lazy", "empirical_sinkhorn2_lazy", "empirical_sinkhorn_divergence_lazy",
           "sinkhorn_knopp_unbalanced", "solve_unbalanced", "sinkhorn_stabilized", "sinkhorn_epsilon_scaling",
           "entropic_gromov_wasserstein", "solve_gromov_batch"]


def _kl_div(p, q, eps=1e-16):
    """nx.kl_div with mass=False (ot/backend.py:1542-1546)."""
    return np.sum(p * np.log(p / q + eps))


def solve(M, a=None, b=None, reg=None, reg_type="KL", max_iter=None, tol=None, grad="autodiff"):
    """Balanced entropic branch of ot.solve (ot/solvers.py:300-304, :429-494).

    Returns a dict with plan, potentials (log-scalings), value_linear, value and the envelope gradients
    (dM, da, db) = (plan, reg (f - mean f), reg (g - mean g)) that grad="envelope" attaches (:485-494).
    grad="last_step": max_iter - 1 iterations, then one step in which BOTH updates start from the loop's
    potentials (:463-474).
    """
    M = np.asarray(M)
    n, m = M.shape
    if a is None:
        a = (np.ones(n) / n).astype(M.dtype)
    if b is None:
        b = (np.ones(m) / m).astype(M.dtype)
    rt = reg_type.lower()
    if rt not in ("entropy", "kl"):
        raise NotImplementedError('Not implemented reg_type="{}"'.format(reg_type))
    if max_iter is None:
        max_iter = 1000
    if tol is None:
        tol = 1e-9
    if grad == "last_step":
        if max_iter == 0:
            raise ValueError("The maximum number of iterations must be greater than 0 when using grad=last_step.")
        max_iter = max_iter - 1
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        plan, log = so.sinkhorn_log(a, b, M, reg, numItermax=max_iter, stopThr=tol, log=True)
    f, g = log["log_u"], log["log_v"]
    if max_iter == 0:  # the reference loop body never runs: potentials stay at zero
        f, g = np.zeros(n, dtype=M.dtype), np.zeros(m, dtype=M.dtype)
    if grad == "last_step":
        from scipy.special import logsumexp

        loga, logb = np.log(a), np.log(b)
        v = logb - logsumexp(-M / reg + f[:, None], 0)
        u = loga - logsumexp(-M / reg + g[None, :], 1)
        plan = np.exp(-M / reg + u[:, None] + v[None, :])
        f, g = u, v
    value_linear = np.sum(M * plan)
    if rt == "entropy":
        value = value_linear + reg * np.sum(plan * np.log(plan + 1e-16))
    else:
        value = value_linear + reg * _kl_div(plan, a[:, None] * b[None, :])
    return {"plan": plan, "potentials": (f, g), "value_linear": value_linear, "value": value,
            "grad_M": plan, "grad_a": reg * (f - f.mean()), "grad_b": reg * (g - g.mean())}


def empirical_sinkhorn2_lazy(X_s, X_t, reg, a=None, b=None, numIterMax=10000, stopThr=1e-9, batchSize=100,
                             log=False, warn=True, warmstart=None):
    """Lazy branch of empirical_sinkhorn2 (ot/bregman/_empirical.py:413-483): solve, then
    loss = sum over source blocks of sum(M_block * exp(f_i + g_j - M_block / reg)) (:466-478)."""
    X_s, X_t = np.asarray(X_s), np.asarray(X_t)
    ns, nt = X_s.shape[0], X_t.shape[0]
    if a is None:
        a = (np.ones((ns,)) / ns).astype(X_s.dtype)
    if b is None:
        b = (np.ones((nt,)) / nt).astype(X_s.dtype)
    f, g, hist = so.empirical_sinkhorn_lazy(X_s, X_t, reg, a, b, numIterMax=numIterMax, stopThr=stopThr,
                                            batchSize=batchSize, log=True, warn=warn, warmstart=warmstart)
    bs = batchSize if isinstance(batchSize, int) else batchSize[0]
    loss = 0
    for i in range(0, ns, bs):
        Mb = so.sqeuclidean(X_s[i:i + bs, :], X_t).astype(a.dtype, copy=False)
        loss += np.sum(Mb * np.exp(f[i:i + bs, None] + g[None, :] - Mb / reg))
    return (loss, hist) if log else loss


def empirical_sinkhorn_divergence_lazy(X_s, X_t, reg, a=None, b=None, numIterMax=10000, stopThr=1e-9, warn=True,
                                       warmstart=None, batchSize=100):
    """empirical_sinkhorn_divergence with isLazy=True (ot/bregman/_empirical.py:640-763):
    max(0, W_ab - (W_aa + W_bb) / 2); warm starts (u, u) and (v, v) for the symmetric terms (:646-648)."""
    if warmstart is None:
        ws_a = ws_b = None
    else:
        u, v = warmstart
        ws_a, ws_b = (u, u), (v, v)
    kw = dict(numIterMax=numIterMax, stopThr=stopThr, batchSize=batchSize, warn=warn)
    l_ab = empirical_sinkhorn2_lazy(X_s, X_t, reg, a, b, warmstart=warmstart, **kw)
    l_a = empirical_sinkhorn2_lazy(X_s, X_s, reg, a, a, warmstart=ws_a, **kw)
    l_b = empirical_sinkhorn2_lazy(X_t, X_t, reg, b, b, warmstart=ws_b, **kw)
    return np.maximum(0, l_ab - 0.5 * (l_a + l_b)), (l_ab, l_a, l_b)


def solve_sample_lazy(X_a, X_b, a=None, b=None, reg=None, max_iter=None, tol=None, batch_size=None):
    """Lazy regularised branch of ot.solve_sample (ot/solvers.py:1977-2002): defaults 1000 / 1e-9 / 100;
    value_linear from empirical_sinkhorn2, potentials = (log["u"], log["v"]) = (f, g)."""
    if max_iter is None:
        max_iter = 1000
    if tol is None:
        tol = 1e-9
    if batch_size is None:
        batch_size = 100
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        value_linear, log = empirical_sinkhorn2_lazy(X_a, X_b, reg, a, b, numIterMax=max_iter, stopThr=tol,
                                                     batchSize=batch_size, log=True)
    return {"value_linear": value_linear, "potentials": (log["u"], log["v"]), "niter": log["niter"]}


def sinkhorn_knopp_unbalanced(a, b, M, reg, reg_m, reg_type="kl", c=None, warmstart=None, numItermax=1000,
                              stopThr=1e-6, log=False):
    """Scaling-domain unbalanced Sinkhorn (ot/unbalanced/_sinkhorn.py:690-804).

    K = exp(-M/reg) * c with c = a b^T by default, the one matrix for reg_type "entropy" (:727-740; no c factor for
    2-D b); per iteration u = (a / Kv)^fi_1, v = (b / K^T u)^fi_2 (:749-758); guard -> previous (u, v) and stop
    (:760-772); err = (|u - uprev|_inf / max(|u|_inf, |uprev|_inf, 1) + same for v) / 2 every iteration, `<` stop
    test (:774-787); log: err, logu = log(u + 1e-300), logv, cost, total_cost (:789-814)."""
    a, b, M = (np.asarray(x) for x in (a, b, M))
    dim_a, dim_b = M.shape
    if len(a) == 0:
        a = (np.ones(dim_a) / dim_a).astype(M.dtype)
    if len(b) == 0:
        b = (np.ones(dim_b) / dim_b).astype(M.dtype)
    n_hists = b.shape[1] if b.ndim > 1 else 0
    if isinstance(reg_m, (tuple, list)):
        reg_m1, reg_m2 = (reg_m[0], reg_m[0]) if len(reg_m) == 1 else (reg_m[0], reg_m[1])
    else:
        reg_m1 = reg_m2 = reg_m
    hist = {"err": []}
    if warmstart is None:
        if n_hists:
            u = np.ones((dim_a, 1), dtype=M.dtype)
            v = np.ones((dim_b, n_hists), dtype=M.dtype)
            a = a.reshape(dim_a, 1)
        else:
            u = np.ones(dim_a, dtype=M.dtype)
            v = np.ones(dim_b, dtype=M.dtype)
    else:
        u, v = np.exp(warmstart[0]), np.exp(warmstart[1])
    if reg_type.lower() == "entropy":
        c = np.ones((dim_a, dim_b), dtype=M.dtype)
    if n_hists:
        K = np.exp(-M / reg)
    else:
        c = a[:, None] * b[None, :] if c is None else c
        K = np.exp(-M / reg) * c
    fi_1 = reg_m1 / (reg_m1 + reg) if reg_m1 != float("inf") else 1
    fi_2 = reg_m2 / (reg_m2 + reg) if reg_m2 != float("inf") else 1
    for i in range(numItermax):
        uprev, vprev = u, v
        Kv = np.dot(K, v)
        u = (a / Kv) ** fi_1
        Ktu = np.dot(K.T, u)
        v = (b / Ktu) ** fi_2
        if (np.any(Ktu == 0.0) or np.any(np.isnan(u)) or np.any(np.isnan(v)) or np.any(np.isinf(u))
                or np.any(np.isinf(v))):
            warnings.warn("Numerical errors at iteration %s" % i)
            u, v = uprev, vprev
            break
        err_u = np.max(np.abs(u - uprev)) / max(np.max(np.abs(u)), np.max(np.abs(uprev)), 1.0)
        err_v = np.max(np.abs(v - vprev)) / max(np.max(np.abs(v)), np.max(np.abs(vprev)), 1.0)
        err = 0.5 * (err_u + err_v)
        hist["err"].append(err)
        if err < stopThr:
            break
    hist["logu"] = np.log(u + 1e-300)
    hist["logv"] = np.log(v + 1e-300)
    if n_hists:
        res = np.einsum("ik,ij,jk,ij->k", u, K, v, M)
        return (res, hist) if log else res
    plan = u[:, None] * K * v[None, :]
    if log:
        linear_cost = np.sum(plan * M)
        hist["cost"] = linear_cost
        total = linear_cost + reg * _kl_div(plan, c)
        if reg_m1 != float("inf"):
            total = total + reg_m1 * _kl_div(np.sum(plan, 1), a)
        if reg_m2 != float("inf"):
            total = total + reg_m2 * _kl_div(np.sum(plan, 0), b)
        hist["total_cost"] = total
        return plan, hist
    return plan


def solve_unbalanced(M, a=None, b=None, reg=None, unbalanced=None, max_iter=None, tol=None):
    """Unbalanced KL branch of ot.solve (ot/solvers.py:300-306, :524-556): defaults 1000 / 1e-9,
    c = a b^T, value_linear = log["cost"], value = log["total_cost"], potentials = (logu, logv)."""
    M = np.asarray(M)
    n, m = M.shape
    if a is None:
        a = (np.ones(n) / n).astype(M.dtype)
    if b is None:
        b = (np.ones(m) / m).astype(M.dtype)
    c = a[:, None] * b[None, :]
    if max_iter is None:
        max_iter = 1000
    if tol is None:
        tol = 1e-9
    plan, log = sinkhorn_knopp_unbalanced(a, b, M, reg, unbalanced, reg_type="kl", c=c, numItermax=max_iter,
                                          stopThr=tol, log=True)
    return {"plan": plan, "potentials": (log["logu"], log["logv"]), "value_linear": log["cost"],
            "value": log["total_cost"], "err": log["err"]}


def sinkhorn_stabilized(a, b, M, reg, numItermax=1000, tau=1e3, stopThr=1e-9, warmstart=None, print_period=20,
                        log=False, warn=True):
    """ot.bregman.sinkhorn_stabilized (ot/bregman/_sinkhorn.py:1105-1382), 1-D and 2-D b.

    Scalings start at 1/n, 1/m (:1252-1254); each iteration is v = b / K^T u, u = a / K v (:1270-1271); when
    max|u| or max|v| exceeds tau the scalings are folded into the potentials and K is rebuilt (:1274-1288); every
    print_period iterations err = || sum_i Gamma_ij - b ||_2 (:1303-1304); stop on err <= stopThr (:1316); nan in
    u or v -> previous (u, v) and stop (:1319-1325).  Returns the plan Gamma (:1349 / :1360) and, with log=True,
    the dict of :1332-1340.

    2-D b (n_hists columns, :1230-1234): u, v are (dim, n_hists) and share ONE kernel K.  The absorption is joint and
    keeps the scalings (:1275-1279: alpha += reg * max over the histograms of log u, beta += reg * the SCALAR max of
    log v; the reset of u, v sits in the 1-D branch only); the error is the relative max-change of u and v
    (:1297-1301); the result is the vector of losses <M, Gamma_h> (:1342-1348).
    """
    a, b, M = so._as_arrays(a, b, M)
    n_hists = b.shape[1] if b.ndim > 1 else 0
    if n_hists:
        a = a[:, None]  # :1232
    dim_a, dim_b = len(a), len(b)
    hist = {"err": []}
    if warmstart is None:
        alpha, beta = np.zeros(dim_a, dtype=M.dtype), np.zeros(dim_b, dtype=M.dtype)
    else:
        alpha, beta = warmstart
    if n_hists:
        u = np.ones((dim_a, n_hists), dtype=M.dtype) / dim_a
        v = np.ones((dim_b, n_hists), dtype=M.dtype) / dim_b
    else:
        u = np.ones(dim_a, dtype=M.dtype) / dim_a
        v = np.ones(dim_b, dtype=M.dtype) / dim_b

    def get_K(alpha, beta):
        return np.exp(-(M - alpha.reshape((dim_a, 1)) - beta.reshape((1, dim_b))) / reg)  # :1256-1258

    def get_Gamma(alpha, beta, u, v):
        return np.exp(-(M - alpha.reshape((dim_a, 1)) - beta.reshape((1, dim_b))) / reg
                      + np.log(u.reshape((dim_a, 1))) + np.log(v.reshape((1, dim_b))))  # :1260-1266

    K = get_K(alpha, beta)
    err = 1
    converged = False
    ii = 0
    n_absorb = 0  # not in the reference's log: lets the tests assert that a case really exercises the absorption
    for ii in range(numItermax):
        uprev, vprev = u, v
        v = b / np.dot(K.T, u)
        u = a / np.dot(K, v)
        if np.max(np.abs(u)) > tau or np.max(np.abs(v)) > tau:
            if n_hists:
                alpha, beta = alpha + reg * np.max(np.log(u), 1), beta + reg * np.max(np.log(v))
            else:
                alpha, beta = alpha + reg * np.log(u), beta + reg * np.log(v)
                u = np.ones(dim_a, dtype=M.dtype) / dim_a
                v = np.ones(dim_b, dtype=M.dtype) / dim_b
            K = get_K(alpha, beta)
            n_absorb += 1
        if ii % print_period == 0:
            if n_hists:
                err_u = np.max(np.abs(u - uprev)) / max(np.max(np.abs(u)), np.max(np.abs(uprev)), 1.0)
                err_v = np.max(np.abs(v - vprev)) / max(np.max(np.abs(v)), np.max(np.abs(vprev)), 1.0)
                err = 0.5 * (err_u + err_v)
            else:
                transp = get_Gamma(alpha, beta, u, v)
                err = np.linalg.norm(np.sum(transp, axis=0) - b)
            hist["err"].append(err)
        if err <= stopThr:
            converged = True
            break
        if np.any(np.isnan(u)) or np.any(np.isnan(v)):
            warnings.warn("Numerical errors at iteration %d" % ii)
            u, v = uprev, vprev
            converged = True  # the reference leaves through `break`: no convergence warning
            break
    if not converged and warn:
        warnings.warn("Sinkhorn did not converge. You might want to increase the number of iterations "
                      "`numItermax` or the regularization parameter `reg`.")
    if n_hists:
        G = np.stack([np.sum(get_Gamma(alpha, beta, u[:, h], v[:, h]) * M) for h in range(n_hists)])
        alpha, beta = alpha[:, None], beta[:, None]  # :1333-1334
    else:
        G = get_Gamma(alpha, beta, u, v)
    if not log:
        return G
    hist["n_iter"] = ii
    hist["n_absorb"] = n_absorb
    hist["logu"] = alpha / reg + np.log(u)
    hist["logv"] = beta / reg + np.log(v)
    hist["alpha"] = alpha + reg * np.log(u)
    hist["beta"] = beta + reg * np.log(v)
    hist["warmstart"] = (hist["alpha"], hist["beta"])
    return G, hist


def sinkhorn_epsilon_scaling(a, b, M, reg, numItermax=100, epsilon0=1e4, numInnerItermax=100, tau=1e3,
                             stopThr=1e-9, warmstart=None, print_period=10, log=False, warn=True):
    """ot.bregman.sinkhorn_epsilon_scaling (ot/bregman/_sinkhorn.py:1385-1573): stabilized solves with
    reg_i = (epsilon0 - reg) exp(-i) + reg, warm-started from the previous potentials (:1509-1530); every
    print_period outer iterations err = ||G^T 1 - b||^2 + ||G 1 - a||^2 (:1536-1541); stop when
    err <= stopThr and i > 35 (:1550)."""
    a, b, M = so._as_arrays(a, b, M)
    dim_a, dim_b = len(a), len(b)
    numItermin = 35
    numItermax = max(numItermin, numItermax)
    hist = {"err": []}
    if warmstart is None:
        alpha, beta = np.zeros(dim_a, dtype=M.dtype), np.zeros(dim_b, dtype=M.dtype)
    else:
        alpha, beta = warmstart
    err = 1
    converged = False
    ii = 0
    G = None
    for ii in range(numItermax):
        regi = (epsilon0 - reg) * np.exp(-ii) + reg
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            G, logi = sinkhorn_stabilized(a, b, M, regi, numItermax=numInnerItermax, stopThr=stopThr,
                                          warmstart=(alpha, beta), print_period=20, tau=tau, log=True)
        alpha, beta = logi["alpha"], logi["beta"]
        if ii % print_period == 0:
            err = np.linalg.norm(np.sum(G, axis=0) - b) ** 2 + np.linalg.norm(np.sum(G, axis=1) - a) ** 2
            hist["err"].append(err)
        if err <= stopThr and ii > numItermin:
            converged = True
            break
    if not converged and warn:
        warnings.warn("Sinkhorn did not converge. You might want to increase the number of iterations "
                      "`numItermax` or the regularization parameter `reg`.")
    if not log:
        return G
    hist["alpha"] = alpha
    hist["beta"] = beta
    hist["warmstart"] = (alpha, beta)
    hist["niter"] = ii
    return G, hist


def entropic_gromov_wasserstein(C1, C2, p=None, q=None, loss_fun="square_loss", epsilon=0.1, symmetric=None, G0=None,
                                max_iter=1000, tol=1e-9, solver="PGD", warmstart=False, log=False, **kwargs):
    """ot.gromov.entropic_gromov_wasserstein (ot/gromov/_bregman.py:25-252) on top of the oracle's sinkhorn_knopp.

    init_matrix / tensor_product / gwggrad / gwloss follow ot/gromov/_utils.py:38-354: L(a, b) = f1(a) + f2(b) - h1(a) h2(b),
    constC = f1(C1) p 1^T + 1 q^T f2(C2)^T, tens = constC - h1(C1) T h2(C2)^T, gradient = 2 tens.
    """
    C1, C2 = np.asarray(C1), np.asarray(C2)
    N1, N2 = C1.shape[0], C2.shape[0]
    p = np.ones(N1, dtype=C1.dtype) / N1 if p is None else np.asarray(p)
    q = np.ones(N2, dtype=C2.dtype) / N2 if q is None else np.asarray(q)

    def init_matrix(C1, C2):
        if loss_fun == "square_loss":
            fC1, fC2, hC1, hC2 = C1 ** 2, C2 ** 2, C1, 2 * C2
        else:
            fC1, fC2, hC1, hC2 = C1 * np.log(C1 + 1e-18) - C1, C2, C1, np.log(C2 + 1e-18)
        constC1 = np.dot(np.dot(fC1, p.reshape(-1, 1)), np.ones((1, len(q)), dtype=q.dtype))
        constC2 = np.dot(np.ones((len(p), 1), dtype=p.dtype), np.dot(q.reshape(1, -1), fC2.T))
        return constC1 + constC2, hC1, hC2

    def tensor_product(constC, hC1, hC2, T):
        return constC + (-np.dot(np.dot(hC1, T), hC2.T))

    T = np.outer(p, q) if G0 is None else np.asarray(G0)
    constC, hC1, hC2 = init_matrix(C1, C2)
    if symmetric is None:
        symmetric = np.allclose(C1, C1.T, atol=1e-10) and np.allclose(C2, C2.T, atol=1e-10)
    if not symmetric:
        constCt, hC1t, hC2t = init_matrix(C1.T, C2.T)
    cpt, err = 0, 1
    if warmstart:
        mu = np.zeros(N1, dtype=C1.dtype) - np.log(N1)
        nu = np.zeros(N2, dtype=C2.dtype) - np.log(N2)
    hist = {"err": []}
    while err > tol and cpt < max_iter:
        Tprev = T
        if symmetric:
            tens = 2 * tensor_product(constC, hC1, hC2, T)
        else:
            tens = 0.5 * (2 * tensor_product(constC, hC1, hC2, T) + 2 * tensor_product(constCt, hC1t, hC2t, T))
        if solver == "PPA":
            tens = tens - epsilon * np.log(T)
        if warmstart:
            T, loginn = so.sinkhorn_knopp(p, q, tens, epsilon, log=True, warmstart=(mu, nu), **kwargs)
            mu = epsilon * np.log(loginn["u"])
            nu = epsilon * np.log(loginn["v"])
        else:
            T = so.sinkhorn_knopp(p, q, tens, epsilon, **kwargs)
        if cpt % 10 == 0:
            err = np.linalg.norm(T - Tprev)
            hist["err"].append(err)
        cpt += 1
    if abs(np.sum(T) - 1) > 1e-5:
        warnings.warn("Solver failed to produce a transport plan. You might "
                      "want to increase the regularization parameter `epsilon`.")
    if log:
        hist["gw_dist"] = np.sum(tensor_product(constC, hC1, hC2, T) * T)
        return T, hist
    return T


def solve_gromov_batch(C1, C2, reg=1e-2, a=None, b=None, loss="sqeuclidean", symmetric=None, M=None, alpha=None,
                       T_init=None, max_iter=50, tol=1e-5, max_iter_inner=50, tol_inner=1e-5, logits=None):
    """ot.batch.solve_gromov_batch (ot/batch/_quadratic.py:269-498) on the oracle's bregman_log_projection_batch.

    tensor_batch :17-152 / compute_tensor_batch :504-532, tensor_product_batch :535-572, proximal loop :456-470,
    values :479-495.  Returns a dict with plan, value, value_linear (x (1 - alpha)), value_quad (x alpha), n_iter.
    """
    C1, C2 = np.asarray(C1), np.asarray(C2)
    B, n, m = C1.shape[0], C1.shape[1], C2.shape[1]
    a = np.ones((B, n), dtype=C1.dtype) / n if a is None else np.asarray(a)
    b = np.ones((B, m), dtype=C2.dtype) / m if b is None else np.asarray(b)

    def tr(C):
        return np.transpose(C, (0, 2, 1)) if C.ndim == 3 else np.transpose(C, (0, 2, 1, 3))

    if symmetric is None:
        symmetric = np.allclose(C1, tr(C1), atol=1e-10) and np.allclose(C2, tr(C2), atol=1e-10)
    if loss == "sqeuclidean":
        fC1 = np.sum(C1 ** 2, axis=-1) if C1.ndim == 4 else C1 ** 2
        fC2 = np.sum(C2 ** 2, axis=-1) if C2.ndim == 4 else C2 ** 2
        hC1 = 2 * (C1[..., None] if C1.ndim == 3 else C1)
        hC2 = C2[..., None] if C2.ndim == 3 else C2
    elif loss == "kl":
        fC1 = np.zeros((B, n, n), dtype=C1.dtype)
        fC2 = np.sum(C2 * np.log(C2 + 1e-15), axis=-1)
        hC1 = C1 if logits else np.log(C1 + 1e-15)
        hC2 = C2
    else:
        raise ValueError(f"Unknown loss function: {loss}")
    if not symmetric:
        fC1 = 0.5 * (fC1 + tr(fC1))
        fC2 = 0.5 * (fC2 + tr(fC2))

    def const_from(pa, pb):
        return np.einsum("bij,bj->bi", fC1, pa)[:, :, None] + np.einsum("bij,bj->bi", fC2, pb)[:, None, :]

    constC = const_from(a, b)

    def product(T, recompute_const):
        const = const_from(np.sum(T, axis=2), np.sum(T, axis=1)) if recompute_const else constC
        dot = np.einsum("bijd,bjk->bikd", hC1, T)
        dot = np.sum(np.einsum("bikd,bjkd->bijd", dot, hC2), axis=-1)
        if not symmetric:
            dot_t = np.einsum("bijd,bjk->bikd", tr(hC1), T)
            dot_t = np.sum(np.einsum("bikd,bjkd->bijd", dot_t, tr(hC2)), axis=-1)
            dot = (dot + dot_t) / 2
        return const - dot

    if M is None and alpha is None:
        M = np.zeros((B, n, m), dtype=C1.dtype)
        alpha = 1.0
    if T_init is None:
        const = np.sqrt(np.sum(a, axis=1) * np.sum(b, axis=1))
        T_init = np.einsum("bi,bj->bij", a, b) / const[:, None, None]
    T = T_init
    log_T = np.log(T + 1e-15)
    it = 0
    for it in range(max_iter):
        T_prev = T
        Mk = (1 - alpha) * M + 2 * alpha * product(T, False)
        K = -Mk / reg + log_T
        out = so.bregman_log_projection_batch(K, a, b, max_iter=max_iter_inner, tol=tol_inner)
        T, log_T = out["T"], out["log_T"]
        if np.max(np.mean(np.abs(T_prev - T), axis=(1, 2))) < tol:
            break
    value_linear = np.sum(M * T, axis=(1, 2))
    value_quadratic = np.sum(product(T, True) * T, axis=(1, 2))
    return {"plan": T, "value": (1 - alpha) * value_linear + alpha * value_quadratic,
            "value_linear": value_linear * (1 - alpha), "value_quad": value_quadratic * alpha, "n_iter": it,
            "potentials": out["potentials"]}
