"""CPU restatement of the ROW-SHARDED solve protocol (TEST INFRASTRUCTURE ONLY).

Mirrors, in numpy, what each rank's kernels do in a row-sharded run (pot_b200/csrc/dense_kernels.cuh:
fused sweep -> reduce_partials -> all-reduce of m+1 values -> finalize) so that the exchange pattern and
the on-device stop / guard logic can be exercised with world_size > 1 on CPU (gloo).  `allreduce(x, op)`
sums (or maxes) a numpy vector across ranks in place.  Parity target: the unsharded oracle
(oracle/sinkhorn_oracle.py), i.e. the reference loops ot/bregman/_sinkhorn.py:626-670 and :880-906.
"""

import numpy as np
from scipy.special import logsumexp


def knopp_sharded(a_loc, b, M_loc, reg, n_total, numItermax, stopThr, allreduce):
    """Returns (u_loc, v, err list, niter, status) with status 1 converged / 2 guard / 3 max-iter."""
    dt = M_loc.dtype
    m = b.shape[0]
    K = np.exp(M_loc / (-reg))
    u = np.full(a_loc.shape[0], 1.0 / n_total, dtype=dt)
    v = np.full(m, 1.0 / m, dtype=dt)
    # seeding pass: c0 = K^T u0 (the first line of the reference loop, :629)
    buf = np.concatenate([K.T @ u, np.zeros(1, dtype=dt)])
    allreduce(buf, "sum")
    c = buf[:m]
    v_prev, u_prev = v, u
    with np.errstate(divide="ignore", invalid="ignore"):
        v = b / c
    cur_flags = bool(np.any(c == 0) or np.any(~np.isfinite(v)))
    errs = []
    for ii in range(numItermax):
        with np.errstate(divide="ignore", invalid="ignore"):
            u_new = a_loc / (K @ v)                       # fused sweep, phase A
        part = K.T @ u_new                                # fused sweep, phase B (this rank's column sums)
        buf = np.concatenate([part, np.array([0.0 if np.all(np.isfinite(u_new)) else 1.0], dtype=dt)])
        allreduce(buf, "sum")                             # the ONE exchange of the iteration
        c, uflag = buf[:m], buf[m] != 0
        if cur_flags or uflag:                            # finalize: guard -> revert and stop (:633-645)
            return u_prev, v_prev, errs, ii, 2
        u_prev, u = u, u_new
        if ii % 10 == 0:
            err = float(np.linalg.norm(v * c - b))        # column marginal of diag(u) K diag(v)
            errs.append(err)
            if err < stopThr:
                return u, v, errs, ii, 1
        if ii + 1 >= numItermax:
            return u, v, errs, ii, 3
        v_prev = v
        with np.errstate(divide="ignore", invalid="ignore"):
            v = b / c
        cur_flags = bool(np.any(c == 0) or np.any(~np.isfinite(v)))
    raise AssertionError("unreachable")


def log_sharded(a_loc, b, M_loc, reg, numItermax, stopThr, allreduce):
    """Log-domain protocol: exact column LSE (max all-reduce + sum all-reduce) for the seeding pass,
    then one sum all-reduce of the row-frame column marginals per iteration."""
    dt = M_loc.dtype
    m = b.shape[0]
    Mr = -M_loc / reg
    loga, logb = np.log(a_loc), np.log(b)
    u = np.zeros(a_loc.shape[0], dtype=dt)
    # seeding pass: v = logb - LSE_i(Mr + u), combined across ranks as (max, rescaled sum)
    mx_loc = np.max(Mr + u[:, None], axis=0)
    mx = mx_loc.copy()
    allreduce(mx, "max")
    s = np.sum(np.exp(Mr + u[:, None] - mx_loc[None, :]), axis=0) * np.exp(mx_loc - mx)
    allreduce(s, "sum")
    v = logb - (mx + np.log(s))
    errs = []
    for ii in range(numItermax):
        u = loga - logsumexp(Mr + v[None, :], axis=1)                 # local rows: no exchange
        cm = np.sum(np.exp(Mr + u[:, None] + v[None, :]), axis=0)    # row-frame column marginals
        allreduce(cm, "sum")
        if ii % 10 == 0:
            err = float(np.linalg.norm(cm - b))
            errs.append(err)
            if err < stopThr:
                return u, v, errs, ii, 1
        if ii + 1 >= numItermax:
            return u, v, errs, ii, 3
        v = v + (logb - np.log(cm))
    raise AssertionError("unreachable")


def unbalanced_sharded(a_loc, b, K_loc, fi_1, fi_2, numItermax, stopThr, allreduce):
    """Row-sharded unbalanced scaling iteration (pot_b200/csrc/dense.cu::launch_finalize_unbal; reference loop
    ot/unbalanced/_sinkhorn.py:749-787).  K_loc: this rank's rows of exp(-M/reg) * c.  Two exchanges per iteration:
    SUM of the m column sums + the "u not finite" flag, MAX of the three u statistics of the error.
    Returns (u_loc, v, err list, n_iterations_done, status) with status 1 converged / 2 guard / 3 max-iter."""
    dt = K_loc.dtype
    m = b.shape[0]
    u = np.ones(a_loc.shape[0], dtype=dt)
    v = np.ones(m, dtype=dt)
    errs = []
    for ii in range(numItermax):
        with np.errstate(divide="ignore", invalid="ignore"):
            u_new = (a_loc / (K_loc @ v)) ** fi_1                      # fused sweep, phase A (local rows)
            part = K_loc.T @ u_new                                     # fused sweep, phase B
        buf = np.concatenate([part, np.array([0.0 if np.all(np.isfinite(u_new)) else 1.0], dtype=dt)])
        allreduce(buf, "sum")                                          # exchange 1
        c, uflag = buf[:m], buf[m] != 0
        with np.errstate(invalid="ignore"):                            # inf - inf on a guarded iteration
            stats = np.array([np.max(np.abs(u_new - u)), np.max(np.abs(u_new)), np.max(np.abs(u))], dtype=dt)
        stats = np.where(np.isnan(stats), 0.0, stats).astype(dt)       # fmax skips nan on the device
        allreduce(stats, "max")                                        # exchange 2
        with np.errstate(divide="ignore", invalid="ignore"):
            v_new = (b / c) ** fi_2
        if uflag or np.any(c == 0) or np.any(~np.isfinite(v_new)):     # guard: previous (u, v), stop (:760-772)
            return u, v, errs, ii, 2
        err_u = stats[0] / max(stats[1], stats[2], 1.0)
        err_v = np.max(np.abs(v_new - v)) / max(np.max(np.abs(v_new)), np.max(np.abs(v)), 1.0)
        u, v = u_new, v_new
        err = 0.5 * (err_u + err_v)
        errs.append(float(err))
        if err < stopThr:
            return u, v, errs, ii, 1
    return u, v, errs, numItermax - 1, 3
