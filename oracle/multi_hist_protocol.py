"""CPU restatement of the device-side state machine of b2ot_multi_scaling_step (TEST INFRASTRUCTURE ONLY).

Mirrors, in numpy, what the kernels of pot_b200/csrc/multi_hist_kernels.cuh do in one step for H histograms on one
kernel matrix - the two passes and the decisions of mh_stats_kernel (modes KNOPP and UNBALANCED) - so that the host
logic around it (pot_b200._dense.multi_hist_steps / multi_hist_select: probe step, chunked enqueue with no-op steps
after a stop, which buffer holds the result) can be exercised on CPU against the unsharded oracle, i.e. the reference
loops ot/bregman/_sinkhorn.py:626-663 (2-D b) and ot/unbalanced/_sinkhorn.py:749-787 (2-D b).
"""

import numpy as np

KNOPP, UNBALANCED = 1, 2


class State:
    """The 16 doubles of `stats` that matter to the host + the error history."""

    def __init__(self):
        self.status, self.at, self.errs = 0, 0, []


def step(mode, K, a, B, up, vp, it, check_every, stop_thr, fi, st):
    """One enqueued step: returns (u_new, v_new); a no-op (returns None, None) once the solve has ended."""
    if st.status != 0:
        return None, None
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        if mode == UNBALANCED:                       # u-then-v (:756-759)
            un = (a[:, None] / (K @ vp)) ** fi[0]
            c = K.T @ un
            vn = (B / c) ** fi[1]
        else:                                        # v-then-u (:629-631)
            c = K.T @ up
            vn = B / c
            un = a[:, None] / (K @ vn)
        e2 = float(np.sum((vp * c - B) ** 2)) if mode == KNOPP else 0.0
    bad = bool(np.any(c == 0) or np.any(~np.isfinite(un)) or np.any(~np.isfinite(vn)))
    if mode == KNOPP:
        if it >= 1 and (it - 1) % check_every == 0:  # the column pass measured the iterate of step it - 1
            err = float(np.sqrt(e2))
            st.errs.append(err)
            if err < stop_thr:
                st.status, st.at = 1, it - 1
                return un, vn
        if bad:
            st.status, st.at = 2, it
    else:
        if bad:
            st.status, st.at = 2, it
            return un, vn
        fmax = lambda x: float(np.max(np.abs(x)))  # noqa: E731
        err = 0.5 * (fmax(un - up) / max(fmax(un), fmax(up), 1.0) + fmax(vn - vp) / max(fmax(vn), fmax(vp), 1.0))
        st.errs.append(err)
        if err < stop_thr:
            st.status, st.at = 1, it
    return un, vn


def solve(mode, K, a, B, u0, v0, max_iter, stop_thr, check_every, fi, poll_every, steps_fn, select_fn):
    """The host loop of pot_b200._dense.multi_hist_solve with the kernels replaced by `step`."""
    us, vs = [u0.copy(), None], [v0.copy(), None]
    st = State()
    steps = steps_fn(mode == KNOPP, max_iter, check_every)
    k = 0
    while k < steps:
        for _ in range(min(poll_every, steps - k)):
            s = k & 1
            un, vn = step(mode, K, a, B, us[s], vs[s], k, check_every, stop_thr, fi, st)
            if un is not None:
                us[s ^ 1], vs[s ^ 1] = un, vn
            k += 1
        if st.status != 0:
            break
    sel, niter, status = select_fn(st.status, st.at, max_iter)
    return us[sel], vs[sel], st.errs, niter, status
