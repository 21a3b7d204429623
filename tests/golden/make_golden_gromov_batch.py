"""Golden vectors for ot.batch.solve_gromov_batch from the UNMODIFIED reference.

Build-container only (the GPU box has no /root/reference); same recipe as make_golden.py:

    PYTHONPATH=/tmp/potref python tests/golden/make_golden_gromov_batch.py

Fixed work: tol = 0 and tol_inner = 0 (max_iter proximal steps of max_iter_inner projections each).
Output: tests/golden/golden_gromov_batch.npz
"""

import os
import warnings

import numpy as np

import ot  # the reference
import ot.batch

HERE = os.path.dirname(os.path.abspath(__file__))
G = {}


def put(case, **fields):
    for k, val in fields.items():
        G[f"{case}/{k}"] = np.asarray(val)


def structures(seed, B, n, m, dtype, d=None, symmetric=True, positive=False):
    rng = np.random.RandomState(seed)
    if d is None:
        xs, xt = rng.randn(B, n, 2), rng.randn(B, m, 3)
        C1 = ((xs[:, :, None] - xs[:, None]) ** 2).sum(-1)
        C2 = ((xt[:, :, None] - xt[:, None]) ** 2).sum(-1)
        C1, C2 = C1 / C1.max(), C2 / C2.max()
        if not symmetric:
            C1 = C1 + 0.1 * rng.rand(B, n, n)
            C2 = C2 + 0.1 * rng.rand(B, m, m)
    else:
        C1, C2 = rng.rand(B, n, n, d), rng.rand(B, m, m, d)
        if symmetric:
            C1, C2 = 0.5 * (C1 + C1.transpose(0, 2, 1, 3)), 0.5 * (C2 + C2.transpose(0, 2, 1, 3))
        if positive:  # edge-label distributions for the KL loss
            C1, C2 = C1 / C1.sum(-1, keepdims=True), C2 / C2.sum(-1, keepdims=True)
    wa, wb = rng.rand(B, n) + 0.5, rng.rand(B, m) + 0.5
    M = rng.rand(B, n, m)
    return (C1.astype(dtype), C2.astype(dtype), (wa / wa.sum(1, keepdims=True)).astype(dtype),
            (wb / wb.sum(1, keepdims=True)).astype(dtype), M.astype(dtype))


FIX = dict(tol=0.0, tol_inner=0.0)
# name, dtype, B, n, m, d, symmetric, positive, fused alpha, kwargs
CASES = [
    ("gw_f64", np.float64, 4, 24, 24, None, True, False, None, dict(reg=0.05, max_iter=8, max_iter_inner=40)),
    ("gw_rect_f64", np.float64, 3, 20, 28, None, True, False, None, dict(reg=0.05, max_iter=6, max_iter_inner=30)),
    ("fgw_f64", np.float64, 3, 24, 24, None, True, False, 0.5, dict(reg=0.05, max_iter=6, max_iter_inner=40)),
    ("gw_asym_f64", np.float64, 3, 18, 22, None, False, False, None, dict(reg=0.05, max_iter=6, max_iter_inner=30)),
    ("gw_feat_f64", np.float64, 2, 16, 20, 3, True, False, None, dict(reg=0.1, max_iter=5, max_iter_inner=30)),
    ("gw_kl_f64", np.float64, 2, 16, 16, 4, True, True, None, dict(reg=0.1, max_iter=5, max_iter_inner=30, loss="kl",
                                                                   logits=False)),
    ("gw_f32", np.float32, 4, 24, 24, None, True, False, None, dict(reg=0.1, max_iter=6, max_iter_inner=40)),
]


def main():
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for name, dtype, B, n, m, d, sym, pos, alpha, kw in CASES:
            C1, C2, a, b, M = structures(B * n + m, B, n, m, dtype, d, sym, pos)
            extra = dict(M=M, alpha=alpha) if alpha is not None else {}
            res = ot.batch.solve_gromov_batch(C1, C2, a=a, b=b, grad="detach", **extra, **kw, **FIX)
            put(f"gwb_{name}", C1=C1, C2=C2, a=a, b=b, M=M, alpha=-1.0 if alpha is None else alpha, plan=res.plan,
                value=res.value, value_linear=res.value_linear, value_quad=res.value_quad, n_iter=res.log["n_iter"],
                **{f"kw_{k}": v for k, v in kw.items()})
        # default weights, convergence-driven loops
        C1, C2, a, b, M = structures(5, 3, 20, 20, np.float64)
        res = ot.batch.solve_gromov_batch(C1, C2, reg=0.1, grad="detach", max_iter=50, tol=1e-6, max_iter_inner=50,
                                          tol_inner=1e-7)
        put("gwb_default_conv_f64", C1=C1, C2=C2, plan=res.plan, value=res.value, n_iter=res.log["n_iter"])
    out = os.path.join(HERE, "golden_gromov_batch.npz")
    np.savez_compressed(out, **G)
    print(f"wrote {len(G)} arrays to {out} ({os.path.getsize(out) / 1024:.0f} KiB)")


if __name__ == "__main__":
    main()
