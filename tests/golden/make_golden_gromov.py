"""Golden vectors for ot.gromov.entropic_gromov_wasserstein(2) from the UNMODIFIED reference.

Build-container only (the GPU box has no /root/reference); same recipe as make_golden.py:

    PYTHONPATH=/tmp/potref python tests/golden/make_golden_gromov.py

Fixed work: outer tol = 0 (max_iter outer iterations), inner Sinkhorn with numItermax fixed and stopThr = 0.
Output: tests/golden/golden_gromov.npz
"""

import os
import warnings

import numpy as np

import ot  # the reference
import ot.gromov

HERE = os.path.dirname(os.path.abspath(__file__))
G = {}


def put(case, **fields):
    for k, val in fields.items():
        G[f"{case}/{k}"] = np.asarray(val)


def spaces(seed, n, m, dtype, symmetric=True):
    rng = np.random.RandomState(seed)
    xs, xt = rng.randn(n, 2), rng.randn(m, 3)
    C1, C2 = ot.dist(xs, xs), ot.dist(xt, xt)
    C1, C2 = C1 / C1.max(), C2 / C2.max()
    if not symmetric:
        C1 = C1 + 0.1 * rng.rand(n, n)
        C2 = C2 + 0.1 * rng.rand(m, m)
    wp, wq = rng.rand(n) + 0.5, rng.rand(m) + 0.5
    return C1.astype(dtype), C2.astype(dtype), (wp / wp.sum()).astype(dtype), (wq / wq.sum()).astype(dtype)


# name, dtype, n, m, symmetric, kwargs
CASES = [
    ("pgd_sq_f64", np.float64, 30, 24, True, dict(loss_fun="square_loss", epsilon=0.05, solver="PGD", max_iter=25)),
    ("pgd_kl_f64", np.float64, 26, 31, True, dict(loss_fun="kl_loss", epsilon=0.05, solver="PGD", max_iter=15)),
    ("ppa_sq_f64", np.float64, 30, 24, True, dict(loss_fun="square_loss", epsilon=0.1, solver="PPA", max_iter=25)),
    ("pgd_asym_f64", np.float64, 22, 28, False, dict(loss_fun="square_loss", epsilon=0.05, solver="PGD", max_iter=12)),
    ("pgd_warm_f64", np.float64, 30, 24, True, dict(loss_fun="square_loss", epsilon=0.05, solver="PGD", max_iter=12,
                                                     warmstart=True)),
    ("ppa_warm_f64", np.float64, 25, 25, True, dict(loss_fun="square_loss", epsilon=0.1, solver="PPA", max_iter=12,
                                                     warmstart=True)),
    ("pgd_sq_f32", np.float32, 30, 24, True, dict(loss_fun="square_loss", epsilon=0.1, solver="PGD", max_iter=12)),
]
INNER = dict(numItermax=100, stopThr=0.0)


def main():
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for name, dtype, n, m, sym, kw in CASES:
            C1, C2, p, q = spaces(n * m + len(name), n, m, dtype, sym)
            T, log = ot.gromov.entropic_gromov_wasserstein(C1, C2, p, q, tol=0.0, log=True, **kw, **INNER)
            val = ot.gromov.entropic_gromov_wasserstein2(C1, C2, p, q, tol=0.0, **kw, **INNER)
            put(f"egw_{name}", C1=C1, C2=C2, p=p, q=q, T=T, err=np.asarray(log["err"], dtype=np.float64),
                gw_dist=log["gw_dist"], gw2=val, **{f"kw_{k}": v for k, v in kw.items()})
        # default weights (p = q = None), convergence-driven outer loop
        C1, C2, p, q = spaces(11, 20, 20, np.float64)
        T, log = ot.gromov.entropic_gromov_wasserstein(C1, C2, epsilon=0.1, tol=1e-7, max_iter=200, log=True, **INNER)
        put("egw_default_conv_f64", C1=C1, C2=C2, T=T, err=np.asarray(log["err"], dtype=np.float64),
            gw_dist=log["gw_dist"])
    out = os.path.join(HERE, "golden_gromov.npz")
    np.savez_compressed(out, **G)
    print(f"wrote {len(G)} arrays to {out} ({os.path.getsize(out) / 1024:.0f} KiB)")


if __name__ == "__main__":
    main()
