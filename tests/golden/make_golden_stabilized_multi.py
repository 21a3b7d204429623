"""Golden vectors for ot.bregman.sinkhorn_stabilized with SEVERAL target histograms (2-D b), UNMODIFIED reference.

Build-container only (the GPU box has no /root/reference); the reference is the in-place build of oracle/build_ref.py:

    PYTHONPATH=oracle/_ref python tests/golden/make_golden_stabilized_multi.py

Output: tests/golden/golden_stabilized_multi.npz
"""

import os
import warnings

import numpy as np

import ot  # the reference

HERE = os.path.dirname(os.path.abspath(__file__))
G = {}


def put(case, **fields):
    for k, val in fields.items():
        G[f"{case}/{k}"] = np.asarray(val)


def problem(seed, n, m, h, dtype):
    rng = np.random.RandomState(seed)
    xs, xt = rng.randn(n, 2).astype(dtype), (rng.randn(m, 2) + 1.0).astype(dtype)
    M = ot.dist(xs, xt).astype(dtype)
    M = (M / M.max()).astype(dtype)
    wa, wb = rng.rand(n) + 0.1, rng.rand(m, h) + 0.1
    return M, (wa / wa.sum()).astype(dtype), (wb / wb.sum(0)).astype(dtype)


# name, dtype, n, m, n_hists, reg, kwargs
CASES = [
    ("mild_f64", np.float64, 40, 52, 3, 0.1, dict(numItermax=100, stopThr=0.0)),
    ("absorb_f64", np.float64, 60, 50, 4, 0.005, dict(numItermax=200, stopThr=0.0)),
    ("absorb_tau10_f64", np.float64, 60, 50, 2, 0.01, dict(numItermax=150, stopThr=0.0, tau=10.0)),
    ("period7_f64", np.float64, 33, 47, 5, 0.02, dict(numItermax=90, stopThr=0.0, print_period=7)),
    ("conv_f64", np.float64, 48, 40, 3, 0.05, dict(numItermax=2000, stopThr=1e-8, print_period=5)),
    ("conv_absorb_f64", np.float64, 48, 40, 3, 0.01, dict(numItermax=5000, stopThr=1e-7, tau=50.0)),
    ("wide_f64", np.float64, 17, 300, 9, 0.02, dict(numItermax=120, stopThr=0.0)),
    ("one_hist_f64", np.float64, 30, 31, 1, 0.05, dict(numItermax=60, stopThr=0.0)),
    ("mild_f32", np.float32, 40, 52, 3, 0.1, dict(numItermax=100, stopThr=0.0)),
    ("absorb_f32", np.float32, 60, 50, 3, 0.02, dict(numItermax=150, stopThr=0.0, tau=100.0)),
]


def main():
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for name, dtype, n, m, h, reg, kw in CASES:
            M, a, b = problem(len(name) + n, n, m, h, dtype)
            res, log = ot.bregman.sinkhorn_stabilized(a, b, M, reg, log=True, **kw)
            case = f"stabm_{name}"
            put(case, M=M, a=a, b=b, reg=reg, losses=res, err=np.asarray(log["err"], dtype=np.float64),
                n_iter=log["n_iter"], logu=log["logu"], logv=log["logv"], alpha=log["alpha"], beta=log["beta"],
                **{f"kw_{k}": v for k, v in kw.items()})
            if name == "absorb_f64":  # warm start (1-D potentials, :1245-1248) and the sinkhorn2 route
                ws = (log["alpha"][:, 0].copy(), log["beta"][:, 0].copy())
                r2, log2 = ot.sinkhorn(a, b, M, reg, method="sinkhorn_stabilized", warmstart=ws, numItermax=40,
                                       stopThr=0.0, log=True)
                put(case + "_warm", ws_alpha=ws[0], ws_beta=ws[1], losses=r2,
                    err=np.asarray(log2["err"], dtype=np.float64), n_iter=log2["n_iter"], alpha=log2["alpha"],
                    beta=log2["beta"])
                put(case + "_loss", value=ot.sinkhorn2(a, b, M, reg, method="sinkhorn_stabilized", numItermax=200,
                                                        stopThr=0.0))
    out = os.path.join(HERE, "golden_stabilized_multi.npz")
    np.savez_compressed(out, **G)
    print(f"wrote {len(G)} arrays to {out} ({os.path.getsize(out) / 1024:.0f} KiB)")


if __name__ == "__main__":
    main()
