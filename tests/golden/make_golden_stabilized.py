"""Golden vectors for ot.bregman.sinkhorn_stabilized / sinkhorn_epsilon_scaling from the UNMODIFIED reference.

Build-container only (the GPU box has no /root/reference); same recipe as make_golden.py:

    PYTHONPATH=/tmp/potref python tests/golden/make_golden_stabilized.py

Output: tests/golden/golden_stabilized.npz
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


def problem(seed, n, m, dtype):
    rng = np.random.RandomState(seed)
    xs, xt = rng.randn(n, 2).astype(dtype), (rng.randn(m, 2) + 1.0).astype(dtype)
    M = ot.dist(xs, xt).astype(dtype)
    M = (M / M.max()).astype(dtype)
    wa, wb = rng.rand(n) + 0.1, rng.rand(m) + 0.1
    return M, (wa / wa.sum()).astype(dtype), (wb / wb.sum()).astype(dtype)


# name, dtype, n, m, reg, kwargs
STAB = [
    ("mild_f64", np.float64, 40, 52, 0.1, dict(numItermax=100, stopThr=0.0)),
    ("absorb_f64", np.float64, 60, 50, 0.005, dict(numItermax=200, stopThr=0.0)),
    ("absorb_tau10_f64", np.float64, 60, 50, 0.01, dict(numItermax=150, stopThr=0.0, tau=10.0)),
    ("period7_f64", np.float64, 33, 47, 0.02, dict(numItermax=90, stopThr=0.0, print_period=7)),
    ("conv_f64", np.float64, 48, 40, 0.05, dict(numItermax=2000, stopThr=1e-8, print_period=5)),
    ("conv_absorb_f64", np.float64, 48, 40, 0.01, dict(numItermax=5000, stopThr=1e-7, tau=50.0)),
    ("wide_f64", np.float64, 17, 300, 0.02, dict(numItermax=120, stopThr=0.0)),
    ("mild_f32", np.float32, 40, 52, 0.1, dict(numItermax=100, stopThr=0.0)),
    ("absorb_f32", np.float32, 60, 50, 0.02, dict(numItermax=150, stopThr=0.0, tau=100.0)),
]


def main():
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for name, dtype, n, m, reg, kw in STAB:
            M, a, b = problem(len(name) + n, n, m, dtype)
            P, log = ot.bregman.sinkhorn_stabilized(a, b, M, reg, log=True, **kw)
            case = f"stab_{name}"
            put(case, M=M, a=a, b=b, reg=reg, plan=P, err=np.asarray(log["err"], dtype=np.float64),
                n_iter=log["n_iter"], logu=log["logu"], logv=log["logv"], alpha=log["alpha"], beta=log["beta"],
                **{f"kw_{k}": v for k, v in kw.items()})
            if name == "absorb_f64":  # warm start from the potentials of the solve above, through ot.sinkhorn
                P2, log2 = ot.sinkhorn(a, b, M, reg, method="sinkhorn_stabilized", warmstart=log["warmstart"],
                                       numItermax=40, stopThr=0.0, log=True)
                put(case + "_warm", plan=P2, err=np.asarray(log2["err"], dtype=np.float64), n_iter=log2["n_iter"],
                    alpha=log2["alpha"], beta=log2["beta"])
                put(case + "_loss", value=ot.sinkhorn2(a, b, M, reg, method="sinkhorn_stabilized", numItermax=200,
                                                        stopThr=0.0))
        # epsilon scaling: 40 outer x 100 inner stabilized iterations
        for name, dtype, n, m, reg, kw in [
            ("f64", np.float64, 40, 52, 0.01, dict(numItermax=40, numInnerItermax=100, stopThr=1e-9)),
            ("short_f64", np.float64, 30, 30, 0.05, dict(numItermax=10, numInnerItermax=30, stopThr=0.0, epsilon0=10.0,
                                                        print_period=3)),
        ]:
            M, a, b = problem(7 + n, n, m, dtype)
            P, log = ot.bregman.sinkhorn_epsilon_scaling(a, b, M, reg, log=True, **kw)
            put(f"eps_{name}", M=M, a=a, b=b, reg=reg, plan=P, err=np.asarray(log["err"], dtype=np.float64),
                niter=log["niter"], alpha=log["alpha"], beta=log["beta"], **{f"kw_{k}": v for k, v in kw.items()})
    out = os.path.join(HERE, "golden_stabilized.npz")
    np.savez_compressed(out, **G)
    print(f"wrote {len(G)} arrays to {out} ({os.path.getsize(out) / 1024:.0f} KiB)")


if __name__ == "__main__":
    main()
