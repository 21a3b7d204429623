"""Golden vectors for the callers right after the Sinkhorn path, from the UNMODIFIED reference (PythonOT/POT).

Build-container only (the GPU box has no /root/reference); same recipe as make_golden.py:

    PYTHONPATH=/tmp/potref python tests/golden/make_golden_solvers.py

Covers ot.solve (balanced entropic: KL / entropy, grad detach / envelope / last_step; unbalanced KL),
ot.unbalanced.sinkhorn_knopp_unbalanced, lazy ot.bregman.empirical_sinkhorn2,
ot.bregman.empirical_sinkhorn_divergence (isLazy=True) and ot.solve_sample(lazy=True).

Output: tests/golden/golden_solvers.npz
"""

import os
import warnings

import numpy as np
import torch

import ot  # the reference
import ot.unbalanced

HERE = os.path.dirname(os.path.abspath(__file__))
G = {}


def put(case, **fields):
    for k, val in fields.items():
        if isinstance(val, torch.Tensor):
            val = val.detach().cpu().numpy()
        G[f"{case}/{k}"] = np.asarray(val)


def clouds(seed, n, m, d, dtype):
    rng = np.random.RandomState(seed)
    return rng.randn(n, d).astype(dtype), (rng.randn(m, d) + 1.0).astype(dtype)


def weights(seed, n, dtype):
    rng = np.random.RandomState(seed)
    w = rng.rand(n) + 0.1
    return (w / w.sum()).astype(dtype)


def cost(seed, n, m, dtype):
    xs, xt = clouds(seed, n, m, 2, dtype)
    M = ot.dist(xs, xt).astype(dtype)
    return (M / M.max()).astype(dtype)


def solve_cases():
    for dtype, tag in ((np.float64, "f64"), (np.float32, "f32")):
        for (n, m, reg, its) in ((40, 52, 0.1, 60), (128, 96, 0.05, 150)):
            M = cost(n + m, n, m, dtype)
            a, b = weights(1, n, dtype), weights(2, m, dtype)
            base = f"solve_{tag}_{n}x{m}"
            put(base, M=M, a=a, b=b, reg=reg, its=its)
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                for rt in ("KL", "entropy"):
                    r = ot.solve(M, a, b, reg=reg, reg_type=rt, max_iter=its, tol=0)
                    put(f"{base}/{rt}", plan=r.plan, f=r.potentials[0], g=r.potentials[1], value=r.value,
                        value_linear=r.value_linear)
                    r = ot.solve(M, a, b, reg=reg, reg_type=rt, max_iter=its, tol=0, grad="last_step")
                    put(f"{base}/{rt}_last_step", plan=r.plan, f=r.potentials[0], g=r.potentials[1], value=r.value,
                        value_linear=r.value_linear)
                    # envelope gradients through the reference's torch backend (CPU)
                    Mt = torch.tensor(M, requires_grad=True)
                    at = torch.tensor(a, requires_grad=True)
                    bt = torch.tensor(b, requires_grad=True)
                    r = ot.solve(Mt, at, bt, reg=reg, reg_type=rt, max_iter=its, tol=0, grad="envelope")
                    r.value.backward()
                    put(f"{base}/{rt}_envelope", value=r.value, gM=Mt.grad, ga=at.grad, gb=bt.grad)
                # uniform weights by default, converged
                r = ot.solve(M, reg=reg, tol=1e-9 if dtype == np.float64 else 1e-6, max_iter=3000)
                put(f"{base}/default", plan=r.plan, f=r.potentials[0], g=r.potentials[1], value=r.value,
                    value_linear=r.value_linear)
                # unbalanced KL
                for um in (1.0, (0.5, 2.0)):
                    r = ot.solve(M, a, b, reg=reg, unbalanced=um, max_iter=its, tol=0)
                    utag = "unb1" if um == 1.0 else "unb_pair"
                    put(f"{base}/{utag}", plan=r.plan, f=r.potentials[0], g=r.potentials[1], value=r.value,
                        value_linear=r.value_linear)


def unbalanced_cases():
    for dtype, tag in ((np.float64, "f64"), (np.float32, "f32")):
        for (n, m, reg, its) in ((33, 47, 0.1, 40), (128, 160, 0.05, 100)):
            M = cost(3 * n + m, n, m, dtype)
            a, b = weights(3, n, dtype), weights(4, m, dtype)
            a *= dtype(1.3)  # unbalanced masses
            base = f"unb_{tag}_{n}x{m}"
            put(base, M=M, a=a, b=b, reg=reg, its=its)
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                variants = {
                    "scalar": dict(reg_m=1.0),
                    "pair": dict(reg_m=(0.5, 2.0)),
                    "semi": dict(reg_m=(float("inf"), 1.0)),
                    "entropy": dict(reg_m=1.0, reg_type="entropy"),
                }
                for name, kw in variants.items():
                    P, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, reg, numItermax=its, stopThr=0,
                                                                    log=True, **kw)
                    put(f"{base}/{name}", plan=P, err=lg["err"], logu=lg["logu"], logv=lg["logv"], cost=lg["cost"],
                        total_cost=lg["total_cost"])
                # converged run (default threshold)
                P, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, reg, 1.0, numItermax=5000,
                                                                stopThr=1e-6 if dtype == np.float32 else 1e-9,
                                                                log=True)
                put(f"{base}/conv", plan=P, nerr=len(lg["err"]), err_last=lg["err"][-1], logu=lg["logu"],
                    logv=lg["logv"], total_cost=lg["total_cost"])
                # custom reference matrix c and a warm start
                rng = np.random.RandomState(5)
                c = (rng.rand(n, m) + 0.5).astype(dtype)
                c /= c.sum()
                ws = (rng.randn(n).astype(dtype) * dtype(0.1), rng.randn(m).astype(dtype) * dtype(0.1))
                P, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, reg, 1.0, c=c, warmstart=ws,
                                                                numItermax=its, stopThr=0, log=True)
                put(f"{base}/c_warm", c=c, ws_u=ws[0], ws_v=ws[1], plan=P, err=lg["err"], logu=lg["logu"],
                    logv=lg["logv"], cost=lg["cost"], total_cost=lg["total_cost"])
                # 2-D b: losses only
                b2 = np.stack([weights(6, m, dtype), weights(7, m, dtype) * dtype(0.8)], 1)
                L, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b2, M, reg, 1.0, numItermax=its, stopThr=0,
                                                                log=True)
                put(f"{base}/multib", b2=b2, loss=L, logu=lg["logu"], logv=lg["logv"], err=lg["err"])
    # numerical guard: tiny reg underflows K
    n = m = 20
    xs, xt = clouds(5, n, m, 2, np.float64)
    M = ot.dist(xs, xt) + 50.0
    a = np.ones(n) / n
    b = np.ones(m) / m
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        P, lg = ot.unbalanced.sinkhorn_knopp_unbalanced(a, b, M, 5e-2, 1.0, numItermax=50, stopThr=0, log=True)
        msgs = [str(x.message) for x in w]
    put("unb_guard", M=M, a=a, b=b, reg=5e-2, plan=P, logu=lg["logu"], logv=lg["logv"], nerr=len(lg["err"]),
        warned=int(any("Numerical errors" in s for s in msgs)))


def empirical_cases():
    for dtype, tag in ((np.float64, "f64"), (np.float32, "f32")):
        n, m, d, reg, its = 70, 90, 3, 0.5, 40
        xs, xt = clouds(11, n, m, d, dtype)
        a, b = weights(8, n, dtype), weights(9, m, dtype)
        base = f"emp2_{tag}"
        put(base, xs=xs, xt=xt, a=a, b=b, reg=reg, its=its)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            L, lg = ot.bregman.empirical_sinkhorn2(xs, xt, reg, a, b, numIterMax=its, stopThr=-1, isLazy=True,
                                                   batchSize=32, log=True)
            put(base + "/loss", loss=L, f=lg["u"], g=lg["v"], niter=lg["niter"])
            D, lg = ot.bregman.empirical_sinkhorn_divergence(xs, xt, reg, a, b, numIterMax=its, stopThr=-1,
                                                             isLazy=True, batchSize=32, log=True)
            put(base + "/div", div=D, l_ab=lg["sinkhorn_loss_ab"], l_a=lg["sinkhorn_loss_a"],
                l_b=lg["sinkhorn_loss_b"])
            r = ot.solve_sample(xs, xt, a, b, reg=reg, lazy=True, max_iter=its, tol=-1, batch_size=32)
            put(base + "/solve_sample", value_linear=r.value_linear, f=r.potentials[0], g=r.potentials[1],
                block=r.lazy_plan[3:9, 5:12])
            r = ot.solve_sample(xs, xt, a, b, reg=reg, max_iter=its, tol=0)
            put(base + "/solve_sample_dense", value=r.value, value_linear=r.value_linear, plan=r.plan,
                f=r.potentials[0], g=r.potentials[1])


if __name__ == "__main__":
    solve_cases()
    unbalanced_cases()
    empirical_cases()
    out = os.path.join(HERE, "golden_solvers.npz")
    np.savez_compressed(out, **G)
    print(f"wrote {out}: {len(G)} arrays, {os.path.getsize(out) / 1024:.0f} KiB")
