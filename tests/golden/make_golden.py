"""Generate golden vectors by running the UNMODIFIED reference (PythonOT/POT).

Run in the build container only (the GPU box has no /root/reference):

    D=/tmp/potref; mkdir -p $D && cp -r /root/reference/{ot,setup.py,README.md,pyproject.toml,setup.cfg} $D/
    mkdir -p $D/deps && ln -s /root/reference/deps/eigen $D/deps/eigen && chmod -R u+w $D
    (cd $D && python setup.py build_ext --inplace -j8)       # ~4 min, only to make `import ot` work
    PYTHONPATH=/tmp/potref python tests/golden/make_golden.py

The hot path itself is pure Python in the reference (ot/bregman/_sinkhorn.py,
ot/bregman/_empirical.py, ot/batch/_utils.py, ot/batch/_linear.py); the build
step exists only because ``ot/__init__.py`` imports the Cython EMD modules.

Outputs: tests/golden/golden_small.npz  (inputs + reference outputs, small n)
         tests/golden/golden_c1.npz     (BASELINE config 1: n=m=1000 fp64, seeds + u, v, err)
"""

import os
import sys
import warnings

import numpy as np

import ot  # the reference
import ot.batch

HERE = os.path.dirname(os.path.abspath(__file__))
G = {}


def put(case, **fields):
    for k, val in fields.items():
        G[f"{case}/{k}"] = np.asarray(val)


def clouds(seed, n, m, d, dtype):
    rng = np.random.RandomState(seed)
    xs = rng.randn(n, d).astype(dtype)
    xt = (rng.randn(m, d) + 1.0).astype(dtype)
    return xs, xt


def weights(seed, n, dtype, uniform):
    if uniform:
        return (np.ones(n) / n).astype(dtype)
    rng = np.random.RandomState(seed)
    w = rng.rand(n) + 0.1
    return (w / w.sum()).astype(dtype)


def dense_cases():
    for dtype, tag in ((np.float64, "f64"), (np.float32, "f32")):
        for (n, m, reg, uni, its) in ((50, 60, 0.1, True, 100), (33, 47, 0.5, False, 57), (128, 96, 0.05, False, 200)):
            xs, xt = clouds(n + m, n, m, 2, dtype)
            M = ot.dist(xs, xt).astype(dtype)
            M = (M / M.max()).astype(dtype)
            a = weights(1, n, dtype, uni)
            b = weights(2, m, dtype, uni)
            base = f"dense_{tag}_{n}x{m}"
            put(base, a=a, b=b, M=M, reg=reg, its=its)
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                # fixed iteration count (stopThr=0 never triggers: strict <)
                P, lg = ot.bregman.sinkhorn_knopp(a, b, M, reg, numItermax=its, stopThr=0, log=True)
                put(base + "/knopp_fixed", plan=P, u=lg["u"], v=lg["v"], err=lg["err"], niter=lg["niter"])
                P, lg = ot.bregman.sinkhorn_log(a, b, M, reg, numItermax=its, stopThr=0, log=True)
                put(base + "/log_fixed", plan=P, log_u=lg["log_u"], log_v=lg["log_v"], err=lg["err"], niter=lg["niter"])
                # converged with the default-ish threshold
                thr = 1e-9 if dtype == np.float64 else 1e-6
                P, lg = ot.bregman.sinkhorn_knopp(a, b, M, reg, numItermax=2000, stopThr=thr, log=True)
                put(base + "/knopp_conv", plan=P, u=lg["u"], v=lg["v"], err=lg["err"], niter=lg["niter"], thr=thr)
                P, lg = ot.bregman.sinkhorn_log(a, b, M, reg, numItermax=2000, stopThr=thr, log=True)
                put(base + "/log_conv", plan=P, log_u=lg["log_u"], log_v=lg["log_v"], err=lg["err"], niter=lg["niter"], thr=thr)
                # warm start = (log u, log v) of a short run
                _, lg0 = ot.bregman.sinkhorn_knopp(a, b, M, reg, numItermax=7, stopThr=0, log=True)
                ws = (np.log(lg0["u"]), np.log(lg0["v"]))
                P, lg = ot.bregman.sinkhorn_knopp(a, b, M, reg, numItermax=20, stopThr=0, log=True, warmstart=ws)
                put(base + "/knopp_warm", ws_u=ws[0], ws_v=ws[1], plan=P, u=lg["u"], v=lg["v"], err=lg["err"], niter=lg["niter"])
                P, lg = ot.bregman.sinkhorn_log(a, b, M, reg, numItermax=20, stopThr=0, log=True, warmstart=ws)
                put(base + "/log_warm", plan=P, log_u=lg["log_u"], log_v=lg["log_v"], err=lg["err"], niter=lg["niter"])
                # sinkhorn2 losses
                put(base + "/loss", knopp=ot.sinkhorn2(a, b, M, reg, numItermax=its, stopThr=0),
                    log=ot.sinkhorn2(a, b, M, reg, method="sinkhorn_log", numItermax=its, stopThr=0))

    # multi-histogram b (2-D): losses only (ot/bregman/_sinkhorn.py:676-681, :819-854)
    n, m = 40, 30
    xs, xt = clouds(7, n, m, 2, np.float64)
    M = ot.dist(xs, xt)
    M /= M.max()
    a = weights(3, n, np.float64, False)
    rng = np.random.RandomState(11)
    b2 = rng.rand(m, 3) + 0.1
    b2 /= b2.sum(0, keepdims=True)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        L, lg = ot.bregman.sinkhorn_knopp(a, b2, M, 0.1, numItermax=100, stopThr=0, log=True)
        put("multib", a=a, b=b2, M=M, reg=0.1, its=100, knopp_loss=L, knopp_u=lg["u"], knopp_v=lg["v"])
        L, lg = ot.bregman.sinkhorn_log(a, b2, M, 0.1, numItermax=100, stopThr=0, log=True)
        put("multib", log_loss=L, log_u=lg["log_u"], log_v=lg["log_v"])

    # numerical guard: tiny reg underflows K (ot/bregman/_sinkhorn.py:633-645)
    n, m = 20, 20
    xs, xt = clouds(5, n, m, 2, np.float64)
    M = ot.dist(xs, xt)
    a = weights(0, n, np.float64, True)
    b = weights(0, m, np.float64, True)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        P, lg = ot.bregman.sinkhorn_knopp(a, b, M, 1e-3, numItermax=50, stopThr=0, log=True)
        msgs = [str(x.message) for x in w]
    put("guard", a=a, b=b, M=M, reg=1e-3, plan=P, u=lg["u"], v=lg["v"], niter=lg["niter"],
        err=lg["err"], warned=int(any("numerical errors" in s for s in msgs)))

    # empty a / b mean uniform (ot/bregman/_sinkhorn.py:592-595)
    n, m = 25, 35
    xs, xt = clouds(9, n, m, 2, np.float64)
    M = ot.dist(xs, xt)
    M /= M.max()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = ot.sinkhorn([], [], M, 0.2, numItermax=30, stopThr=0, log=True)
    put("empty", M=M, reg=0.2, plan=P, u=lg["u"], v=lg["v"])


def empirical_cases():
    for dtype, tag in ((np.float64, "f64"), (np.float32, "f32")):
        for (ns, nt, d, reg, bsz, its) in ((40, 50, 3, 0.5, (1, 3), 30), (64, 32, 8, 1.0, 5, 40), (37, 41, 2, 0.3, 100, 20)):
            xs, xt = clouds(ns * nt, ns, nt, d, dtype)
            a = weights(4, ns, dtype, False)
            b = weights(5, nt, dtype, False)
            base = f"emp_{tag}_{ns}x{nt}x{d}"
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                # `err <= stopThr` (ot/bregman/_empirical.py:255): -1 never stops
                f, g, lg = ot.bregman.empirical_sinkhorn(xs, xt, reg, a=a, b=b, numIterMax=its, stopThr=-1,
                                                         isLazy=True, batchSize=bsz, log=True)
                put(base, xs=xs, xt=xt, a=a, b=b, reg=reg, its=its,
                    bs=np.asarray(bsz if isinstance(bsz, tuple) else (bsz, bsz)),
                    f=f, g=g, err=lg["err"], niter=lg["niter"], plan=lg["lazy_plan"][:],
                    plan_block=lg["lazy_plan"][3:11, 5:20])
                thr = 1e-9 if dtype == np.float64 else 1e-5
                f, g, lg = ot.bregman.empirical_sinkhorn(xs, xt, reg, a=a, b=b, numIterMax=3000, stopThr=thr,
                                                         isLazy=True, batchSize=bsz, log=True)
                put(base + "/conv", f=f, g=g, err=lg["err"], niter=lg["niter"], thr=thr)
    # default (uniform) weights + non-lazy path == sinkhorn on dist
    xs, xt = clouds(77, 30, 20, 2, np.float64)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P = ot.bregman.empirical_sinkhorn(xs, xt, 0.7, numIterMax=50, stopThr=0)
        f, g = ot.bregman.empirical_sinkhorn(xs, xt, 0.7, numIterMax=50, stopThr=-1, isLazy=True, batchSize=7)
    put("emp_default", xs=xs, xt=xt, reg=0.7, plan=P, f=f, g=g)


def batch_cases():
    for dtype, tag in ((np.float64, "f64"), (np.float32, "f32")):
        rng = np.random.RandomState(3)
        B, n, m, d = 5, 24, 31, 3
        X = rng.randn(B, n, d).astype(dtype)
        Y = (rng.randn(B, m, d) + 0.5).astype(dtype)
        M = ot.batch.dist_batch(X, Y)
        M = (M / M.max()).astype(dtype)
        a = (rng.rand(B, n) + 0.1).astype(dtype)
        a /= a.sum(1, keepdims=True)
        b = (rng.rand(B, m) + 0.1).astype(dtype)
        b /= b.sum(1, keepdims=True)
        base = f"batch_{tag}"
        put(base, X=X, Y=Y, M=M, a=a, b=b, reg=0.1)
        for solver in ("log_sinkhorn", "sinkhorn"):
            for reg_type in ("entropy", "kl"):
                # fixed count: `err < tol` with tol=0 never stops (ot/batch/_utils.py:148, :277)
                r = ot.batch.solve_batch(M, 0.1, a, b, max_iter=40, tol=0, solver=solver, reg_type=reg_type)
                put(f"{base}/{solver}_{reg_type}_fixed", plan=r.plan, value=r.value, value_linear=r.value_linear,
                    pa=r.potentials[0], pb=r.potentials[1], n_iter=r.log["n_iter"])
                r = ot.batch.solve_batch(M, 0.1, a, b, max_iter=1000, tol=1e-5, solver=solver, reg_type=reg_type)
                put(f"{base}/{solver}_{reg_type}_conv", plan=r.plan, value=r.value, value_linear=r.value_linear,
                    pa=r.potentials[0], pb=r.potentials[1], n_iter=r.log["n_iter"])
        # uniform default weights
        r = ot.batch.solve_batch(M, 0.1, max_iter=25, tol=0)
        put(f"{base}/uniform", plan=r.plan, value=r.value, pa=r.potentials[0], pb=r.potentials[1])


def doctest_vectors():
    a = [0.5, 0.5]
    b = [0.5, 0.5]
    M = [[0.0, 1.0], [1.0, 0.0]]
    put("doctest", sinkhorn_2x2=ot.sinkhorn(a, b, M, 1),                     # _sinkhorn.py:123-129
        sinkhorn2_2x2=ot.sinkhorn2(a, b, M, 1),                              # _sinkhorn.py:342-347
        knopp_2x2=ot.bregman.sinkhorn_knopp(a, b, M, 1),                     # _sinkhorn.py:563-569
        log_2x2=ot.bregman.sinkhorn_log(a, b, M, 1))                         # _sinkhorn.py:766-772
    xs = np.reshape(np.arange(2, dtype=np.float64), (2, 1))
    xt = np.reshape(np.arange(0, 2, dtype=np.float64), (2, 1))
    put("doctest", emp_2x2=ot.bregman.empirical_sinkhorn(xs, xt, reg=0.1))    # _empirical.py:159-167


def config1():
    """BASELINE config 1: n=m=1000 2-D Gaussian clouds, sqeuclidean, reg=0.05, 1000 fixed its, fp64."""
    rng = np.random.RandomState(0)
    x = rng.randn(1000, 2)
    y = rng.randn(1000, 2) + 1
    M = ot.dist(x, y)
    a = ot.unif(1000)
    b = ot.unif(1000)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        P, lg = ot.sinkhorn(a, b, M, 0.05, numItermax=1000, stopThr=0, log=True)
    np.savez_compressed(os.path.join(HERE, "golden_c1.npz"), seed=0, n=1000, m=1000, reg=0.05, its=1000,
                        u=lg["u"], v=lg["v"], err=np.asarray(lg["err"]), niter=lg["niter"],
                        plan_row0=P[0], plan_col0=P[:, 0], plan_sum=P.sum(), loss=np.sum(P * M),
                        M_checksum=M.sum(), M_max=M.max())


if __name__ == "__main__":
    assert "potref" in ot.__file__ or "reference" in ot.__file__, ot.__file__
    dense_cases()
    empirical_cases()
    batch_cases()
    doctest_vectors()
    config1()
    np.savez_compressed(os.path.join(HERE, "golden_small.npz"), **G)
    print("wrote", len(G), "arrays;", ot.__version__, "numpy", np.__version__, file=sys.stderr)
