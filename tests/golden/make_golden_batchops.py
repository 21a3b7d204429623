"""Golden vectors of the UNMODIFIED reference for the batched helpers of ot.batch (dist_batch, loss_linear_batch,
loss_linear_samples_batch, entropy_batch, solve_sample_batch).

    python tests/golden/make_golden_batchops.py        # needs oracle/_ref (oracle/build_ref.py)

  golden_batchops.npz, per case `{dtype}_{metric}`: X, Y, p, M = dist_batch(X, Y, metric, p), T (a random plan),
  loss = loss_linear_batch(M, T), loss_samples = loss_linear_samples_batch(X, Y, T, metric), ent = entropy_batch(T),
  and solve_sample_batch(X, Y, reg=0.1, metric, max_iter=20, tol=0): ssb_value, ssb_value_linear, ssb_plan.
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle.build_ref import import_ref  # noqa: E402

ot = import_ref()
import ot.batch as ob  # noqa: E402

warnings.simplefilter("ignore")
G = {}
for dt, dn in ((np.float64, "f64"), (np.float32, "f32")):
    for metric, p in (("sqeuclidean", 2), ("euclidean", 2), ("minkowski", 3), ("minkowski", 1), ("kl", 2)):
        rng = np.random.RandomState(5)
        B, n, m, d = 2, 37, 26, 5
        X, Y = rng.rand(B, n, d).astype(dt), rng.rand(B, m, d).astype(dt) + dt(0.3)
        if metric == "kl":
            X, Y = X / X.sum(-1, keepdims=True), Y / Y.sum(-1, keepdims=True)
        c = f"{dn}_{metric}" + ("_p1" if p == 1 else "")
        M = ob.dist_batch(X, Y, metric=metric, p=p)
        T = rng.rand(B, n, m).astype(dt)
        T /= T.sum(axis=(1, 2), keepdims=True)
        G[f"{c}/X"], G[f"{c}/Y"], G[f"{c}/p"], G[f"{c}/M"], G[f"{c}/T"] = X, Y, np.asarray(p), M, T
        G[f"{c}/loss"] = ob.loss_linear_batch(M, T)
        G[f"{c}/ent"] = ob.entropy_batch(T)
        if p == 2:
            G[f"{c}/loss_samples"] = ob.loss_linear_samples_batch(X, Y, T, metric=metric)
        res = ob.solve_sample_batch(X, Y, 0.1, metric=metric, p=p, max_iter=20, tol=0)
        G[f"{c}/ssb_value"], G[f"{c}/ssb_value_linear"] = np.asarray(res.value), np.asarray(res.value_linear)
        G[f"{c}/ssb_plan"] = np.asarray(res.plan)
        assert M.dtype == dt
out = os.path.join(HERE, "golden_batchops.npz")
np.savez_compressed(out, **G)
print(len(G), "arrays ->", out, os.path.getsize(out), "bytes")
