"""Golden gradients from the UNMODIFIED reference (torch backend, CPU) for the loss-returning entry points.

Run in the build container (needs oracle/_ref, built from /root/reference by oracle/build_ref.py):

    python tests/golden/make_golden_gradients.py

  golden_gradients.npz, per case `pc_{dtype}_{metric}[_same]`:
    X, Y, a, b, reg, its
    dense/gX, gY, ga, gb   : X, Y -> M = ot.dist(X, Y, metric) -> ot.solve(M, a, b, reg, max_iter=its, tol=0,
                             grad="envelope").value.backward()  (the envelope gradient of ot/solvers.py:485-494 chained
                             through the reference's differentiable dist)
    dense/unrolled_gM      : d sinkhorn2 / dM by the reference's own unrolled autodiff (ot.sinkhorn2 on torch tensors,
                             the path test/test_bregman.py:203-225 pins), informative: the two agree up to the entropy
                             term's gradient; the test only reports the distance
    lazy/f, g              : ot.bregman.empirical_sinkhorn(isLazy=True, numIterMax=its, stopThr=-1) on numpy
    lazy/gX, gY, ga, gb    : envelope gradient built from the reference's (f, g): autograd of
                             sum_ij C(X, Y)_ij P_ij with P = exp(f_i + g_j - C_ij / reg) held fixed (C = ot.dist on torch),
                             and reg (f - mean f), reg (g - mean g)
"""
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle.build_ref import import_ref  # noqa: E402

ot = import_ref()
warnings.simplefilter("ignore")
G = {}


def case(name, dtype, metric, same, n=40, m=30, d=3, reg=0.5, its=200, seed=0):
    rng = np.random.RandomState(seed)
    X = rng.randn(n, d).astype(dtype)
    Y = X if same else (rng.randn(m, d) + 0.5).astype(dtype)
    m = Y.shape[0]
    a = rng.rand(n).astype(dtype) + 0.5
    a /= a.sum()
    b = rng.rand(m).astype(dtype) + 0.5
    b /= b.sum()
    G[f"{name}/X"], G[f"{name}/Y"], G[f"{name}/a"], G[f"{name}/b"] = X, Y, a, b
    G[f"{name}/reg"], G[f"{name}/its"] = np.asarray(reg), np.asarray(its)

    # dense: envelope through the reference's differentiable dist
    Xt = torch.tensor(X, requires_grad=True)
    Yt = Xt if same else torch.tensor(Y, requires_grad=True)
    at, bt = torch.tensor(a, requires_grad=True), torch.tensor(b, requires_grad=True)
    M = ot.dist(Xt, Yt, metric=metric)
    res = ot.solve(M, at, bt, reg=reg, max_iter=its, tol=0, grad="envelope")
    res.value.backward()
    G[f"{name}/dense/gX"] = Xt.grad.numpy()
    if not same:
        G[f"{name}/dense/gY"] = Yt.grad.numpy()
    G[f"{name}/dense/ga"], G[f"{name}/dense/gb"] = at.grad.numpy(), bt.grad.numpy()

    if n <= 64:
        Mt = torch.tensor(ot.dist(X, Y, metric=metric), requires_grad=True)
        val = ot.sinkhorn2(torch.tensor(a), torch.tensor(b), Mt, reg, method="sinkhorn_log", numItermax=its, stopThr=0)
        val.backward()
        G[f"{name}/dense/unrolled_gM"] = Mt.grad.numpy()
        G[f"{name}/dense/plan"] = res.plan.detach().numpy()

    # lazy: the reference's potentials, envelope gradient by autograd with the plan held fixed
    f, g = ot.bregman.empirical_sinkhorn(X, Y, reg, a, b, metric=metric, numIterMax=its, stopThr=-1.0, isLazy=True,
                                         batchSize=16, warn=False)
    G[f"{name}/lazy/f"], G[f"{name}/lazy/g"] = np.asarray(f), np.asarray(g)
    Xt = torch.tensor(X.astype(np.float64), requires_grad=True)
    Yt = Xt if same else torch.tensor(Y.astype(np.float64), requires_grad=True)
    ft, gt = torch.tensor(np.asarray(f, dtype=np.float64)), torch.tensor(np.asarray(g, dtype=np.float64))
    # no X-is-Y diagonal zeroing here: the lazy loop of the reference calls dist on row blocks (never `X is Y`)
    Yd = Yt if not same else Xt + 0.0
    C = ot.dist(Xt, Yd, metric=metric)
    P = torch.exp(ft[:, None] + gt[None, :] - C.detach() / reg)
    if metric == "euclidean":  # sqrt(0) has no derivative: the zero-cost entries contribute nothing
        C2 = ot.dist(Xt, Yd, metric="sqeuclidean")
        Csafe = torch.sqrt(torch.where(C2 > 0, C2, torch.ones_like(C2)))
        loss = (torch.where(C2 > 0, Csafe, torch.zeros_like(C2)) * P).sum()
    else:
        loss = (C * P).sum()
    loss.backward()
    G[f"{name}/lazy/gX"] = Xt.grad.numpy()
    if not same:
        G[f"{name}/lazy/gY"] = Yt.grad.numpy()
    G[f"{name}/lazy/ga"] = (reg * (ft - ft.mean())).numpy()
    G[f"{name}/lazy/gb"] = (reg * (gt - gt.mean())).numpy()
    G[f"{name}/lazy/loss"] = np.asarray(float((C.detach() * P).sum()))


for dt, dn in ((np.float64, "f64"), (np.float32, "f32")):
    for metric in ("sqeuclidean", "euclidean"):
        case(f"pc_{dn}_{metric}", dt, metric, same=False)
    case(f"pc_{dn}_sqeuclidean_same", dt, "sqeuclidean", same=True)
    case(f"pc_{dn}_sqeuclidean_big", dt, "sqeuclidean", same=False, n=300, m=170, d=16, reg=2.0, its=50, seed=3)

out = os.path.join(HERE, "golden_gradients.npz")
np.savez_compressed(out, **G)
print(len(G), "arrays ->", out, os.path.getsize(out), "bytes")
