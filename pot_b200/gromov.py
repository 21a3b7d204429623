"""Entropic Gromov-Wasserstein on the device Sinkhorn solver.

Mirrors (same names, argument meaning, return values, warnings and errors):

    ot.gromov.entropic_gromov_wasserstein    ot/gromov/_bregman.py:25-252
    ot.gromov.entropic_gromov_wasserstein2   ot/gromov/_bregman.py:255-413

Every outer iteration (projected gradient or proximal point, :197-244) is a gradient of the GW energy followed by one
entropic OT solve on it.  The solve is `pot_b200.bregman.sinkhorn_knopp` (Gibbs kernel, fused sweep, device-side
stopping, plan kernel: libb2ot.so); it is the repeated Sinkhorn consumer this row is about.  The gradient
`constC - hC1 T hC2^T` (ot/gromov/_utils.py:244-280) runs on the library's own strided matrix-product kernel
(`b2ot_strided_gemm`, pot_b200/_gemm.py: two launches, the second fuses the `constC -`); the rest (`init_matrix` - matrix
times vector and outer products -, the `- epsilon log T` term of PPA, the error every 10th outer iteration) is n x m
elementwise glue on the device.  Inputs must be torch CUDA tensors; there is no CPU path.
"""

import math
import warnings

import torch

from . import _gemm
from .bregman import _require_cuda_matrix, _weights, sinkhorn

__all__ = ["entropic_gromov_wasserstein", "entropic_gromov_wasserstein2"]


def _transform_matrix(C1, C2, loss_fun):
    """ot/gromov/_utils.py:38-150: L(a, b) = f1(a) + f2(b) - h1(a) h2(b)."""
    if loss_fun == "square_loss":
        return C1 ** 2, C2 ** 2, C1, 2 * C2
    return C1 * torch.log(C1 + 1e-18) - C1, C2, C1, torch.log(C2 + 1e-18)  # kl_loss


def _init_matrix(C1, C2, p, q, loss_fun):
    """ot/gromov/_utils.py:153-241."""
    fC1, fC2, hC1, hC2 = _transform_matrix(C1, C2, loss_fun)
    constC1 = torch.mm(torch.mm(fC1, p.reshape(-1, 1)), torch.ones((1, q.shape[0]), dtype=q.dtype, device=q.device))
    constC2 = torch.mm(torch.ones((p.shape[0], 1), dtype=p.dtype, device=p.device),
                       torch.mm(q.reshape(1, -1), fC2.t()))
    return constC1 + constC2, hC1, hC2


def _tensor_product(constC, hC1, hC2, T):
    """constC - hC1 T hC2^T (ot/gromov/_utils.py:244-280) on the library's own strided matrix-product kernel
    (b2ot_strided_gemm, two launches; the second one fuses the `constC -`)."""
    return _gemm.contract(constC, hC1, T, hC2)


def entropic_gromov_wasserstein(C1, C2, p=None, q=None, loss_fun="square_loss", epsilon=0.1, symmetric=None, G0=None,
                                max_iter=1000, tol=1e-9, solver="PGD", warmstart=False, verbose=False, log=False,
                                **kwargs):
    """Drop-in for ot.gromov.entropic_gromov_wasserstein (ot/gromov/_bregman.py:25-252).

    Returns the coupling T; with log=True also {"err", "gw_dist"}.  `**kwargs` go to the inner Sinkhorn solve
    (numItermax, stopThr, ... of ot.sinkhorn), like the reference.
    """
    if solver not in ["PGD", "PPA"]:
        raise ValueError("Unknown solver '%s'. Pick one in ['PGD', 'PPA']." % solver)  # :149-150
    if loss_fun not in ("square_loss", "kl_loss"):
        raise ValueError(f"Unknown `loss_fun='{loss_fun}'`. Use one of: {'square_loss', 'kl_loss'}.")  # :152-155
    C1 = _require_cuda_matrix(C1, "C1").detach()
    C2 = _require_cuda_matrix(C2, "C2").detach().to(C1.dtype)
    if C2.device != C1.device:
        raise ValueError(f"pot_b200: `C2` is on {C2.device} but C1 is on {C1.device}")
    N1, N2 = C1.shape[0], C2.shape[0]
    p = _weights(p if p is not None else [], N1, C1, "p")  # unif(N1) when None (:160-165)
    q = _weights(q if q is not None else [], N2, C1, "q")
    with torch.cuda.device(C1.device):
        T = torch.outer(p, q) if G0 is None else _require_cuda_matrix(G0, "G0").detach().to(C1.dtype)  # :172-175
        constC, hC1, hC2 = _init_matrix(C1, C2, p, q, loss_fun)
        if symmetric is None:  # :178-181
            symmetric = bool(torch.allclose(C1, C1.t(), atol=1e-10)) and bool(torch.allclose(C2, C2.t(), atol=1e-10))
        if not symmetric:
            constCt, hC1t, hC2t = _init_matrix(C1.t(), C2.t(), p, q, loss_fun)
        cpt = 0
        err = 1.0
        if warmstart:  # potentials matching ot.sinkhorn's initial scalings 1/N (:188-192)
            mu = torch.zeros(N1, dtype=C1.dtype, device=C1.device) - math.log(N1)
            nu = torch.zeros(N2, dtype=C1.dtype, device=C1.device) - math.log(N2)
        errs = []
        while err > tol and cpt < max_iter:
            Tprev = T
            if symmetric:
                tens = 2 * _tensor_product(constC, hC1, hC2, T)  # gwggrad, _utils.py:323-354
            else:
                tens = 0.5 * (2 * _tensor_product(constC, hC1, hC2, T) + 2 * _tensor_product(constCt, hC1t, hC2t, T))
            if solver == "PPA":
                tens = tens - epsilon * torch.log(T)
            if warmstart:
                T, loginn = sinkhorn(p, q, tens, epsilon, method="sinkhorn", log=True, warmstart=(mu, nu), **kwargs)
                mu = epsilon * torch.log(loginn["u"])
                nu = epsilon * torch.log(loginn["v"])
            else:
                T = sinkhorn(p, q, tens, epsilon, method="sinkhorn", **kwargs)
            if cpt % 10 == 0:  # :232-243: the only host read of the outer loop
                e = torch.linalg.norm(T - Tprev)
                err = float(e)
                errs.append(e)
                if verbose:
                    if cpt % 200 == 0:
                        print("{:5s}|{:12s}".format("It.", "Err") + "\n" + "-" * 19)
                    print("{:5d}|{:8e}|".format(cpt, err))
            cpt += 1
        if abs(float(torch.sum(T)) - 1) > 1e-5:  # :246-250
            warnings.warn("Solver failed to produce a transport plan. You might "
                          "want to increase the regularization parameter `epsilon`.")
        if log:
            return T, {"err": errs, "gw_dist": torch.sum(_tensor_product(constC, hC1, hC2, T) * T)}  # gwloss
    return T


def entropic_gromov_wasserstein2(C1, C2, p=None, q=None, loss_fun="square_loss", epsilon=0.1, symmetric=None, G0=None,
                                 max_iter=1000, tol=1e-9, solver="PGD", warmstart=False, verbose=False, log=False,
                                 **kwargs):
    """Drop-in for ot.gromov.entropic_gromov_wasserstein2 (ot/gromov/_bregman.py:255-413): the GW loss of the entropic
    coupling; with log=True also the dict of the solve plus "T"."""
    T, logv = entropic_gromov_wasserstein(C1, C2, p, q, loss_fun, epsilon, symmetric, G0, max_iter, tol, solver,
                                          warmstart, verbose, log=True, **kwargs)
    logv["T"] = T
    if log:
        return logv["gw_dist"], logv
    return logv["gw_dist"]
