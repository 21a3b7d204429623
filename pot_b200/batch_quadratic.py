"""Batched (fused) Gromov-Wasserstein by proximal gradient, every inner projection solved on chip by libb2ot.so.

Mirrors (same names, argument meaning, return values and errors):

    ot.batch.solve_gromov_batch             ot/batch/_quadratic.py:269-498
    ot.batch.tensor_batch                   ot/batch/_quadratic.py:17-152
    ot.batch.loss_quadratic_batch           ot/batch/_quadratic.py:155-198
    ot.batch.loss_quadratic_samples_batch   ot/batch/_quadratic.py:201-266

The outer loop (:456-470) is: cost of the current plan `Mk = (1 - alpha) M + 2 alpha L x T`, then a log-domain Bregman
projection of `K = -Mk/reg + log T` onto the marginals.  The projection is `pot_b200.batch.bregman_log_projection_batch`
(one thread-block cluster per problem, the whole inner solve in shared memory); the tensor product `L x T` (`einsum`, :551-562) runs on the
library's own strided, batched matrix-product kernel (`b2ot_strided_gemm`, pot_b200/_gemm.py: the feature axis and the
transposes are strides, the `const -` and the feature sum are fused); only the one differentiable evaluation of the
envelope value goes through torch autograd.
Inputs must be torch CUDA tensors; there is no CPU path.
"""

import warnings

import torch

from . import _gemm
from .batch import _check3, _weights, bregman_log_projection_batch
from .utils import OTResult

__all__ = ["solve_gromov_batch", "tensor_batch", "tensor_product_batch", "loss_quadratic_batch",
           "loss_quadratic_samples_batch"]


def _transpose(C):
    return C.transpose(1, 2)  # (0, 2, 1) or (0, 2, 1, 3), :575-578


def _const_from_marginals(fC1, fC2, a, b):
    """f1(C1) a 1^T + 1 b^T f2(C2)^T (:581-590)."""
    return torch.einsum("bij,bj->bi", fC1, a)[:, :, None] + torch.einsum("bij,bj->bi", fC2, b)[:, None, :]


def tensor_batch(a, b, C1, C2, symmetric=True, nx=None, loss="sqeuclidean", logits=None):
    """Cost tensor L of the decomposable loss l(x, y) = f1(x) + f2(y) - <h1(x), h2(y)> as the dict
    {"constC", "hC1", "hC2", "fC1", "fC2"} (ot/batch/_quadratic.py:17-152, :504-532)."""
    if loss == "sqeuclidean":
        fC1 = torch.sum(C1 ** 2, dim=-1) if C1.dim() == 4 else C1 ** 2
        fC2 = torch.sum(C2 ** 2, dim=-1) if C2.dim() == 4 else C2 ** 2
        hC1 = 2 * (C1.unsqueeze(-1) if C1.dim() == 3 else C1)
        hC2 = C2.unsqueeze(-1) if C2.dim() == 3 else C2
    elif loss == "kl":
        assert logits in [True, False], "logits must be either True or False for KL loss"
        assert C2.dim() == 4, "C2 must be a bxnxnxd tensor"
        fC1 = torch.zeros((C1.shape[0], C1.shape[1], C1.shape[2]), dtype=C1.dtype, device=C1.device)
        fC2 = torch.sum(C2 * torch.log(C2 + 1e-15), dim=-1)
        hC1 = C1 if logits else torch.log(C1 + 1e-15)
        hC2 = C2
    else:
        raise ValueError(f"Unknown loss function: {loss}")
    if not symmetric:
        fC1 = 0.5 * (fC1 + _transpose(fC1))
        fC2 = 0.5 * (fC2 + _transpose(fC2))
    return {"constC": _const_from_marginals(fC1, fC2, a, b), "hC1": hC1, "hC2": hC2, "fC1": fC1, "fC2": fC2}


def tensor_product_batch(L, T, nx=None, recompute_const=False, symmetric=True):
    """L x T = const - hC1 T hC2^T (summed over the feature axis), :535-572."""
    if recompute_const:
        const = _const_from_marginals(L["fC1"], L["fC2"], torch.sum(T, dim=2), torch.sum(T, dim=1))
    else:
        const = L["constC"]
    hC1, hC2 = L["hC1"], L["hC2"]
    if not (torch.is_grad_enabled() and any(t.requires_grad for t in (hC1, hC2, T, const))):
        # the solver's loop: the library's own strided matrix-product kernel (forward-only; b2ot_strided_gemm)
        return _gemm.contract_batch(const, hC1, T, hC2, symmetric=symmetric)
    # differentiable evaluation (the envelope value of solve_gromov_batch, once per solve): torch autograd
    dot = torch.einsum("bijd,bjk->bikd", hC1, T)
    dot = torch.einsum("bikd,bjkd->bijd", dot, hC2)
    dot = torch.sum(dot, dim=-1)
    if not symmetric:
        dot_t = torch.einsum("bijd,bjk->bikd", _transpose(hC1), T)
        dot_t = torch.einsum("bikd,bjkd->bijd", dot_t, _transpose(hC2))
        dot_t = torch.sum(dot_t, dim=-1)
        dot = (dot + dot_t) / 2
    return const - dot


def loss_quadratic_batch(L, T, recompute_const=False, symmetric=True, nx=None):
    """<L x T, T> per problem (:155-198)."""
    LT = tensor_product_batch(L, T, recompute_const=recompute_const, symmetric=symmetric)
    return torch.sum(LT * T, dim=(1, 2))


def loss_quadratic_samples_batch(a, b, C1, C2, T, loss="sqeuclidean", symmetric=None, nx=None, logits=None,
                                 recompute_const=False):
    """Quadratic loss straight from the samples' structure matrices (:201-266)."""
    if not isinstance(loss, str):
        raise ValueError(f"Unknown loss function: {loss}")
    L = tensor_batch(a, b, C1, C2, symmetric=symmetric, loss=loss, logits=logits)
    return loss_quadratic_batch(L, T, recompute_const=recompute_const, symmetric=symmetric)


def _structure(C, name):
    if not isinstance(C, torch.Tensor):
        raise TypeError(f"pot_b200: `{name}` must be a torch CUDA tensor (got {type(C).__name__}); no numpy/CPU backend")
    if not C.is_cuda:
        raise ValueError(f"pot_b200: `{name}` must live on a CUDA device (got {C.device}); there is no CPU fallback")
    if C.dim() not in (3, 4) or C.dtype not in (torch.float32, torch.float64):
        raise ValueError(f"pot_b200: `{name}` must be a float32/float64 (B, n, n) or (B, n, n, d) tensor")
    return C


def solve_gromov_batch(C1, C2, reg=1e-2, a=None, b=None, loss="sqeuclidean", symmetric=None, M=None, alpha=None,
                       T_init=None, max_iter=50, tol=1e-5, max_iter_inner=50, tol_inner=1e-5, grad="envelope",
                       logits=None):
    """Drop-in for ot.batch.solve_gromov_batch (ot/batch/_quadratic.py:269-498).

    Returns an OTResult with (B, n, m) `plan`, (B,) `value`, `value_linear` (weighted by 1 - alpha), `value_quad`
    (weighted by alpha), `potentials` of the last inner projection and log {"n_iter"}.
    grad: "detach" and "envelope" as the reference; "autodiff" (differentiation through the unrolled inner loops) is
    replaced by "envelope" with a warning: the kernels are forward-only.
    """
    C1, C2 = _structure(C1, "C1"), _structure(C2, "C2")
    Bn, n, m = C1.shape[0], C1.shape[1], C2.shape[1]
    a = _weights(a, Bn, n, C1)
    b = _weights(b, Bn, m, C1)
    if grad == "autodiff":
        warnings.warn("pot_b200: grad='autodiff' is replaced by grad='envelope' (forward-only kernels)")
        grad = "envelope"
    with torch.cuda.device(C1.device):
        if symmetric is None:  # :410-413
            symmetric = bool(torch.allclose(C1, _transpose(C1), atol=1e-10)) and \
                bool(torch.allclose(C2, _transpose(C2), atol=1e-10))
        if not isinstance(loss, str):
            raise ValueError(f"Unknown loss function: {loss}")
        L = tensor_batch(a, b, C1, C2, symmetric=symmetric, loss=loss, logits=logits)
        if M is None and alpha is None:  # :426-446
            M = torch.zeros((Bn, n, m), dtype=C1.dtype, device=C1.device)
            alpha = 1.0
        elif M is not None and alpha is None:
            raise ValueError("If M is provided, alpha must also be provided for Fused Gromov-Wasserstein problem")
        elif M is None and alpha is not None:
            raise ValueError("If alpha is provided, M must also be provided for Fused Gromov-Wasserstein problem")
        else:
            M = _check3(M, "M")
            if M.shape[0] != Bn or M.shape[1] != n or M.shape[2] != m:
                raise ValueError(f"Shape of M {M.shape} does not match the batch size {Bn} and dimensions {n}, {m}")
            if alpha < 0 or alpha > 1:
                raise ValueError("Alpha must be in [0, 1] for Fused Gromov-Wasserstein problem")
        if T_init is None:  # :450-452
            const = torch.sqrt(torch.sum(a, dim=1) * torch.sum(b, dim=1))
            T_init = torch.einsum("bi,bj->bij", a, b) / const[:, None, None]
        T = T_init.detach()
        log_T = torch.log(T + 1e-15)
        Ld = {k: v.detach() for k, v in L.items()}
        Md = M.detach()
        it = 0
        out = None
        for it in range(max_iter):  # :458-470
            T_prev = T
            Mk = (1 - alpha) * Md + 2 * alpha * tensor_product_batch(Ld, T, recompute_const=False, symmetric=symmetric)
            K = -Mk / reg + log_T
            out = bregman_log_projection_batch(K, a, b, max_iter=max_iter_inner, tol=tol_inner, grad="detach")
            T, log_T = out["T"], out["log_T"]
            max_err = torch.max(torch.mean(torch.abs(T_prev - T), dim=(1, 2)))
            if float(max_err) < tol:  # the outer loop's host read, once per proximal step
                break
        if grad == "detach":
            M, L = Md, Ld
        value_linear = torch.sum(M * T, dim=(1, 2))  # loss_linear_batch, ot/batch/_linear.py:123-150
        value_quadratic = loss_quadratic_batch(L, T, recompute_const=True, symmetric=symmetric)
        value = (1 - alpha) * value_linear + alpha * value_quadratic
    potentials = out["potentials"] if out is not None else None
    return OTResult(value=value, value_linear=value_linear * (1 - alpha), value_quad=value_quadratic * alpha, plan=T,
                    potentials=potentials, log={"n_iter": it})
