"""The Gromov-Wasserstein contraction on the library's own strided matrix-product kernel (b2ot_strided_gemm).

    contract(const, hC1, T, hC2)        const - hC1 T hC2^T                      ot/gromov/_utils.py:244-280
    contract_batch(const, hC1, T, hC2)  const - sum_d hC1[..., d] T hC2[..., d]^T  ot/batch/_quadratic.py:535-572

The kernels are forward-only: when an operand requires grad the callers keep the reference's torch expression (autograd),
see batch_quadratic.tensor_product_batch.
"""
import torch

from . import _lib
from ._dense import dtype_code, stream_ptr


def _gemm(Zb, Zf, Sf, M, N, K, A, sA, B, sB, C0, sC0, Cout, sC, alpha, beta):
    lib = _lib.load()
    with torch.cuda.device(Cout.device):
        _lib.check(lib.b2ot_strided_gemm(dtype_code(Cout.dtype), Zb, Zf, Sf, M, N, K, A.data_ptr(), *sA, B.data_ptr(), *sB,
                                         C0.data_ptr() if C0 is not None else None, *sC0, Cout.data_ptr(), *sC,
                                         float(alpha), float(beta), stream_ptr(Cout.device)), "b2ot_strided_gemm")


def contract(const, hC1, T, hC2, transpose=False, alpha=-1.0, out=None):
    """out = const + alpha * op(hC1) T op(hC2)^T for (n, n), (n, m), (m, m) operands; op = transpose when asked."""
    n, m = T.shape
    dt, dev = T.dtype, T.device
    hC1, hC2, T = hC1.to(dt), hC2.to(dt), T.contiguous()
    P = torch.empty((n, m), dtype=dt, device=dev)
    s1 = (hC1.stride(1), hC1.stride(0)) if transpose else (hC1.stride(0), hC1.stride(1))
    _gemm(1, 1, 1, n, m, n, hC1, (0, 0) + s1, T, (0, 0, m, 1), None, (0, 0), P, (0, 0, m), 1.0, 0.0)
    if out is None:
        out = torch.empty((n, m), dtype=dt, device=dev)
    const = const.to(dt)
    if const.stride(-1) != 1:
        const = const.contiguous()
    # B(k, j) = op(hC2)(j, k)
    s2 = (hC2.stride(0), hC2.stride(1)) if transpose else (hC2.stride(1), hC2.stride(0))
    _gemm(1, 1, 1, n, m, m, P, (0, 0, m, 1), hC2, (0, 0) + s2, const, (0, const.stride(0)), out, (0, 0, m), alpha, 1.0)
    return out


def contract_batch(const, hC1, T, hC2, symmetric=True):
    """const - sum_d hC1[b, :, :, d] T[b] hC2[b, :, :, d]^T, averaged with the transposed structures when not symmetric.
    hC1 (B, n, n, D), hC2 (B, m, m, D), T (B, n, m), const (B, n, m)."""
    Bn, n, m = T.shape
    D = hC1.shape[3]
    dt, dev = T.dtype, T.device
    hC1, hC2, T, const = hC1.to(dt).contiguous(), hC2.to(dt).contiguous(), T.contiguous(), const.to(dt).contiguous()
    P = torch.empty((Bn, D, n, m), dtype=dt, device=dev)
    out = torch.empty((Bn, n, m), dtype=dt, device=dev)
    passes = ((False, -1.0),) if symmetric else ((False, -0.5), (True, -0.5))
    src = const
    for tr, alpha in passes:
        a_i, a_k = (D, n * D) if tr else (n * D, D)           # op(hC1)[b, :, :, d](i, k)
        _gemm(Bn, D, 1, n, m, n, hC1, (n * n * D, 1, a_i, a_k), T, (n * m, 0, m, 1), None, (0, 0), P,
              (D * n * m, n * m, m), 1.0, 0.0)
        b_k, b_j = (m * D, D) if tr else (D, m * D)           # B(k, j) = op(hC2)[b, :, :, d](j, k)
        _gemm(Bn, 1, D, n, m, m, P, (D * n * m, n * m, m, 1), hC2, (m * m * D, 1, b_k, b_j), src, (n * m, m), out,
              (n * m, 0, m), alpha, 1.0)
        src = out
    return out
