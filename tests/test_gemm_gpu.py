"""b2ot_strided_gemm (the Gromov-Wasserstein contraction, pot_b200/_gemm.py) against torch in fp64.

Reference expressions: `constC - hC1 T hC2^T` (ot/gromov/_utils.py:244-280) and the einsums of tensor_product_batch
(ot/batch/_quadratic.py:551-566)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from pot_b200 import _gemm

DEV = "cuda"


def _rand(*shape, dtype, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return torch.randn(*shape, device=DEV, dtype=dtype, generator=g)


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-12), (torch.float32, 2e-5)])
@pytest.mark.parametrize("n,m", [(1, 1), (5, 3), (64, 64), (65, 130), (200, 77), (513, 300)])
def test_contract_dense(n, m, dtype, tol):
    hC1, hC2 = _rand(n, n, dtype=dtype, seed=1), _rand(m, m, dtype=dtype, seed=2)
    T, const = _rand(n, m, dtype=dtype, seed=3), _rand(n, m, dtype=dtype, seed=4)
    ref = const.double() - hC1.double() @ T.double() @ hC2.double().t()
    scale = float(ref.abs().max()) + 1.0
    out = _gemm.contract(const, hC1, T, hC2)
    assert out.dtype == dtype and float((out.double() - ref).abs().max()) <= tol * scale * max(n, m)
    # transposed structures (the asymmetric branch of the solver) and non-contiguous views
    out_t = _gemm.contract(const, hC1, T, hC2, transpose=True)
    ref_t = const.double() - hC1.double().t() @ T.double() @ hC2.double()
    assert float((out_t.double() - ref_t).abs().max()) <= tol * scale * max(n, m)
    out_v = _gemm.contract(const, hC1.t(), T, hC2.t())
    assert float((out_v.double() - ref_t).abs().max()) <= tol * scale * max(n, m)


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-12), (torch.float32, 2e-5)])
@pytest.mark.parametrize("B,n,m,D", [(1, 4, 6, 1), (3, 70, 33, 1), (2, 65, 65, 3), (5, 20, 129, 4)])
@pytest.mark.parametrize("symmetric", [True, False])
def test_contract_batch(B, n, m, D, symmetric, dtype, tol):
    hC1, hC2 = _rand(B, n, n, D, dtype=dtype, seed=5), _rand(B, m, m, D, dtype=dtype, seed=6)
    T, const = _rand(B, n, m, dtype=dtype, seed=7), _rand(B, n, m, dtype=dtype, seed=8)
    h1, h2, Td = hC1.double(), hC2.double(), T.double()
    dot = torch.einsum("bikd,bjkd->bijd", torch.einsum("bijd,bjk->bikd", h1, Td), h2).sum(-1)
    if not symmetric:
        dot_t = torch.einsum("bikd,bjkd->bijd", torch.einsum("bijd,bjk->bikd", h1.transpose(1, 2), Td),
                             h2.transpose(1, 2)).sum(-1)
        dot = (dot + dot_t) / 2
    ref = const.double() - dot
    out = _gemm.contract_batch(const, hC1, T, hC2, symmetric=symmetric)
    assert out.shape == (B, n, m) and out.dtype == dtype
    scale = float(ref.abs().max()) + 1.0
    assert float((out.double() - ref).abs().max()) <= tol * scale * max(n, m) * D


def test_contract_is_deterministic():
    hC1, hC2 = _rand(300, 300, dtype=torch.float32, seed=1), _rand(200, 200, dtype=torch.float32, seed=2)
    T, const = _rand(300, 200, dtype=torch.float32, seed=3), _rand(300, 200, dtype=torch.float32, seed=4)
    a = _gemm.contract(const, hC1, T, hC2)
    b = _gemm.contract(const, hC1, T, hC2)
    assert torch.equal(a, b)
