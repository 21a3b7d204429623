"""b2ot_colsum_static: the column marginal without any allocation (scratch in a file-scope device array).

The entry point has no counterpart in the reference (include/b2ot.h says why it exists); what it computes is
nx.sum(plan, 0) (ot/unbalanced/_sinkhorn.py:795) like b2ot_colsum, with the same kernels and summation order, so the
two entry points must agree to the bit.  The file name keeps this module last in pytest's collection order.
"""
import ctypes as C

import pytest
import torch

from pot_b200 import _lib
from pot_b200._dense import dtype_code, stream_ptr

pytestmark = pytest.mark.gpu

MAX_COLS = 16384  # B2OT_COLSUM_STATIC_MAX_COLS (include/b2ot.h)
TOL = {torch.float32: dict(rtol=1e-5, atol=1e-6), torch.float64: dict(rtol=1e-12, atol=1e-14)}


def _colsum_scratch(P):
    lib = _lib.load()
    n, m = P.shape
    out = torch.empty(m, dtype=P.dtype, device=P.device)
    scratch = torch.empty(lib.b2ot_colsum_scratch_bytes(n, m), dtype=torch.uint8, device=P.device)
    with torch.cuda.device(P.device):
        _lib.check(lib.b2ot_colsum(dtype_code(P.dtype), P.data_ptr(), n, m, P.stride(0), out.data_ptr(),
                                   scratch.data_ptr(), stream_ptr(P.device)), "b2ot_colsum")
    return out


def _colsum_static(P, out):
    lib = _lib.load()
    n, m = P.shape
    with torch.cuda.device(P.device):
        return lib.b2ot_colsum_static(dtype_code(P.dtype), P.data_ptr(), n, m, P.stride(0), out.data_ptr(),
                                      stream_ptr(P.device))


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_static_scratch_matches_existing_entry_point_twice(dtype):
    dev = torch.device("cuda:0")
    gen = torch.Generator(device="cpu").manual_seed(1234)
    # two calls in a row on the shared array: the second is smaller (stale partials of the first lie beside its own)
    # and ragged (m not a multiple of the block, fewer rows than row blocks); the third is larger again
    shapes = [(777, 5000), (19, 333), (2048, MAX_COLS)]
    for n, m in shapes:
        P = torch.rand(n, m, generator=gen, dtype=torch.float64).to(dtype).to(dev)
        want = _colsum_scratch(P)
        out = torch.full((m,), float("nan"), dtype=dtype, device=dev)
        before = torch.cuda.memory_allocated(dev)
        rc = _colsum_static(P, out)
        assert rc == 0, _lib.load().b2ot_last_error()
        assert torch.cuda.memory_allocated(dev) == before
        torch.cuda.synchronize(dev)
        torch.testing.assert_close(out, want, **TOL[dtype])
        assert torch.equal(out, want)  # same kernels, same order
        ref = P.double().sum(0)
        torch.testing.assert_close(out.double(), ref, rtol=1e-5 if dtype == torch.float32 else 1e-12, atol=0)


def test_static_scratch_strided_rows():
    dev = torch.device("cuda:0")
    base = torch.rand(300, 1024, dtype=torch.float64, device=dev)
    P = base[:, :700]  # ld = 1024 > m
    out = torch.empty(700, dtype=torch.float64, device=dev)
    assert _colsum_static(P, out) == 0
    assert torch.equal(out, _colsum_scratch(P))


def test_static_scratch_rejects_beyond_the_maximum():
    dev = torch.device("cuda:0")
    lib = _lib.load()
    P = torch.ones(2, MAX_COLS + 1, dtype=torch.float32, device=dev)
    out = torch.zeros(MAX_COLS + 1, dtype=torch.float32, device=dev)
    rc = _colsum_static(P, out)
    assert rc == -1
    msg = lib.b2ot_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "exceeds the static scratch" in msg and str(MAX_COLS) in msg
    torch.cuda.synchronize(dev)
    assert float(out.abs().sum()) == 0.0  # nothing was launched
    # the next valid call is unaffected
    Q = torch.rand(64, 100, dtype=torch.float32, device=dev)
    o2 = torch.empty(100, dtype=torch.float32, device=dev)
    assert _colsum_static(Q, o2) == 0
    assert torch.equal(o2, _colsum_scratch(Q))
