"""Seeded inputs of the BASELINE-size parity cases (shared by tests/golden/make_golden_fullsize.py, which feeds them to
the unmodified reference, and tests/test_fullsize_gpu.py, which feeds the SAME arrays to the CUDA path).

Costs are formed in float64 and cast to float32 last, so the host BLAS of the box that regenerates them can move an
entry by at most one float32 ulp relative to the arrays the golden run saw."""
import numpy as np

C2_N, C2_REG, C2_ITS = 16384, 0.01, 20
C4_N, C4_D, C4_REG, C4_ITS = 16384, 64, 0.1, 10
C5_B, C5_N, C5_REG, C5_ITS = 4096, 256, 0.1, 20


def sqeuclid64(xs, xt):
    xs, xt = xs.astype(np.float64), xt.astype(np.float64)
    M = (xs ** 2).sum(1)[:, None] + (xt ** 2).sum(1)[None, :] - 2.0 * xs @ xt.T
    np.maximum(M, 0.0, out=M)
    return M


def config2_inputs(n=C2_N):
    """BASELINE config 2: 2-D Gaussian clouds seed 0, M = sqeuclidean / max, uniform weights, fp32."""
    rng = np.random.RandomState(0)
    xs = rng.randn(n, 2).astype(np.float32)
    xt = (rng.randn(n, 2) + 1.0).astype(np.float32)
    M = sqeuclid64(xs, xt)
    M /= M.max()
    a = np.full(n, 1.0 / n, dtype=np.float32)
    return a, a.copy(), M.astype(np.float32)


def config4_inputs(n=C4_N, d=C4_D):
    """BASELINE config 4's clouds (X /= sqrt(d), costs O(1)) at a size the reference finishes in minutes."""
    rng = np.random.RandomState(4)
    X = (rng.randn(n, d) / np.sqrt(d)).astype(np.float32)
    Y = (rng.randn(n, d) / np.sqrt(d) + 1.0 / np.sqrt(d)).astype(np.float32)
    return X, Y


def config5_inputs(B=C5_B, n=C5_N):
    """BASELINE config 5 at full size: B problems of n x n, 2-D clouds per problem, M / global max, fp32."""
    rng = np.random.RandomState(5)
    X = rng.randn(B, n, 2)
    Y = rng.randn(B, n, 2) + 1.0
    M = (X ** 2).sum(-1)[:, :, None] + (Y ** 2).sum(-1)[:, None, :] - 2.0 * np.einsum("bid,bjd->bij", X, Y)
    M /= M.max()
    return M.astype(np.float32)
