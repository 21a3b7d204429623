"""pot_b200 — B200-native entropic Sinkhorn hot path with PythonOT/POT's signatures.

    import pot_b200 as ot
    G = ot.sinkhorn(a, b, M, reg)                      # torch CUDA tensors in, torch CUDA tensors out
    G = ot.bregman.sinkhorn_log(a, b, M, reg)
    f, g = ot.bregman.empirical_sinkhorn(Xs, Xt, reg, isLazy=True)
    res = ot.batch.solve_batch(M, reg)

All arithmetic runs in libb2ot.so (hand-written sm_100a CUDA behind the C ABI of include/b2ot.h).
Importing the package loads that library and fails if it has not been built; there is no CPU path.
"""

from . import _lib

_lib.load()  # fail loudly when the CUDA library is missing

from . import batch, bregman, empirical, gromov, host, sharded, solvers, stabilized, unbalanced, utils  # noqa: E402
from . import batch_quadratic  # noqa: E402
from .batch import solve_batch  # noqa: E402

# the reference exposes the quadratic solvers under ot.batch as well
for _name in batch_quadratic.__all__:
    setattr(batch, _name, getattr(batch_quadratic, _name))
from .empirical import dist, empirical_sinkhorn, empirical_sinkhorn2, empirical_sinkhorn_divergence  # noqa: E402
from .solvers import solve, solve_sample  # noqa: E402
from .unbalanced import sinkhorn_unbalanced  # noqa: E402
from .utils import LazyTensor, OTResult  # noqa: E402

# the reference exposes the empirical solvers under ot.bregman as well
bregman.empirical_sinkhorn = empirical_sinkhorn
bregman.empirical_sinkhorn2 = empirical_sinkhorn2
bregman.get_sinkhorn_lazytensor = empirical.get_sinkhorn_lazytensor
bregman.empirical_sinkhorn_divergence = empirical_sinkhorn_divergence
bregman.sinkhorn_stabilized = stabilized.sinkhorn_stabilized
bregman.sinkhorn_epsilon_scaling = stabilized.sinkhorn_epsilon_scaling
from .bregman import sinkhorn, sinkhorn2, sinkhorn_knopp, sinkhorn_log  # noqa: E402

__version__ = "0.1.0"
__all__ = ["batch", "bregman", "empirical", "host", "sharded", "utils", "sinkhorn", "sinkhorn2", "sinkhorn_knopp",
           "sinkhorn_log", "solve_batch", "dist", "empirical_sinkhorn", "empirical_sinkhorn2",
           "empirical_sinkhorn_divergence", "solve", "solve_sample", "solvers", "stabilized", "gromov", "unbalanced", "sinkhorn_unbalanced",
           "LazyTensor", "OTResult"]
