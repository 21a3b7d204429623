"""entropic_gromov_wasserstein (PGD, square loss) on one B200 next to the numpy oracle on the host cores:
n = m = N points, `outer` projected-gradient iterations of `inner` fixed Sinkhorn iterations each (fp64).
The GPU call is the public pot_b200.gromov entry (inner solves in libb2ot.so, the two gradient products on b2ot_strided_gemm).
Prints one JSON line."""
import json
import os
import sys
import time
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402
from oracle import solvers_oracle as sorc  # noqa: E402  (CPU leg only)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    outer = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    inner = int(sys.argv[3]) if len(sys.argv) > 3 else 200
    warnings.simplefilter("ignore")
    rng = np.random.RandomState(0)
    xs, xt = rng.randn(n, 2), rng.randn(n, 3)
    C1 = ((xs[:, None] - xs[None]) ** 2).sum(-1)
    C2 = ((xt[:, None] - xt[None]) ** 2).sum(-1)
    C1, C2 = C1 / C1.max(), C2 / C2.max()
    kw = dict(epsilon=0.05, max_iter=outer, tol=0.0, numItermax=inner, stopThr=0.0)
    d1, d2 = torch.as_tensor(C1, device="cuda:0"), torch.as_tensor(C2, device="cuda:0")
    ot.gromov.entropic_gromov_wasserstein(d1, d2, **kw)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    T = ot.gromov.entropic_gromov_wasserstein(d1, d2, **kw)
    torch.cuda.synchronize()
    gpu_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    Tr = sorc.entropic_gromov_wasserstein(C1, C2, **{**kw, "max_iter": 1})
    cpu_s = (time.perf_counter() - t0) * outer
    Tr = sorc.entropic_gromov_wasserstein(C1, C2, **{**kw, "max_iter": 2})
    T2 = ot.gromov.entropic_gromov_wasserstein(d1, d2, **{**kw, "max_iter": 2})
    rel = float(np.abs(T2.cpu().numpy() - Tr).max() / Tr.max())
    print(json.dumps({"workload": f"entropic GW PGD square_loss n=m={n} fp64, {outer} outer x {inner} inner iterations",
                      "gpu_ms": round(1e3 * gpu_s, 2), "gpu_ms_per_outer": round(1e3 * gpu_s / outer, 3),
                      "cpu_oracle_ms_per_outer": round(1e3 * cpu_s / outer, 1), "cpu_threads": os.cpu_count(),
                      "max_rel_diff_vs_oracle_after_2_outer": rel}))


if __name__ == "__main__":
    main()
