"""ot.batch.solve_gromov_batch on one B200 next to the numpy oracle on the host cores: B problems of n x n (fp32),
`outer` proximal steps of `inner` log-domain projections each (fixed work).  The GPU call is the public
pot_b200.batch.solve_gromov_batch (projections in libb2ot.so, tensor product in cuBLAS).  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402
from oracle import solvers_oracle as sorc  # noqa: E402  (CPU leg only)


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    outer = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    inner = int(sys.argv[4]) if len(sys.argv) > 4 else 50
    rng = np.random.RandomState(0)
    xs, xt = rng.randn(B, n, 2).astype(np.float32), rng.randn(B, n, 3).astype(np.float32)
    C1 = ((xs[:, :, None] - xs[:, None]) ** 2).sum(-1)
    C2 = ((xt[:, :, None] - xt[:, None]) ** 2).sum(-1)
    C1, C2 = C1 / C1.max(), C2 / C2.max()
    kw = dict(reg=0.05, max_iter=outer, tol=0.0, max_iter_inner=inner, tol_inner=0.0)
    d1, d2 = torch.as_tensor(C1, device="cuda:0"), torch.as_tensor(C2, device="cuda:0")
    ot.batch.solve_gromov_batch(d1, d2, grad="detach", **kw)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = ot.batch.solve_gromov_batch(d1, d2, grad="detach", **kw)
    torch.cuda.synchronize()
    gpu_s = time.perf_counter() - t0
    bs = min(B, 64)
    t0 = time.perf_counter()
    r = sorc.solve_gromov_batch(C1[:bs], C2[:bs], **kw)
    cpu_s = (time.perf_counter() - t0) * B / bs
    rel = float(np.abs(res.plan[:bs].cpu().numpy() - r["plan"]).max() / r["plan"].max())
    print(json.dumps({"workload": f"solve_gromov_batch B={B} n=m={n} fp32, {outer} proximal steps x {inner} projections",
                      "gpu_ms": round(1e3 * gpu_s, 2), "problems_per_s": round(B / gpu_s, 1),
                      "cpu_oracle_ms_scaled_from_%d_problems" % bs: round(1e3 * cpu_s, 1), "cpu_threads": os.cpu_count(),
                      "max_rel_diff_vs_oracle": rel}))


if __name__ == "__main__":
    main()
