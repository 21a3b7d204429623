"""Golden vectors of the UNMODIFIED reference at (or near) the BASELINE.json sizes of configs 2, 4 and 5.

Run in the build container (needs oracle/_ref, built from /root/reference by oracle/build_ref.py):

    python tests/golden/make_golden_fullsize.py            # ~10 min of CPU

The inputs are NOT stored (they are GiB-sized): `tests/fullsize_inputs.py` regenerates them from the seeds with the
same numpy calls on the test side; the cost matrices are formed in float64 and then cast to float32, so a different
BLAS on the GPU box can move an entry by at most one float32 ulp.

  golden_fullsize.npz
    c2/log_u, c2/log_v, c2/err : ot.bregman.sinkhorn_log, n = m = 16384 fp32, reg 0.01, 20 fixed iterations (config 2)
    c4/f, c4/g, c4/err         : ot.bregman.empirical_sinkhorn(isLazy=True), n = m = 16384, d = 64, fp32, reg 0.1,
                                 10 fixed iterations (config 4's path, 128 x 128 tiles of the tensor-core sweep)
    c5/{solver}/value, value_linear, f_head, g_head : ot.batch.solve_batch, 4096 x 256 x 256 fp32 (config 5 at FULL
                                 size), reg 0.1, 20 fixed iterations, both solvers; potentials of the first 8 problems
"""
import os
import sys
import time
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from oracle.build_ref import import_ref  # noqa: E402
import fullsize_inputs as fin  # noqa: E402

ot = import_ref()
import ot.batch  # noqa: E402

G = {}
warnings.simplefilter("ignore")

t0 = time.time()
a, b, M = fin.config2_inputs()
_, lg = ot.bregman.sinkhorn_log(a, b, M, fin.C2_REG, numItermax=fin.C2_ITS, stopThr=0, log=True, warn=False)
assert lg["log_u"].dtype == np.float32
G["c2/log_u"], G["c2/log_v"], G["c2/err"] = lg["log_u"], lg["log_v"], np.asarray(lg["err"], dtype=np.float64)
G["c2/niter"] = np.asarray(lg["niter"])
del M
print("config 2 done", time.time() - t0, flush=True)

X, Y = fin.config4_inputs()
f, g, lg = ot.bregman.empirical_sinkhorn(X, Y, fin.C4_REG, numIterMax=fin.C4_ITS, stopThr=-1.0, isLazy=True,
                                         batchSize=512, log=True, warn=False)
G["c4/f"], G["c4/g"], G["c4/err"] = np.asarray(f), np.asarray(g), np.asarray(lg["err"], dtype=np.float64)
print("config 4 done", time.time() - t0, f.dtype, flush=True)

Mb = fin.config5_inputs()
for solver in ("log_sinkhorn", "sinkhorn"):
    val, vlin, fh, gh = [], [], [], []
    for c0 in range(0, Mb.shape[0], 256):  # chunks of 256 problems bound the (B, n, m) temporaries of the reference
        res = ot.batch.solve_batch(Mb[c0:c0 + 256], fin.C5_REG, max_iter=fin.C5_ITS, tol=0, solver=solver,
                                   reg_type="entropy")
        val.append(np.asarray(res.value))
        vlin.append(np.asarray(res.value_linear))
        if c0 == 0:
            fh, gh = np.asarray(res.potentials[0])[:8], np.asarray(res.potentials[1])[:8]
    G[f"c5/{solver}/value"] = np.concatenate(val)
    G[f"c5/{solver}/value_linear"] = np.concatenate(vlin)
    G[f"c5/{solver}/f_head"], G[f"c5/{solver}/g_head"] = fh, gh
    print("config 5", solver, "done", time.time() - t0, flush=True)

np.savez_compressed(os.path.join(HERE, "golden_fullsize.npz"), **G)
print("wrote golden_fullsize.npz:", {k: (v.shape, str(v.dtype)) for k, v in G.items()})
