"""BASELINE config 2: ot.bregman.sinkhorn_log dense fp32 n = m = 16384 (M = 1 GiB), reg = 0.01, 500 fixed iterations,
one B200.  Reports the whole public-API call (Mr = -M/reg, 500 iterations, plan) and the steady-state iteration rate."""
import json
import os
import sys
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402
from pot_b200 import _dense, _lib  # noqa: E402

n = int(os.environ.get("N", 16384))
m = int(os.environ.get("M", n))
reg, iters = 0.01, 500
dev = torch.device("cuda:0")
rng = np.random.RandomState(0)
xs = torch.from_numpy(rng.randn(n, 2).astype(np.float32)).to(dev)
xt = torch.from_numpy((rng.randn(m, 2) + 1.0).astype(np.float32)).to(dev)
M = ot.dist(xs, xt)
M /= M.max()
a = torch.full((n,), 1.0 / n, device=dev)
b = torch.full((m,), 1.0 / m, device=dev)
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    ot.sinkhorn_log(a, b, M, reg, numItermax=20, stopThr=0)  # warm-up
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    G, lg = ot.sinkhorn_log(a, b, M, reg, numItermax=iters, stopThr=0, log=True)
    e1.record()
    torch.cuda.synchronize()
api_ms = e0.elapsed_time(e1)
# steady state
lib = _lib.load()
Mr = torch.empty_like(M)
_lib.check(lib.b2ot_negscale(M.data_ptr(), Mr.data_ptr(), n, m, M.stride(0), Mr.stride(0), reg, _lib.F32,
                             _dense.stream_ptr(dev)), "negscale")
s = _dense.DenseSolver("log", Mr, torch.log(a), b, torch.log(b), max_iter=10 ** 7, check_every=10, stop_thr=0.0)
s.init(torch.zeros(n, device=dev), torch.zeros(m, device=dev))
s.iterate(20)
torch.cuda.synchronize()
e0.record()
s.iterate(iters)
e1.record()
torch.cuda.synchronize()
it_ms = e0.elapsed_time(e1) / iters
sweep_ms, _ = s.profile(20)
info = s.info()
s.close()
alg = 4.0 * n * m + 4.0 * (2 * n + 3 * m)
peak = 6550.0
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
print(json.dumps({"config": 2, "workload": f"ot.bregman.sinkhorn_log dense fp32 n=m={n}, reg={reg}, {iters} fixed iterations",
                  "api_call_ms": api_ms, "api_iterations_per_s": iters / (api_ms * 1e-3), "ms_per_iteration": it_ms,
                  "it_per_s": 1e3 / it_ms, "sweep_ms": sweep_ms, "algorithmic_GBps": alg / (sweep_ms * 1e-3) / 1e9,
                  "hbm_peak_GBps": peak, "frac": alg / (sweep_ms * 1e-3) / 1e9 / peak, "niter": lg["niter"],
                  "marginal_err_last": float(lg["err"][-1]), "sweep": info}))
