"""Where does the time of a whole ot.bregman.sinkhorn_log call go (config 2: 16384^2 fp32, reg 0.01, 500 iterations)?
Counts library launches and times the call with torch events; prints the per-phase estimate."""
import os
import sys
import time
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402
from pot_b200 import _lib  # noqa: E402

n = m = int(os.environ.get("N", 16384))
dev = torch.device("cuda:0")
rng = np.random.RandomState(0)
xs = torch.from_numpy(rng.randn(n, 2).astype(np.float32)).to(dev)
xt = torch.from_numpy((rng.randn(m, 2) + 1.0).astype(np.float32)).to(dev)
M = ot.dist(xs, xt)
M /= M.max()
a = torch.full((n,), 1.0 / n, device=dev)
b = torch.full((m,), 1.0 / m, device=dev)
lib = _lib.load()
for fn, reg, name in ((ot.sinkhorn_log, 0.01, "sinkhorn_log"), (ot.bregman.sinkhorn_knopp, 0.05, "sinkhorn_knopp")):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        fn(a, b, M, reg, numItermax=20, stopThr=0)
        for its in (100, 500):
            torch.cuda.synchronize()
            c0 = lib.b2ot_launch_count()
            t0 = time.perf_counter()
            G, lg = fn(a, b, M, reg, numItermax=its, stopThr=0, log=True)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            print(f"{name} its={its}: wall {1e3 * (t1 - t0):.2f} ms, launches {lib.b2ot_launch_count() - c0}, "
                  f"niter {lg['niter']}, nerr {len(lg['err'])}", flush=True)

# ---- phase breakdown of the log-domain call ----
from pot_b200 import _dense  # noqa: E402


def tsync():
    torch.cuda.synchronize()
    return time.perf_counter()


for mode, reg in (("log", 0.01), ("knopp", 0.05)):
    for rep in range(2):
        t = [tsync()]
        Kt = _dense.cost_transform(M, reg, "negscale" if mode == "log" else "kexp")
        t.append(tsync())
        av = torch.log(a) if mode == "log" else a
        logb = torch.log(b) if mode == "log" else None
        solver = _dense.DenseSolver(mode, Kt, av, b, logb, max_iter=100, check_every=10, stop_thr=0.0)
        t.append(tsync())
        u0 = torch.zeros(n, device=dev) if mode == "log" else a.clone()
        v0 = torch.zeros(m, device=dev) if mode == "log" else b.clone()
        solver.init(u0, v0)
        t.append(tsync())
        solver.iterate(100)
        t.append(tsync())
        st = solver.poll()
        t.append(tsync())
        u, v, errs = solver.export()
        t.append(tsync())
        G = _dense.plan(mode, Kt, u, v)
        t.append(tsync())
        solver.close()
        t.append(tsync())
        names = ["cost_transform", "create", "init", "iterate(100)", "poll", "export", "plan", "close"]
        print(mode, rep, {k: round(1e3 * (t[i + 1] - t[i]), 3) for i, k in enumerate(names)}, flush=True)

# ---- BASELINE config 1 on the GPU path: n = m = 1000 fp64, reg 0.05, 1000 fixed iterations ----
rng = np.random.RandomState(0)
x1 = torch.from_numpy(rng.randn(1000, 2)).to(dev)
y1 = torch.from_numpy(rng.randn(1000, 2) + 1.0).to(dev)
M1 = ot.dist(x1, y1)
a1 = torch.full((1000,), 1e-3, device=dev, dtype=torch.float64)
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    ot.sinkhorn(a1, a1, M1, 0.05, numItermax=50, stopThr=0)
    for its in (1000, 1000, 1000):
        t0 = tsync()
        G1 = ot.sinkhorn(a1, a1, M1, 0.05, numItermax=its, stopThr=0)
        t1 = tsync()
        print(f"config 1 (1000^2 fp64, {its} its): wall {1e3 * (t1 - t0):.2f} ms = {its / (t1 - t0):.0f} it/s", flush=True)

# ---- per-kernel split at config 1 (sweep alone vs whole iteration, CUDA events on the launch stream) ----
K1 = _dense.cost_transform(M1, 0.05, "kexp")
s1 = _dense.DenseSolver("knopp", K1, a1, a1, None, max_iter=100000, check_every=10, stop_thr=0.0)
s1.init(a1.clone(), a1.clone())
s1.iterate(20)
torch.cuda.synchronize()
sw, itms = s1.profile(200)
t0 = tsync()
s1.iterate(1000)
t1 = tsync()
print(f"config 1 kernels: sweep {1e3 * sw:.2f} us, iteration (events around every launch) {1e3 * itms:.2f} us, "
      f"1000 graph-replayed iterations {1e3 * (t1 - t0):.2f} ms, info {s1.info()}", flush=True)
s1.close()
