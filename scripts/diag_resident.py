"""Phase timing of the resident solver (diagnosis build: make DIAG=1 OUT=../libb2ot_diag.so; B2OT_LIB=... B2OT_RES_TRACE=1)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pot_b200 import _dense

n = m = int(os.environ.get("N", "1000"))
dt = torch.float64 if os.environ.get("DT", "f64") == "f64" else torch.float32
rng = np.random.RandomState(0)
K = torch.as_tensor(np.exp(-rng.rand(n, m) / 0.3), device="cuda", dtype=dt)
a = torch.full((n,), 1.0 / n, dtype=dt, device="cuda")
b = torch.full((m,), 1.0 / m, dtype=dt, device="cuda")
s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=100000, check_every=10, stop_thr=0.0)
print(s.info())
s.init(a.clone(), b.clone())
s.iterate(200)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); s.iterate(1000); e1.record(); torch.cuda.synchronize()
print("us per iteration:", e0.elapsed_time(e1))
s.iterate(64)
torch.cuda.synchronize()
tr = s.ws[256:256 + 64 * 8 * 8].view(torch.int64).cpu().numpy().reshape(64, 8)
d = np.diff(tr[:, :6], axis=1)
print("median clocks per phase [rowpass, gatherA, colpass, gatherB, decide]:", np.median(d[8:], axis=0))
print("iteration period (clk):", np.median(np.diff(tr[8:, 0])))
