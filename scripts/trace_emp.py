"""Pipeline trace of the tcgen05 sweep (CTA (0,0) of the last launch): clock64 stamps per tile.
slots: 0 producer got b_empty | 1 MMA got acc_empty | 2 MMA got b_full | 3 MMA issued+committed |
       4 epilogue iteration start | 5 colvec stored (prefetched loads landed) | 6 epilogue got acc_full | 7 tile done"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pot_b200.empirical import EmpiricalSolver

n = int(os.environ.get("N", "32768")); d = 64; dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
xs = torch.randn(n, d, device=dev, generator=g) / d ** 0.5
xt = torch.randn(n, d, device=dev, generator=g) / d ** 0.5 + 1.0 / d ** 0.5
a = torch.full((n,), 1.0 / n, device=dev); b = torch.full((n,), 1.0 / n, device=dev)
tr = torch.zeros(256 * 8, dtype=torch.int64, device=dev)
s = EmpiricalSolver(xs, xt, a, b, 0.1, max_iter=10 ** 6, stop_thr=-1.0, variant="tc")
s.init(torch.zeros(n, device=dev), torch.zeros(n, device=dev))
s.iterate(2)
torch.cuda.synchronize()
os.environ["B2OT_EMP_TRACE"] = hex(tr.data_ptr())
s.iterate(1)
torch.cuda.synchronize()
t = tr.view(256, 8).cpu()
t0 = int(t[t > 0].min())
print("info", s.info(), "dbg", os.environ.get("B2OT_EMP_DBG"))
print("tile  prod_bempty  mma_accempty  mma_bfull  mma_issued | epi_start colvec  acc_full  done")
for i in range(0, 40):
    r = [int(x) - t0 if x > 0 else -1 for x in t[i]]
    print(f"{i:3d} " + " ".join(f"{x:9d}" for x in r))
s.close()
