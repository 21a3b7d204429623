"""A/B of the finalize variant (B2OT_FIN_QSPLIT) at 16384^2: knopp fp32 / fp64 and log fp32, iterate() timing."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pot_b200 import _dense, _lib
n = 16384
dev = "cuda:0"
out = {"qsplit": os.environ.get("B2OT_FIN_QSPLIT", "default")}
CASES = (("knopp", torch.float32), ("knopp", torch.float64), ("log", torch.float32))
if os.environ.get("AB_CASES") == "log64":
    CASES = (("log", torch.float64),)
for mode, dt in CASES:
    g = torch.Generator(device=dev).manual_seed(0)
    M = torch.rand(n, n, device=dev, generator=g, dtype=dt)
    K = _dense.cost_transform(M, 0.05, "kexp" if mode == "knopp" else "negscale")
    del M
    a = torch.full((n,), 1.0 / n, device=dev, dtype=dt)
    if mode == "knopp":
        s = _dense.DenseSolver("knopp", K, a, a, None, max_iter=10 ** 6, check_every=10, stop_thr=0.0)
        s.init(a.clone(), a.clone())
    else:
        s = _dense.DenseSolver("log", K, torch.log(a), a, torch.log(a), max_iter=10 ** 6, check_every=10, stop_thr=0.0)
        s.init(torch.zeros_like(a), torch.zeros_like(a))
    s.iterate(50)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); s.iterate(300); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 300)
    out[f"{mode}_{str(dt).split('.')[-1]}_us"] = round(1e3 * best, 1)
    s.close(); del K
print(json.dumps(out))
