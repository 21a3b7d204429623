"""Per-rank shape of the 8-GPU headline (8192 x 65536 fp32 knopp, K = 2 GiB) on one GPU: sweep time under the
B2OT_* tuning knobs (read at handle creation).  profile() = CUDA events around every sweep launch."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pot_b200 import _dense, _lib  # noqa: E402

n, m = int(os.environ.get("ROWS", 8192)), 65536
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
K = torch.rand(n, m, device=dev, generator=g) + 0.5
a = torch.full((n,), 1.0 / n, device=dev)
b = torch.full((m,), 1.0 / m, device=dev)
KNOBS = [{}, {"B2OT_OCC": "2"}, {"B2OT_WARPS": "16"}, {"B2OT_CLUSTERS": "14"}, {"B2OT_CLUSTERS": "12"},
         {"B2OT_L2_HINT": "0"}, {"B2OT_STAGES": "2"}, {}]
if os.environ.get("DEFAULT_ONLY"):
    KNOBS = [{}]
peak = 6535.1
for kn in KNOBS:
    for k in ("B2OT_OCC", "B2OT_WARPS", "B2OT_CLUSTERS", "B2OT_L2_HINT", "B2OT_STAGES"):
        os.environ.pop(k, None)
    os.environ.update(kn)
    s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=10 ** 6, check_every=10, stop_thr=0.0)
    s.init(a.clone(), b.clone())
    s.iterate(20)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        sw, it = s.profile(100)
        best = min(best, sw)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); s.iterate(200); e1.record(); torch.cuda.synchronize()
    nbytes = 4.0 * n * m + 4.0 * (2 * n + 3 * m)
    print(json.dumps({"knobs": kn, "info": s.info(), "sweep_us": round(1e3 * best, 1), "sweep_GBps": round(nbytes / best / 1e6, 1),
                      "frac": round(nbytes / best / 1e6 / peak, 3), "iterate_us_per_iter": round(1e3 * e0.elapsed_time(e1) / 200, 1)}))
    s.close()
