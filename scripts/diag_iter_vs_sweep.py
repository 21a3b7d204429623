"""Headline shape (65536^2 fp32 knopp): whole-iteration time through iterate() (graph replays) next to the sweep-only and
per-iteration times of profile() (events around every launch), same handle, alternating.  A/B aid: run it from two
checkouts (sys.path[0] decides which pot_b200 / libb2ot.so is used)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pot_b200 import _dense, _lib  # noqa: E402

n = 65536
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
xs = torch.randn(n, 2, device=dev, generator=g)
xt = torch.randn(n, 2, device=dev, generator=g) + 1.0
M = torch.cdist(xs, xt).pow_(2)
M /= M.max()
lib = _lib.load()
_lib.check(lib.b2ot_kexp(M.data_ptr(), M.data_ptr(), n, n, M.stride(0), M.stride(0), 0.05, _lib.F32,
                         _dense.stream_ptr(M.device)), "b2ot_kexp")
a = torch.full((n,), 1.0 / n, device=dev)
b = torch.full((n,), 1.0 / n, device=dev)
s = _dense.DenseSolver("knopp", M, a, b, None, max_iter=10 ** 6, check_every=10, stop_thr=0.0)
s.init(a.clone(), b.clone())
s.iterate(20)
torch.cuda.synchronize()
out = {"lib": _lib.LIB_PATH, "runs": []}
for r in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s.iterate(100)
    e1.record()
    torch.cuda.synchronize()
    it_ms = e0.elapsed_time(e1) / 100
    sweep_ms, prof_iter_ms = s.profile(100)
    out["runs"].append({"iterate_ms_per_iter": round(it_ms, 4), "profile_sweep_ms": round(sweep_ms, 4),
                        "profile_iter_ms": round(prof_iter_ms, 4)})
print(json.dumps(out))
