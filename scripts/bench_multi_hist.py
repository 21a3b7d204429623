"""Time of b2ot_multi_scaling_step (sinkhorn_stabilized with a 2-D b: H histograms on one kernel matrix).

    python scripts/bench_multi_hist.py

Synthetic K; 20 steps between device events after 3 warm-up steps; algorithmic bytes = two passes over K."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pot_b200 import _dense, _lib  # noqa: E402


def main():
    lib = _lib.load()
    dev = torch.device("cuda:0")
    cases = ((torch.float32, 16384, 16384), (torch.float64, 8192, 16384), (torch.float64, 1000, 1000))
    if len(sys.argv) > 1:  # e.g. "0" or "0,2": a subset of the cases (short runs under ncu)
        cases = tuple(cases[int(x)] for x in sys.argv[1].split(","))
    for dt, n, m in cases:
        K = torch.rand(n, m, device=dev, dtype=dt) + 0.1
        a = torch.full((n,), 1.0 / n, device=dev, dtype=dt)
        for H in (tuple(int(x) for x in sys.argv[2].split(",")) if len(sys.argv) > 2 else (1, 4, 8)):
            B = torch.full((m, H), 1.0 / m, device=dev, dtype=dt)
            u = [torch.full((n, H), 1.0 / n, device=dev, dtype=dt) for _ in range(2)]
            v = [torch.full((m, H), 1.0 / m, device=dev, dtype=dt) for _ in range(2)]
            stats = torch.zeros(16, dtype=torch.float64, device=dev)
            eh = torch.zeros(64, dtype=torch.float64, device=dev)
            scratch = torch.empty(lib.b2ot_multi_scratch_bytes(n, m, H), dtype=torch.uint8, device=dev)

            def step(k):
                s, d = k & 1, (k & 1) ^ 1
                rc = lib.b2ot_multi_scaling_step(_dense.dtype_code(dt), _lib.MULTI_KNOPP, K.data_ptr(), n, m, K.stride(0),
                                                 H, a.data_ptr(), B.data_ptr(), u[s].data_ptr(), v[s].data_ptr(),
                                                 u[d].data_ptr(), v[d].data_ptr(), k, 10, 0.0, 1.0, 1.0,
                                                 stats.data_ptr(), eh.data_ptr(), 64, scratch.data_ptr(),
                                                 _dense.stream_ptr(dev))
                assert rc == 0

            for k in range(3):
                step(k)
            assert stats[12].item() == 0.0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for k in range(3, 23):
                step(k)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1000.0 / 20
            gbs = 2.0 * n * m * K.element_size() / (us * 1e-6) / 1e9
            print(json.dumps({"dtype": str(dt).split(".")[1], "n": n, "m": m, "H": H, "us_per_step": round(us, 1),
                              "GBps_two_passes_over_K": round(gbs, 0), "finite": bool(torch.isfinite(u[0]).all())}),
                  flush=True)


if __name__ == "__main__":
    main()
