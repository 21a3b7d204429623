"""sinkhorn_stabilized vs sinkhorn_knopp through the public call (pot_b200.bregman) on one B200: fp32 / fp64,
n = m = N, small reg so that absorptions happen.  Reports wall time per iteration of the whole call (CUDA events around
the call, K build and final plan included), the number of absorptions and the time of the knopp call with the same
iteration count (same sweep kernel, no tau test).  Prints one JSON line per case.
"""
import json
import os
import sys
import warnings

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402
from pot_b200 import stabilized  # noqa: E402


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e30
    for _ in range(reps):
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best, out


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    its = int(sys.argv[2]) if len(sys.argv) > 2 else 500
    dev = "cuda:0"
    warnings.simplefilter("ignore")
    for dt, reg, tau in ((torch.float32, 0.01, 1e3), (torch.float64, 0.002, 1e3)):
        if dt == torch.float64 and n > 32768:
            continue
        g = torch.Generator(device=dev).manual_seed(0)
        xs = torch.randn(n, 2, device=dev, generator=g, dtype=dt)
        xt = torch.randn(n, 2, device=dev, generator=g, dtype=dt) + 1.0
        M = torch.cdist(xs, xt).pow_(2)
        M /= M.max()
        a = torch.full((n,), 1.0 / n, device=dev, dtype=dt)
        b = torch.full((n,), 1.0 / n, device=dev, dtype=dt)
        ms_s, core = timed(lambda: stabilized._stabilized_core(a, b, M, reg, its, tau, 0.0, None, 20))
        ms_full, _ = timed(lambda: ot.bregman.sinkhorn_stabilized(a, b, M, reg, numItermax=its, tau=tau, stopThr=0.0))
        ms_k, _ = timed(lambda: ot.bregman.sinkhorn_knopp(a, b, M, max(reg, 0.02), numItermax=its, stopThr=0.0))
        es = M.element_size()
        sweep_bytes = es * n * n * (its + 2 * core[8])  # sweep reads K; an absorption reads Mr and writes K
        print(json.dumps({
            "workload": f"sinkhorn_stabilized n=m={n} {str(dt).split('.')[-1]} reg={reg} tau={tau} {its} iterations",
            "absorptions": core[8], "n_iter": core[6], "loop_ms": round(ms_s, 2), "call_ms": round(ms_full, 2),
            "loop_us_per_iter": round(1e3 * ms_s / its, 2),
            "loop_GBps_3_passes_per_absorbing_iteration": round(sweep_bytes / ms_s / 1e6, 1),
            "knopp_call_ms_same_iterations": round(ms_k, 2)}))


if __name__ == "__main__":
    main()
