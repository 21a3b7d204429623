"""Timing of the online-cost (empirical) solver on one GPU: ms per iteration (two sweeps) for the tensor-core and
the register-tiled variants at growing n = m, d = 64 fp32 (BASELINE config 4 is n = m = 262144 over 8 GPUs)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402
from pot_b200.empirical import EmpiricalSolver  # noqa: E402


def run(n, m, d, variant, iters):
    dev = "cuda:0"
    g = torch.Generator(device=dev).manual_seed(0)
    xs = torch.randn(n, d, device=dev, generator=g) / d ** 0.5
    xt = torch.randn(m, d, device=dev, generator=g) / d ** 0.5 + 1.0 / d ** 0.5
    a = torch.full((n,), 1.0 / n, device=dev)
    b = torch.full((m,), 1.0 / m, device=dev)
    s = EmpiricalSolver(xs, xt, a, b, 0.1, max_iter=10 ** 6, stop_thr=-1.0, variant=variant)
    s.init(torch.zeros(n, device=dev), torch.zeros(m, device=dev))
    s.iterate(2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s.iterate(iters)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    f, gg, errs = s.export()
    info = s.info()
    s.close()
    flops = 2 * 2.0 * n * m * d  # two sweeps
    return {"n": n, "m": m, "d": d, "variant": info["variant"], "ms_per_iter": ms, "it_per_s": 1e3 / ms,
            "gemm_tflops_fp32_equiv": flops / (ms * 1e-3) / 1e12, "tf32_tflops_3x": 3 * flops / (ms * 1e-3) / 1e12,
            "exp_per_s": 2.0 * n * m / (ms * 1e-3), "finite": bool(torch.isfinite(f).all() and torch.isfinite(gg).all()),
            "info": info}


if __name__ == "__main__":
    sizes = [int(x) for x in os.environ.get("SIZES", "8192,32768,65536").split(",")]
    for n in sizes:
        for variant in os.environ.get("VARIANTS", "tc,simt").split(","):
            if variant == "simt" and n > 32768:
                continue
            iters = max(2, min(20, int(2e11 / (n * n * (1 if variant == "tc" else 8)))))
            print(json.dumps(run(n, n, 64, variant, iters)), flush=True)
