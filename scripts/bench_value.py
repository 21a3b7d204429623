"""Times the plan / objective epilogue (b2ot_dense_plan_value) against its HBM roofline.

    python scripts/bench_value.py [n] [m]      # default 16384 x 16384 fp32

Algorithmic bytes per launch: s*n*m (M read once) + s*n*m (plan written once); values-only: s*n*m.
Timed with CUDA events on the launch stream after warm-up; prints one JSON line per variant.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402
from pot_b200 import _dense  # noqa: E402


def timeit(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    m = int(sys.argv[2]) if len(sys.argv) > 2 else n
    peak = 6550.0
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    for dtype in (torch.float32, torch.float64):
        s = 4 if dtype == torch.float32 else 8
        if s * n * m * 3 > 120e9:
            continue
        g = torch.Generator(device="cuda").manual_seed(0)
        M = torch.rand((n, m), device="cuda", dtype=dtype, generator=g)
        a = torch.full((n,), 1.0 / n, device="cuda", dtype=dtype)
        b = torch.full((m,), 1.0 / m, device="cuda", dtype=dtype)
        f = torch.log(a) + 0.1 * torch.rand(n, device="cuda", dtype=dtype)
        gg = torch.log(b) + 0.1 * torch.rand(m, device="cuda", dtype=dtype)
        for name, terms, want_plan in (("plan+linear+kl", ("linear", "kl"), True),
                                       ("plan+linear+entropy", ("linear", "entropy"), True),
                                       ("plan only", (), True),
                                       ("linear only (no plan)", ("linear",), False)):
            ms = timeit(lambda: _dense.plan_value("log", f, gg, M=M, reg=0.1, a=a, b=b, terms=terms,
                                                  want_plan=want_plan))
            nbytes = s * n * m * (2 if want_plan else 1)
            print(json.dumps({"kernel": "plan_value_kernel", "variant": name, "dtype": str(dtype), "n": n, "m": m,
                              "ms": round(ms, 4), "GBps": round(nbytes / ms / 1e6, 1),
                              "frac_of_measured_hbm": round(nbytes / ms / 1e6 / peak, 3)}))
        # the separate kernels this replaces in the sinkhorn_log + sinkhorn2 path
        Mr = _dense.cost_transform(M, 0.1, "negscale")
        ms_plan = timeit(lambda: _dense.plan("log", Mr, f, gg))
        ms_loss = timeit(lambda: _dense.loss("log", Mr, M, f, gg))
        print(json.dumps({"kernel": "plan_kernel", "dtype": str(dtype), "ms": round(ms_plan, 4),
                          "GBps": round(2 * s * n * m / ms_plan / 1e6, 1)}))
        print(json.dumps({"kernel": "loss_kernel", "dtype": str(dtype), "ms": round(ms_loss, 4),
                          "GBps": round(2 * s * n * m / ms_loss / 1e6, 1)}))
        for op in ("kexp", "negscale"):
            ms_t = timeit(lambda: _dense.cost_transform(M, 0.1, op))
            print(json.dumps({"kernel": f"cost_transform ({op})", "dtype": str(dtype), "ms": round(ms_t, 4),
                              "GBps": round(2 * s * n * m / ms_t / 1e6, 1),
                              "frac_of_measured_hbm": round(2 * s * n * m / ms_t / 1e6 / peak, 3)}))
        del M, Mr


if __name__ == "__main__":
    main()
