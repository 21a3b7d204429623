"""Gromov-Wasserstein contraction const - hC1 T hC2^T: the library's strided matrix-product kernel next to torch.mm
(cuBLAS) on the same operands.  Prints one JSON line per case."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pot_b200 import _gemm  # noqa: E402


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for dtype in (torch.float64, torch.float32):
    for n in (1000, 4000):
        g = torch.Generator(device="cuda").manual_seed(0)
        h1, h2 = (torch.randn(n, n, device="cuda", dtype=dtype, generator=g) for _ in range(2))
        T, c = (torch.randn(n, n, device="cuda", dtype=dtype, generator=g) for _ in range(2))
        ours = timed(lambda: _gemm.contract(c, h1, T, h2))
        lib = timed(lambda: c - torch.mm(torch.mm(h1, T), h2.t()))
        err = float((_gemm.contract(c, h1, T, h2) - (c - torch.mm(torch.mm(h1, T), h2.t()))).abs().max())
        fl = 4.0 * n ** 3
        print(json.dumps({"n": n, "dtype": str(dtype).split(".")[-1], "ours_ms": round(ours, 3),
                          "ours_tflops": round(fl / ours / 1e9, 2), "torch_mm_ms": round(lib, 3),
                          "torch_mm_tflops": round(fl / lib / 1e9, 2), "max_abs_diff": err}), flush=True)
