"""Timing of BASELINE config 5 (4096 x 256 x 256 fp32, batched solver) on one GPU: per-iteration cost from the
difference of two fixed-iteration solves (the load / epilogue cost cancels)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402
from pot_b200 import batch as otb  # noqa: E402


def timed(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    B, n, m = int(os.environ.get("B", 4096)), 256, 256
    dev = "cuda:0"
    g = torch.Generator(device=dev).manual_seed(0)
    X = torch.randn(B, n, 2, device=dev, generator=g)
    Y = torch.randn(B, m, 2, device=dev, generator=g) + 1.0
    M = (X * X).sum(-1)[:, :, None] + (Y * Y).sum(-1)[:, None, :] - 2 * torch.einsum("bid,bjd->bij", X, Y)
    M = (M / M.max()).float().contiguous()
    a = torch.full((B, n), 1.0 / n, device=dev)
    b = torch.full((B, m), 1.0 / m, device=dev)
    out = {}
    for solver, mode, tr in (("sinkhorn", "knopp", 1), ("log_sinkhorn", "log", 2)):
        def run(its):
            return otb._run(mode, M, tr, 0.1, M, a, b, its, 0.0, True, False, 1, sync_n_iters=False)
        run(10)
        t10, t110 = timed(lambda: run(10)), timed(lambda: run(110))
        per_it = (t110 - t10) / 100.0
        out[solver] = {"ms_10_its": t10, "ms_110_its": t110, "us_per_iteration_all_problems": per_it * 1e3,
                       "problem_iterations_per_s": B / (per_it * 1e-3), "config": otb.batch_config(mode, torch.float32, n, m)}
    print(json.dumps({"B": B, "n": n, "m": m, "dtype": "f32", "results": out}))


if __name__ == "__main__":
    main()
