"""BASELINE configs 4 and 5 on N GPUs of one box (torchrun, one rank per GPU):

  config 4: ot.bregman.empirical_sinkhorn, online cost, log domain, n = m = 262144 points, d = 64, fp32, source points
            sharded across the ranks (one exchange of the column log-sum-exp partials per iteration), fixed iterations.
  config 5: ot.batch batched Sinkhorn, 4096 independent 256 x 256 fp32 problems sharded across the ranks with no
            communication; per-iteration cost from the difference of two fixed-iteration solves.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/bench_configs45.py [--n 262144] [--iters 20] [--exchange peer|nccl]
Prints one JSON line per config from rank 0 (times: CUDA events, max over ranks).
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402,F401
from pot_b200 import batch as otb  # noqa: E402
from pot_b200.empirical import EmpiricalSolver  # noqa: E402
from pot_b200.sharded import shard_rows  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=262144)
    ap.add_argument("--d", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--exchange", default="peer")
    ap.add_argument("--B", type=int, default=4096)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD

    def sync():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- config 4 ----------------
    n = m = args.n
    d = args.d
    g = torch.Generator(device=dev).manual_seed(0)  # same seed on every rank: identical clouds, each keeps its block
    xs = torch.randn(n, d, device=dev, generator=g) / d ** 0.5
    xt = torch.randn(m, d, device=dev, generator=g) / d ** 0.5 + 1.0 / d ** 0.5
    r0, r1 = shard_rows(n, world, rank)
    xs_loc = xs[r0:r1].contiguous()
    del xs
    a_loc = torch.full((r1 - r0,), 1.0 / n, device=dev)
    b = torch.full((m,), 1.0 / m, device=dev)
    s = EmpiricalSolver(xs_loc, xt, a_loc, b, 0.1, max_iter=10 ** 6, stop_thr=-1.0, variant="tc", group=group,
                        exchange=args.exchange)
    s.init(torch.zeros(r1 - r0, device=dev), torch.zeros(m, device=dev))
    s.iterate(3)
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s.iterate(args.iters)
    e1.record()
    sync()
    ms = max_ranks(e0.elapsed_time(e1) / args.iters)
    f, gg, errs = s.export()
    info = s.info()
    finite = bool(torch.isfinite(f).all() and torch.isfinite(gg).all())
    s.close()
    if rank == 0:
        flops = 2 * 2.0 * n * m * d
        print(json.dumps({"config": 4, "workload": f"empirical_sinkhorn online-cost log-domain n=m={n} d={d} fp32, reg=0.1, "
                          f"source points sharded over {world} GPU(s), exchange={args.exchange}",
                          "n_gpus": world, "ms_per_iteration": ms, "it_per_s": 1e3 / ms,
                          "fp32_equiv_gemm_tflops": flops / (ms * 1e-3) / 1e12,
                          "tensor_tflops_3x": 3 * flops / (ms * 1e-3) / 1e12,
                          "exp_per_s": 2.0 * n * m / (ms * 1e-3), "finite": finite, "info": info,
                          "kind": os.environ.get("B2OT_EMP_KIND", "f16")}), flush=True)
    del xs_loc, xt
    torch.cuda.empty_cache()

    # ---------------- config 5 ----------------
    B, bn, bm = args.B, 256, 256
    p0, p1 = shard_rows(B, world, rank)
    Bl = p1 - p0
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    X = torch.randn(Bl, bn, 2, device=dev, generator=g)
    Y = torch.randn(Bl, bm, 2, device=dev, generator=g) + 1.0
    M = (X * X).sum(-1)[:, :, None] + (Y * Y).sum(-1)[:, None, :] - 2 * torch.einsum("bid,bjd->bij", X, Y)
    M = (M / M.max()).float().contiguous()
    a = torch.full((Bl, bn), 1.0 / bn, device=dev)
    bb = torch.full((Bl, bm), 1.0 / bm, device=dev)
    out = {}
    for solver, mode, tr in (("sinkhorn", "knopp", 1), ("log_sinkhorn", "log", 2)):
        def run(its):
            return otb._run(mode, M, tr, 0.1, M, a, bb, its, 0.0, True, False, 1, sync_n_iters=False)

        def timed(its, reps=3):
            best = 1e9
            for _ in range(reps):
                sync()
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record()
                run(its)
                t1.record()
                sync()
                best = min(best, max_ranks(t0.elapsed_time(t1)))
            return best

        run(10)
        t10, t110 = timed(10), timed(110)
        per_it = (t110 - t10) / 100.0
        out[solver] = {"ms_solve_100_iterations_incl_load_and_plan": t110 - t10 + t10,
                       "us_per_iteration_all_problems": per_it * 1e3,
                       "problem_iterations_per_s": B / (per_it * 1e-3),
                       "solves_per_s_100_iterations": B / (t110 * 1e-3) * (100.0 / 110.0)}
    if rank == 0:
        print(json.dumps({"config": 5, "workload": f"ot.batch {B} x 256 x 256 fp32, reg=0.1, problems sharded over {world} "
                          "GPU(s), no communication", "n_gpus": world, "results": out}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
