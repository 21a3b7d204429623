#!/usr/bin/env python
"""Headline benchmark: dense fp32 Sinkhorn-Knopp iterations/sec at n=m=65536, row-sharded over N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n 65536]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): 2-D Gaussian point clouds
(seed 0, target cloud shifted by +1), M = squared euclidean cost / max(M), uniform weights, reg = 0.05,
ot.bregman.sinkhorn_knopp with a fixed iteration count (stopThr = 0).  A "step" is one Sinkhorn
iteration = one fused sweep over this rank's rows of K = exp(-M/reg) + the all-reduce of K^T u (N > 1)
+ the finalize kernel.  K (16 GiB) is far larger than the 126 MB L2, so every iteration streams from HBM.

value : iterations/s with K resident in HBM, CUDA events on the launch stream, max over ranks.
e2e   : iterations/s of one whole solve of K iterations through pot_b200.host.sinkhorn_from_host with M in
        pinned HOST memory: host->device upload of M (overlapped with the Gibbs-kernel transform), the solve,
        and the device->host read of (u, v, err) are all inside the timed region.
--impl reference: the CPU restatement of the reference's numpy path (oracle/sinkhorn_oracle.py, same
        np.dot / np.exp calls as ot/bregman/_sinkhorn.py:621-631) on the host cores, on a bounded row sample.
"""

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

REG = 0.05
SEED = 0
D = 2


# ------------------------------------------------------------------------------------------------
def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=65536, help="problem size n = m (default: the headline 65536)")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1: per-iteration exchange of K^T u (peer = fused NVLink peer-memory exchange, "
                         "nccl = torch.distributed all-reduce baseline)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def point_clouds(n, m):
    import numpy as np

    rng = np.random.RandomState(SEED)
    xs = rng.randn(n, D).astype(np.float32)
    xt = (rng.randn(m, D) + 1.0).astype(np.float32)
    return xs, xt


class ClockSampler:
    """nvidia-smi clocks / throttle reasons of one GPU, sampled every 200 ms while a region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the oracle's numpy Sinkhorn-Knopp on a bounded row sample
# ------------------------------------------------------------------------------------------------
def cpu_knopp_sample(n, m, rows, iters, warm):
    """Times `iters` iterations of the reference's scaling-domain loop (K^T u GEMV, Kp v GEMV, guard test;
    marginal check every 10th) on the sub-problem made of the first `rows` source points against all m
    targets; returns full-problem iterations/s (per-iteration work scales with the row count)."""
    import numpy as np

    from oracle import sinkhorn_oracle as orc

    xs, xt = point_clouds(n, m)
    M = orc.sqeuclidean(xs[:rows], xt)
    M = (M / M.max()).astype(np.float32)
    a = np.full(rows, 1.0 / rows, dtype=np.float32)
    b = np.full(m, 1.0 / m, dtype=np.float32)
    import warnings

    K = np.exp(M / (-REG))
    Kp = (1 / a).reshape(-1, 1) * K
    u = np.ones(rows, dtype=np.float32) / rows
    v = np.ones(m, dtype=np.float32) / m
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        u, v, _, _ = orc.knopp_loop(K, Kp, b, u, v, max(1, warm), 0, {"err": []})
        t0 = time.perf_counter()
        orc.knopp_loop(K, Kp, b, u, v, iters, 0, {"err": []})
        per_iter = (time.perf_counter() - t0) / iters
    return (1.0 / per_iter) * (rows / n), per_iter, os.cpu_count()


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = m = args.n
    rows = min(n, 4096)
    its, per_iter, cores = cpu_knopp_sample(n, m, rows, max(1, args.steps), max(1, args.warmup))
    line = {
        "impl": "reference", "metric": "sinkhorn_knopp_iterations_per_sec", "value": its, "unit": "it/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / its,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"ot.bregman.sinkhorn_knopp dense fp32 n=m={n}, reg={REG}, uniform weights, "
                               "2-D Gaussian clouds, fixed iterations (stopThr=0)"},
        "cpu_baseline": {"value": its, "unit": "it/s", "cores": cores, "kind": "port",
                         "sample": f"first {rows} of {n} rows (all {m} columns), numpy/BLAS threads = all cores; "
                                   f"measured {per_iter * 1e3:.2f} ms per sample iteration, scaled by {rows}/{n}"},
        "e2e": {"value": its, "unit": "it/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
def ours(args):
    import torch
    import torch.distributed as dist

    import pot_b200
    from pot_b200 import _dense, _lib
    from pot_b200.host import sinkhorn_from_host
    from pot_b200.sharded import shard_rows

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: pot_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
    lib = _lib.load()
    n = m = args.n
    r0, r1 = shard_rows(n, world, rank)
    nloc = r1 - r0
    K_steps, W = max(1, args.steps), max(3, args.warmup)

    # ---- synthetic inputs: this rank's row block of M, built on the device in row chunks ----
    xs, xt = point_clouds(n, m)
    xs_d, xt_d = torch.from_numpy(xs[r0:r1]).to(dev), torch.from_numpy(xt).to(dev)
    Mdev = torch.empty((nloc, m), dtype=torch.float32, device=dev)
    chunk = max(1, (512 << 20) // (4 * m))
    mx = torch.zeros((), dtype=torch.float32, device=dev)
    scratch = torch.empty(lib.b2ot_dist_scratch_bytes(min(chunk, nloc), m, _lib.F32), dtype=torch.uint8, device=dev)
    for c0 in range(0, nloc, chunk):
        c1 = min(nloc, c0 + chunk)
        blk = Mdev[c0:c1]
        xb = xs_d[c0:c1]
        # squared-euclidean cost block straight into M (the library's own ot.dist kernel, clamped at 0)
        _lib.check(lib.b2ot_dist(_lib.F32, xb.data_ptr(), xt_d.data_ptr(), c1 - c0, m, xb.shape[1], xb.stride(0),
                                 xt_d.stride(0), _lib.METRIC_SQEUCLIDEAN, blk.data_ptr(), blk.stride(0),
                                 scratch.data_ptr(), _dense.stream_ptr(dev)), "b2ot_dist")
        mx = torch.maximum(mx, blk.max())
    if world > 1:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    for c0 in range(0, nloc, chunk):
        Mdev[c0:min(nloc, c0 + chunk)].div_(mx)
    a_loc = torch.full((nloc,), 1.0 / n, dtype=torch.float32, device=dev)
    b = torch.full((m,), 1.0 / m, dtype=torch.float32, device=dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # pinned host copy of M for the end-to-end leg (made before M is turned into K in place)
    M_host = None
    if not args.no_e2e:
        M_host = torch.empty((nloc, m), dtype=torch.float32, pin_memory=True)
        M_host.copy_(Mdev)
        torch.cuda.synchronize(dev)

    # ---- device-resident iterations/s ----
    _lib.check(lib.b2ot_kexp(Mdev.data_ptr(), Mdev.data_ptr(), nloc, m, Mdev.stride(0), Mdev.stride(0), REG,
                             _lib.F32, _dense.stream_ptr(dev)), "b2ot_kexp")
    Kmat = Mdev
    solver = _dense.DenseSolver("knopp", Kmat, a_loc, b, None, max_iter=10 ** 7, check_every=10, stop_thr=0.0,
                                group=group, exchange=args.exchange)
    u0 = torch.full((nloc,), 1.0 / n, dtype=torch.float32, device=dev)
    v0 = torch.full((m,), 1.0 / m, dtype=torch.float32, device=dev)
    solver.init(u0, v0)
    solver.iterate(W)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = lib.b2ot_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    solver.iterate(K_steps)
    ev1.record()
    barrier()
    launches = lib.b2ot_launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    t_steps = max_over_ranks(ev0.elapsed_time(ev1) * 1e-3)
    status, it_done = solver.poll()
    assert status == 0 and it_done == W + K_steps, (status, it_done)

    # ---- roofline of the dominant kernel (the fused sweep), timed alone with CUDA events ----
    sweep_ms, iter_ms = solver.profile(min(20, K_steps))
    sweep_ms = max_over_ranks(sweep_ms)
    u, v, errs = solver.export()
    finite = bool(torch.isfinite(u).all().item() and torch.isfinite(v).all().item())
    info = solver.info()
    solver.close()
    del solver, Kmat, Mdev
    torch.cuda.empty_cache()
    alg_bytes = 4.0 * nloc * m + 4.0 * (2 * nloc + 3 * m)
    peak, peak_src = peaks()
    achieved = alg_bytes / (sweep_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(f"knopp_f32_{nloc}x{m}")
    except Exception:
        pass

    # ---- e2e: whole solve from pinned host memory ----
    e2e = None
    if not args.no_e2e:
        a_host = torch.full((nloc,), 1.0 / n, dtype=torch.float32).pin_memory()
        b_host = torch.full((m,), 1.0 / m, dtype=torch.float32).pin_memory()
        torch.cuda.synchronize(dev)
        sinkhorn_from_host(a_host, b_host, M_host, REG, numItermax=min(W, 5), stopThr=0.0, device=dev, group=group,
                           n_total=n, exchange=args.exchange)  # warm-up solve (allocator, NCCL channels, kernel images)
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        res = sinkhorn_from_host(a_host, b_host, M_host, REG, numItermax=K_steps, stopThr=0.0, device=dev,
                                 group=group, n_total=n, exchange=args.exchange)
        ev1.record()
        barrier()
        t_e2e = max_over_ranks(ev0.elapsed_time(ev1) * 1e-3)
        assert res["niter"] == K_steps - 1, res["niter"]
        e2e = {"value": K_steps / t_e2e, "unit": "it/s",
               "h2d_bytes_per_step": int(res["h2d_bytes"] * world / K_steps),
               "d2h_bytes_per_step": int(res["d2h_bytes"] * world / K_steps),
               "solve_seconds": t_e2e, "iterations_per_solve": K_steps,
               "note": "one solve of `steps` iterations from pinned host M; bytes = whole-job bytes of the solve / steps"}
        del M_host

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        rows = min(n, 4096)
        its, per_iter, cores = cpu_knopp_sample(n, m, rows, 30, 3)
        cpu = {"value": its, "unit": "it/s", "cores": cores, "kind": "port",
               "sample": f"oracle numpy sinkhorn_knopp on the first {rows} of {n} rows (all {m} columns), 30 iterations, "
                         f"{per_iter * 1e3:.2f} ms per sample iteration, scaled by {rows}/{n}"}

    if rank == 0:
        line = {
            "metric": "sinkhorn_knopp_iterations_per_sec", "value": K_steps / t_steps, "unit": "it/s",
            "n_gpus": world, "steps": K_steps, "warmup": W, "ms_per_step": 1e3 * t_steps / K_steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"ot.bregman.sinkhorn_knopp dense fp32 n=m={n} (K = {4 * n * m / 2 ** 30:.0f} GiB), "
                                   f"reg={REG}, uniform weights, 2-D Gaussian clouds (seed {SEED}), fixed iterations "
                                   "(stopThr=0)",
                       "sharding": f"rows split over {world} rank(s), one exchange of K^T u (m+1 floats) per iteration, "
                                   f"exchange={args.exchange}"
                                   if world > 1 else "single GPU",
                       "l2": "inputs larger than L2 (K per rank = %.1f GiB >> 126 MB): no flush needed"
                             % (4 * nloc * m / 2 ** 30),
                       "sweep": info, "finite": finite, "last_err": errs[-1] if errs else None},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "kernel": "fused_sweep_kernel<float, 8, knopp>",
                         "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": sweep_ms, "peak_source": peak_src},
            "cpu_baseline": cpu,
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        reference_arm(a)
    else:
        ours(a)
