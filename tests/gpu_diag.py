"""GPU bring-up diagnostics (run under gpurun): parity of every sweep variant against the CPU oracle
on a grid of shapes, plus a quick throughput probe.  Not part of the pytest suite.

    python tests/gpu_diag.py --group split|fused1|cluster|perf [--quick]
"""

import argparse
import os
import sys
import time
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import pot_b200 as ot  # noqa: E402
from oracle import sinkhorn_oracle as orc  # noqa: E402
from pot_b200 import _dense, _lib  # noqa: E402


def problem(n, m, dtype, seed=0, reg=0.1):
    rng = np.random.RandomState(seed)
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = orc.sqeuclidean(xs, xt)
    M = (M / M.max()).astype(dtype)
    a = rng.rand(n) + 0.1
    b = rng.rand(m) + 0.1
    return (a / a.sum()).astype(dtype), (b / b.sum()).astype(dtype), M, reg


def relerr(x, ref):
    x = np.asarray(x, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.max(np.abs(x - ref) / (np.abs(ref) + 1e-300)))


def run_case(mode, n, m, dtype, variant, its=30, reg=0.1):
    a, b, M, reg = problem(n, m, dtype, reg=reg)
    dev = torch.device("cuda:0")
    at, bt, Mt = (torch.as_tensor(x, device=dev) for x in (a, b, M))
    fn = ot.sinkhorn_knopp if mode == "knopp" else ot.sinkhorn_log
    ofn = orc.sinkhorn_knopp if mode == "knopp" else orc.sinkhorn_log
    t0 = time.time()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        G, lg = fn(at, bt, Mt, reg, numItermax=its, stopThr=0, log=True, b2ot_variant=variant)
        torch.cuda.synchronize()
        t1 = time.time()
        Gr, lgr = ofn(a, b, M, reg, numItermax=its, stopThr=0, log=True)
    ku, kv = ("u", "v") if mode == "knopp" else ("log_u", "log_v")
    if mode == "knopp":
        eu, ev = relerr(lg[ku].cpu().numpy(), lgr[ku]), relerr(lg[kv].cpu().numpy(), lgr[kv])
    else:  # potentials: error relative to the potential range
        sc = max(1.0, float(np.max(np.abs(lgr[ku]))))
        eu = float(np.max(np.abs(lg[ku].cpu().numpy() - lgr[ku]))) / sc
        ev = float(np.max(np.abs(lg[kv].cpu().numpy() - lgr[kv]))) / sc
    Gn = G.cpu().numpy().astype(np.float64)
    eg = float(np.max(np.abs(Gn - Gr) / (np.abs(Gr) + 1e-6 * Gr.max())))
    ee = relerr([float(e) for e in lg["err"]], [float(e) for e in lgr["err"]]) if lgr["err"] else 0.0
    tol = 1e-10 if dtype == np.float64 else 1e-4
    ok = eu < tol and ev < tol and eg < tol and lg["niter"] == lgr["niter"]
    print(f"{'OK ' if ok else 'BAD'} {mode:5s} {np.dtype(dtype).name:7s} n={n:6d} m={m:6d} {variant:5s} "
          f"u {eu:.2e} v {ev:.2e} plan {eg:.2e} err-hist {ee:.2e} niter {lg['niter']}/{lgr['niter']} "
          f"({t1 - t0:.2f}s)", flush=True)
    return ok


def info_line(mode, n, m, dtype):
    dev = torch.device("cuda:0")
    K = torch.rand(n, m, device=dev, dtype=dtype)
    a = torch.full((n,), 1.0 / n, device=dev, dtype=dtype)
    b = torch.full((m,), 1.0 / m, device=dev, dtype=dtype)
    s = _dense.DenseSolver(mode, K, a, b, torch.log(b), max_iter=10)
    print("  config", mode, n, m, dtype, s.info(), flush=True)
    s.close()


def perf_case(mode, n, m, dtype, iters, variant="auto", label=""):
    """Device-resident throughput: K random in (e^-3, 1], fixed iteration count, CUDA events."""
    dev = torch.device("cuda:0")
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    g = torch.Generator(device=dev).manual_seed(1)
    K = torch.empty(n, m, device=dev, dtype=tdt)
    K.uniform_(0.0, 3.0, generator=g)
    if mode == "knopp":
        K.neg_().exp_()
    else:
        K.neg_()
    a = torch.full((n,), 1.0 / n, device=dev, dtype=tdt)
    b = torch.full((m,), 1.0 / m, device=dev, dtype=tdt)
    if mode == "knopp":
        s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=10 ** 6, stop_thr=0.0, variant=variant)
        u0, v0 = a.clone(), b.clone()
    else:
        s = _dense.DenseSolver("log", K, torch.log(a), b, torch.log(b), max_iter=10 ** 6, stop_thr=0.0,
                               variant=variant)
        u0, v0 = torch.zeros_like(a), torch.zeros_like(b)
    s.init(u0, v0)
    s.iterate(5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s.iterate(iters)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    st, it = s.poll()
    esz = K.element_size()
    gbs = (esz * n * m + esz * (2 * n + 3 * m)) / (ms * 1e-3) / 1e9
    print(f"PERF {label}{mode:5s} {np.dtype(dtype).name} n={n} m={m} {s.info()} {ms * 1e3:9.1f} us/iter "
          f"{1e3 / ms:9.1f} it/s {gbs:8.1f} GB/s algorithmic  status={st} iter={it}", flush=True)
    u, v, errs = s.export()
    print("     finite:", bool(torch.isfinite(u).all()), bool(torch.isfinite(v).all()),
          "last err", errs[-1] if errs else None, flush=True)
    s.close()
    del K
    torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--group", required=True)
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    f32, f64 = np.float32, np.float64
    print("device:", torch.cuda.get_device_name(0), "lib", _lib.load().b2ot_version(), flush=True)
    bad = 0
    if args.group == "split":
        for mode in ("knopp", "log"):
            for dt in (f64, f32):
                for (n, m) in ((50, 60), (33, 47), (1, 5), (7, 1), (257, 1031), (1000, 1000)):
                    bad += not run_case(mode, n, m, dt, "split")
    elif args.group == "fused1":
        for mode in ("knopp", "log"):
            for dt in (f64, f32):
                for (n, m) in ((50, 60), (128, 96), (3, 4), (1, 8), (257, 1032), (1000, 1000), (999, 2048),
                               (130, 4100), (77, 8192)):
                    if dt == f64 and m > 4096:
                        continue
                    info_line(mode, n, m, torch.float64 if dt == f64 else torch.float32)
                    bad += not run_case(mode, n, m, dt, "fused")
    elif args.group == "cluster":
        for mode in ("knopp", "log"):
            for dt in (f32, f64):
                shapes = ((64, 16384), (37, 8196), (50, 20000), (40, 32768), (33, 65536), (300, 40000))
                for (n, m) in shapes:
                    if dt == f64 and m > 32768:
                        continue
                    info_line(mode, n, m, torch.float64 if dt == f64 else torch.float32)
                    bad += not run_case(mode, n, m, dt, "fused", its=12)
    elif args.group == "perf":
        perf_case("knopp", 1000, 1000, f64, 500)
        perf_case("knopp", 4096, 4096, f32, 300)
        perf_case("knopp", 16384, 16384, f32, 100)
        perf_case("knopp", 16384, 16384, f32, 50, variant="split", label="split ")
        perf_case("log", 16384, 16384, f32, 100)
        perf_case("log", 16384, 16384, f32, 20, variant="split", label="split ")
        perf_case("knopp", 8192, 65536, f32, 50)
        if not args.quick:
            perf_case("knopp", 65536, 65536, f32, 30)
            perf_case("log", 65536, 65536, f32, 20)
    print("group", args.group, "bad cases:", bad, flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
