"""Worker of the multi-GPU parity tests; launched by tests/test_multigpu.py (or by hand) as

    python -m torch.distributed.run --nnodes=1 --nproc-per-node R --master-addr 127.0.0.1 --master-port P \
        tests/mgpu_worker.py

Every rank checks its own results and exits non-zero on a mismatch.  Checks (exchange = peer, peer_ll (tagged words) and nccl):
  1. the stand-alone peer all-reduce against torch.distributed (sum / max, fp32 / fp64, 300 back-to-back calls);
  2. row-sharded sinkhorn_knopp / sinkhorn_log against the CPU oracle (fp64 1e-10, fp32 1e-4) and against the
     single-GPU solve (same niter, same err history), converged and fixed-iteration runs, ragged row split;
  3. source-sharded empirical Sinkhorn against the oracle;
  4. row-sharded ot.solve (plan rows, value, value_linear, potentials) against the solvers oracle;
  5. problem-sharded batched solver against the oracle;
  6. row-sharded sinkhorn_knopp_unbalanced (scalar / pair / semi-relaxed reg_m, kl / entropy, fixed count, converged,
     numerical guard) against the solvers oracle and against the single-device solve.
MGPU_ONLY=<section number> runs one section (after the all-reduce check).
"""
import os
import sys
import warnings

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import pot_b200 as ot  # noqa: E402
from oracle import sinkhorn_oracle as orc  # noqa: E402
from pot_b200.comm import get_comm  # noqa: E402
from oracle import solvers_oracle as sorc  # noqa: E402
from pot_b200.sharded import (empirical_sinkhorn_sharded, shard_rows, sinkhorn_knopp_sharded,  # noqa: E402
                              sinkhorn_knopp_unbalanced_sharded, sinkhorn_log_sharded, solve_batch_sharded,
                              solve_sharded)


def unbalanced_section(rank, world, dev, group, check, long_its):
    """6. row-sharded unbalanced scaling iteration: two exchanges per iteration (column sums + guard flag: sum; the u
    statistics of the error: max)."""
    rng = np.random.RandomState(6)
    n, m = 203, 160
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M64 = orc.sqeuclidean(xs, xt)
    M64 /= M64.max()
    a64 = rng.rand(n) + 0.5
    a64 /= a64.sum()
    b64 = rng.rand(m) + 0.5
    b64 /= 0.8 * b64.sum()  # different total masses
    r0, r1 = shard_rows(n, world, rank)
    cases = [(1.0, "kl"), ((0.5, 2.0), "kl"), ((float("inf"), 1.0), "kl"), (1.0, "entropy")]
    for exchange in ("peer", "nccl"):
        for dtype, tol in ((np.float64, 1e-10), (np.float32, 1e-4)):
            M, a, b = M64.astype(dtype), a64.astype(dtype), b64.astype(dtype)
            Ml, al, bd = (torch.as_tensor(M[r0:r1], device=dev), torch.as_tensor(a[r0:r1], device=dev),
                          torch.as_tensor(b, device=dev))
            for reg_m, rt in cases:
                for kw in (dict(numItermax=60, stopThr=0.0),
                           dict(numItermax=long_its, stopThr=1e-9 if dtype == np.float64 else 1e-5)):
                    with warnings.catch_warnings():
                        warnings.simplefilter("ignore")
                        G, lg = sinkhorn_knopp_unbalanced_sharded(al, bd, Ml, 0.1, reg_m, reg_type=rt, log=True,
                                                                  group=group, exchange=exchange, **kw)
                        Gr, lr = sorc.sinkhorn_knopp_unbalanced(a, b, M, 0.1, reg_m, reg_type=rt, log=True, **kw)
                    tag = f"unbalanced {exchange} {dtype.__name__} reg_m={reg_m} {rt} {kw['numItermax']}"
                    if kw["stopThr"] == 0.0 or dtype == np.float64:
                        check(len(lg["err"]) == len(lr["err"]), f"{tag}: {len(lg['err'])} vs {len(lr['err'])} iterations")
                    Gn = G.cpu().numpy().astype(np.float64)
                    ref = Gr[r0:r1].astype(np.float64)
                    err = np.abs(Gn - ref) / (np.abs(ref) + 1e-6 * Gr.max())
                    check(err.max() < tol, f"{tag}: plan rel err {err.max():.3e}")
                    scale = max(abs(float(lr["total_cost"])), abs(float(lr["cost"])))
                    check(abs(float(lg["cost"]) - float(lr["cost"])) <= tol * scale, f"{tag}: cost")
                    check(abs(float(lg["total_cost"]) - float(lr["total_cost"])) <= 10 * tol * scale,
                          f"{tag}: total_cost {float(lg['total_cost'])} vs {float(lr['total_cost'])}")
                    k = min(len(lg["err"]), len(lr["err"]))
                    eh = np.abs(np.array([float(e) for e in lg["err"][:k]]) - np.asarray(lr["err"][:k], dtype=np.float64))
                    check(eh.max() <= (1e-9 if dtype == np.float64 else 1e-4), f"{tag}: err history differs by {eh.max():.2e}")
                    v = lg["logv"].clone()
                    v0 = v.clone()
                    dist.broadcast(v0, 0)
                    check(torch.equal(v, v0), f"{tag}: replicated v differs across ranks")
            # the single-device solve of the same problem takes the same decisions (same number of iterations)
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                _, l1 = ot.unbalanced.sinkhorn_knopp_unbalanced(torch.as_tensor(a, device=dev), bd,
                                                                torch.as_tensor(M, device=dev), 0.1, 1.0, log=True,
                                                                numItermax=long_its, stopThr=1e-9 if dtype == np.float64 else 1e-5)
                _, l2 = sinkhorn_knopp_unbalanced_sharded(al, bd, Ml, 0.1, 1.0, log=True, group=group, exchange=exchange,
                                                          numItermax=long_its, stopThr=1e-9 if dtype == np.float64 else 1e-5)
            if dtype == np.float64:
                check(len(l1["err"]) == len(l2["err"]), f"unbalanced {exchange}: sharded {len(l2['err'])} vs single {len(l1['err'])} iterations")
        # numerical guard: K underflows on the rows of ONE rank only -> every rank must stop at the same iteration
        Mbad = M64.copy()
        Mbad[:5] *= 1e7
        with warnings.catch_warnings(record=True) as wr:
            warnings.simplefilter("always")
            _, lr = sorc.sinkhorn_knopp_unbalanced(a64, b64, Mbad, 0.01, 1.0, log=True, numItermax=20, stopThr=0.0)
        with warnings.catch_warnings(record=True) as wg:
            warnings.simplefilter("always")
            _, lg = sinkhorn_knopp_unbalanced_sharded(torch.as_tensor(a64[r0:r1], device=dev), torch.as_tensor(b64, device=dev),
                                                      torch.as_tensor(Mbad[r0:r1], device=dev), 0.01, 1.0, log=True,
                                                      group=group, exchange=exchange, numItermax=20, stopThr=0.0)
        ref_warn = [str(w.message) for w in wr if "Numerical errors" in str(w.message)]
        got_warn = [str(w.message) for w in wg if "Numerical errors" in str(w.message)]
        check(ref_warn == got_warn, f"unbalanced {exchange} guard: {got_warn} vs {ref_warn}")
        check(len(lg["err"]) == len(lr["err"]), f"unbalanced {exchange} guard: err history length")
        if rank == 0:
            print(f"unbalanced sharded ok ({exchange}; guard warnings: {len(ref_warn)})", flush=True)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    one_device = os.environ.get("MGPU_ONE_DEVICE", "0") == "1"
    if one_device:
        # every rank on cuda:0 (a box with ONE GPU): NCCL refuses two ranks per device, so the process group is gloo
        # (host plumbing + the "nccl"-hook baseline exchange staged through the host); the peer exchange maps the
        # other process's buffer with CUDA IPC on the same device and runs exactly the multi-GPU kernels.
        local = 0
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if one_device:
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=dev)
    group = dist.group.WORLD
    long_its = 300 if one_device else 2000  # two processes time-slice one GPU: every exchange costs a context switch
    ar_reps = 40 if one_device else 300

    def check(cond, msg):
        if not cond:
            print(f"[rank {rank}] FAIL: {msg}", flush=True)
            sys.exit(1)

    # ---- 1. stand-alone peer all-reduce ----
    comm = get_comm(group, dev, 1 << 20)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    for dt in (torch.float32, torch.float64):
        for it in range(ar_reps):
            n = 1 + (it * 37) % 5000
            x = torch.randn(n, dtype=dt, device=dev, generator=g)
            ref = x.clone()
            op = "max" if it % 3 == 0 else "sum"
            dist.all_reduce(ref, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
            comm.all_reduce(x, op)
            tol = 0 if op == "max" else (1e-5 if dt == torch.float32 else 1e-13)
            check(torch.allclose(x, ref, rtol=tol, atol=tol), f"peer all-reduce {op} {dt} it {it}")
    # results must be bit-identical on every rank
    x = torch.randn(4097, dtype=torch.float32, device=dev, generator=g)
    comm.all_reduce(x, "sum")
    y = x.clone()
    dist.broadcast(y, 0)
    check(torch.equal(x, y), "peer all-reduce is not bit-identical across ranks")
    if rank == 0:
        print("peer all-reduce ok", flush=True)

    only = os.environ.get("MGPU_ONLY", "")
    if only == "6":
        unbalanced_section(rank, world, dev, group, check, long_its)
        dist.barrier()
        if rank == 0:
            print("MGPU_OK", flush=True)
        dist.destroy_process_group()
        return

    # ---- 2. dense row-sharded solves ----
    rng = np.random.RandomState(0)
    n, m = 203, 160  # 203 rows: ragged split
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M64 = orc.sqeuclidean(xs, xt)
    M64 /= M64.max()
    a64 = rng.rand(n) + 0.5
    a64 /= a64.sum()
    b64 = rng.rand(m) + 0.5
    b64 /= b64.sum()
    r0, r1 = shard_rows(n, world, rank)
    for exchange in ("peer", "peer_ll", "nccl"):
        for dtype, tol in ((np.float64, 1e-10), (np.float32, 1e-4)):
            M, a, b = M64.astype(dtype), a64.astype(dtype), b64.astype(dtype)
            Ml, al, bd = (torch.as_tensor(M[r0:r1], device=dev), torch.as_tensor(a[r0:r1], device=dev),
                          torch.as_tensor(b, device=dev))
            for name, fn, ofn, key in (("knopp", sinkhorn_knopp_sharded, orc.sinkhorn_knopp, "u"),
                                       ("log", sinkhorn_log_sharded, orc.sinkhorn_log, "log_u")):
                for kw in (dict(numItermax=60, stopThr=0.0), dict(numItermax=long_its, stopThr=1e-7 if dtype == np.float64 else 1e-5)):
                    with warnings.catch_warnings():
                        warnings.simplefilter("ignore")
                        G, lg = fn(al, bd, Ml, 0.1, log=True, group=group, n_total=n, exchange=exchange, **kw)
                        Gr, lgr = ofn(a, b, M, 0.1, log=True, **kw)
                    check(lg["niter"] == lgr["niter"], f"{exchange} {name} {dtype.__name__} niter {lg['niter']} vs {lgr['niter']}")
                    Gn = G.cpu().numpy().astype(np.float64)
                    ref = Gr[r0:r1].astype(np.float64)
                    err = np.abs(Gn - ref) / (np.abs(ref) + 1e-6 * Gr.max())
                    check(err.max() < tol, f"{exchange} {name} {dtype.__name__} plan rel err {err.max():.3e}")
                    check(len(lg["err"]) == len(lgr["err"]), f"{exchange} {name} err history length")
                    # every rank holds the same replicated v and took the same decisions
                    v = lg["v" if name == "knopp" else "log_v"].clone()
                    v0 = v.clone()
                    dist.broadcast(v0, 0)
                    check(torch.equal(v, v0), f"{exchange} {name} replicated v differs across ranks")
        if rank == 0:
            print(f"dense sharded ok ({exchange})", flush=True)

    # ---- 2b. slot reuse of the peer exchange across stops and new solves on the cached communicator ----
    # A converged solve leaves no-op iterations enqueued behind its stop; the next solve's first put follows at once.
    # Every sequence number is still exchanged (flag-only puts, un-gated waits: csrc/comm.cuh), so back-to-back solves
    # must reproduce the first result bit for bit, on every rank, however the two ranks drift.
    M, a, b = M64.astype(np.float32), a64.astype(np.float32), b64.astype(np.float32)
    Ml, al, bd = (torch.as_tensor(M[r0:r1], device=dev), torch.as_tensor(a[r0:r1], device=dev),
                  torch.as_tensor(b, device=dev))
    first = {}
    for cycle in range(12 if one_device else 40):
        for name, fn in (("knopp", sinkhorn_knopp_sharded), ("log", sinkhorn_log_sharded)):
            kw = dict(numItermax=400, stopThr=1e-4) if cycle % 2 == 0 else dict(numItermax=7 + cycle % 5, stopThr=0.0)
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                G, lg = fn(al, bd, Ml, 0.1, log=True, group=group, n_total=n,
                           exchange="peer" if (cycle // 2) % 2 == 0 else "peer_ll", **kw)  # both protocols share the slots
            key = (name, cycle % 2, kw["numItermax"])
            if key not in first:
                first[key] = (G.clone(), lg["niter"])
            else:
                check(lg["niter"] == first[key][1] and torch.equal(G, first[key][0]),
                      f"back-to-back sharded solves differ (cycle {cycle}, {name})")
    if rank == 0:
        print("peer exchange slot reuse ok", flush=True)

    # ---- 4. row-sharded ot.solve: plan rows + all-reduced objective terms ----
    for dtype, tol in ((np.float64, 1e-10), (np.float32, 1e-4)):
        M, a, b = M64.astype(dtype), a64.astype(dtype), b64.astype(dtype)
        for rt in ("KL", "entropy"):
            ref = sorc.solve(M, a, b, reg=0.1, reg_type=rt, max_iter=50, tol=0)
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                res = solve_sharded(torch.as_tensor(M[r0:r1], device=dev), torch.as_tensor(a[r0:r1], device=dev),
                                    torch.as_tensor(b, device=dev), reg=0.1, reg_type=rt, max_iter=50, tol=0,
                                    group=group, n_total=n)
            Gn = res.plan.cpu().numpy().astype(np.float64)
            pr = ref["plan"][r0:r1].astype(np.float64)
            err = np.abs(Gn - pr) / (np.abs(pr) + 1e-6 * ref["plan"].max())
            check(err.max() < tol, f"solve_sharded {rt} {dtype.__name__} plan rel err {err.max():.3e}")
            scale = max(abs(float(ref["value"])), abs(float(ref["value_linear"])))
            check(abs(float(res.value_linear) - float(ref["value_linear"])) <= tol * scale, f"solve_sharded {rt} value_linear")
            check(abs(float(res.value) - float(ref["value"])) <= tol * scale,
                  f"solve_sharded {rt} {dtype.__name__} value {float(res.value)} vs {float(ref['value'])}")
            eg = np.abs(res.potentials[1].cpu().numpy() - ref["potentials"][1]).max() / max(1.0, np.abs(ref["potentials"][1]).max())
            check(eg < tol, f"solve_sharded {rt} potentials err {eg:.2e}")
    if rank == 0:
        print("solve sharded ok", flush=True)

    # ---- 3. empirical, source points sharded ----
    n, m, d = 301, 257, 5
    xs = (rng.randn(n, d) / np.sqrt(d))
    xt = (rng.randn(m, d) / np.sqrt(d) + 0.3)
    r0, r1 = shard_rows(n, world, rank)
    for exchange in ("peer", "nccl"):
        for dtype, tol in ((np.float64, 1e-10), (np.float32, 1e-4)):
            X, Y = xs.astype(dtype), xt.astype(dtype)
            for kw in (dict(numIterMax=30, stopThr=-1.0), dict(numIterMax=long_its if one_device else 3000, stopThr=1e-8 if dtype == np.float64 else 1e-4)):
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")
                    f, gg, lg = empirical_sinkhorn_sharded(torch.as_tensor(X[r0:r1], device=dev), torch.as_tensor(Y, device=dev),
                                                           0.5, log=True, group=group, n_total=n, exchange=exchange, **kw)
                    fr, gr, lr = orc.empirical_sinkhorn_lazy(X, Y, 0.5, log=True, batchSize=64, **kw)
                check(lg["niter"] == lr["niter"], f"emp {exchange} {dtype.__name__} niter {lg['niter']} vs {lr['niter']}")
                ef = np.abs(f.cpu().numpy() - fr[r0:r1]).max() / max(1.0, np.abs(fr).max())
                eg = np.abs(gg.cpu().numpy() - gr).max() / max(1.0, np.abs(gr).max())
                check(ef < tol and eg < tol, f"emp {exchange} {dtype.__name__} potentials err {ef:.2e} {eg:.2e}")
        if rank == 0:
            print(f"empirical sharded ok ({exchange})", flush=True)

    # ---- 5. batched solver, whole problems sharded (no communication in the iterations) ----
    B, bn, bm = 11, 40, 33  # 11 problems: ragged split
    X, Y = rng.randn(B, bn, 2), rng.randn(B, bm, 2) + 1.0
    Mb = orc.dist_sqeuclidean_batch(X, Y)
    Mb /= Mb.max()
    # problem-dependent conditioning, so that the shards converge at different iterations
    Mb *= np.linspace(0.5, 2.0, B)[:, None, None]
    p0, p1 = shard_rows(B, world, rank)
    for dtype, tol in ((np.float64, 1e-10), (np.float32, 1e-4)):
        Md = Mb.astype(dtype)
        for solver in ("log_sinkhorn", "sinkhorn"):
            for rt in ("entropy", "kl"):
                for kw in (dict(max_iter=40, tol=0.0), dict(max_iter=500, tol=1e-6 if dtype == np.float64 else 1e-4)):
                    ref = orc.solve_batch(Md, 0.1, solver=solver, reg_type=rt, **kw)
                    res = solve_batch_sharded(torch.as_tensor(Md[p0:p1], device=dev), 0.1, solver=solver, reg_type=rt,
                                              group=group, **kw)
                    check(res.log["n_iter"] == ref["n_iter"],
                          f"batch {solver} {rt} {dtype.__name__} n_iter {res.log['n_iter']} vs {ref['n_iter']} ({kw})")
                    P, Pr = res.plan.cpu().numpy().astype(np.float64), ref["plan"][p0:p1].astype(np.float64)
                    err = np.abs(P - Pr) / (np.abs(Pr) + 1e-6 * ref["plan"].max())
                    check(err.max() < tol, f"batch {solver} {rt} {dtype.__name__} plan rel err {err.max():.3e}")
                    for key, got in (("value", res.value), ("value_linear", res.value_linear)):
                        rv = ref[key][p0:p1]
                        ev = np.abs(got.cpu().numpy() - rv).max() / np.abs(ref[key]).max()
                        check(ev < tol, f"batch {solver} {rt} {dtype.__name__} {key} err {ev:.2e}")
    if rank == 0:
        print("batch sharded ok", flush=True)

    unbalanced_section(rank, world, dev, group, check, long_its)

    dist.barrier()
    if rank == 0:
        print("MGPU_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
