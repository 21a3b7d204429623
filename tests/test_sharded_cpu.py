"""world_size-2 (gloo, CPU) tests of the row-sharded path: row partition helper, the per-iteration
exchange protocol and the replicated stop / guard decision, checked against the unsharded oracle."""
import os
import socket
import warnings

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import sharded_protocol as proto
from oracle import sinkhorn_oracle as orc
from oracle import solvers_oracle as sorc
from pot_b200.sharded import shard_rows


def test_shard_rows_partition():
    for n in (1, 7, 8, 65536, 65537, 1000):
        for w in (1, 2, 3, 4, 8):
            blocks = [shard_rows(n, w, r) for r in range(w)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(w - 1))
            sizes = [b1 - b0 for b0, b1 in blocks]
            assert max(sizes) - min(sizes) <= 1
    assert shard_rows(65536, 8, 3) == (3 * 8192, 4 * 8192)


def _problem(n, m, dtype, reg_scale=1.0):
    rng = np.random.RandomState(11)
    xs, xt = rng.randn(n, 2), rng.randn(m, 2) + 1.0
    M = orc.sqeuclidean(xs, xt)
    M = (M / M.max() * reg_scale).astype(dtype)
    a = rng.rand(n) + 0.1
    b = rng.rand(m) + 0.1
    return (a / a.sum()).astype(dtype), (b / b.sum()).astype(dtype), M


UNBAL = (("fixed", 0.1, (1.0, 1.0), 25, 0.0), ("pair", 0.1, (0.5, 2.0), 2000, 1e-9),
         ("semi", 0.1, (float("inf"), 1.0), 2000, 1e-9), ("guard", 0.01, (1.0, 1.0), 20, 0.0))


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)

    def allreduce(x, op):
        t = torch.from_numpy(x)
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)

    res = {}
    n, m = 37, 29
    a, b, M = _problem(n, m, np.float64)
    r0, r1 = shard_rows(n, world, rank)
    for name, reg, its, thr in (("fixed", 0.1, 25, 0.0), ("conv", 0.1, 2000, 1e-9), ("guard", 1e-3, 50, 0.0)):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            Mc = M * 50 if name == "guard" else M   # unnormalised costs: K underflows, the guard fires
            u, v, errs, niter, st = proto.knopp_sharded(a[r0:r1], b, Mc[r0:r1], reg, n, its, thr, allreduce)
        res["knopp_" + name] = (u, v, errs, niter, st)
    for name, reg, its, thr in (("fixed", 0.05, 25, 0.0), ("conv", 0.1, 2000, 1e-9)):
        u, v, errs, niter, st = proto.log_sharded(a[r0:r1], b, M[r0:r1], reg, its, thr, allreduce)
        res["log_" + name] = (u, v, errs, niter, st)
    for name, reg, reg_m, its, thr in UNBAL:
        fi = [rm / (rm + reg) if rm != float("inf") else 1 for rm in reg_m]
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            Mc = M.copy()
            if name == "guard":
                Mc[:3] *= 1e7   # K underflows on rows of rank 0 only: both ranks must stop together
            Kl = np.exp(-Mc[r0:r1] / reg) * a[r0:r1, None] * b[None, :]
            res["unbal_" + name] = proto.unbalanced_sharded(a[r0:r1], b, Kl, fi[0], fi[1], its, thr, allreduce)
    out.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_protocol_matches_unsharded_oracle():
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n, m = 37, 29
    a, b, M = _problem(n, m, np.float64)
    for name, reg, its, thr in (("fixed", 0.1, 25, 0.0), ("conv", 0.1, 2000, 1e-9), ("guard", 1e-3, 50, 0.0)):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            Mc = M * 50 if name == "guard" else M
            _, lg = orc.sinkhorn_knopp(a, b, Mc, reg, numItermax=its, stopThr=thr, log=True, warn=False)
        u = np.concatenate([got[r]["knopp_" + name][0] for r in range(world)])
        for r in range(world):
            _, v, errs, niter, st = got[r]["knopp_" + name]
            assert niter == lg["niter"], name
            np.testing.assert_allclose(v, lg["v"], rtol=1e-12)
            np.testing.assert_allclose(errs, lg["err"], rtol=1e-6, atol=1e-16)
            assert st == {"fixed": 3, "conv": 1, "guard": 2}[name]
        np.testing.assert_allclose(u, lg["u"], rtol=1e-12)
    for name, reg, its, thr in (("fixed", 0.05, 25, 0.0), ("conv", 0.1, 2000, 1e-9)):
        _, lg = orc.sinkhorn_log(a, b, M, reg, numItermax=its, stopThr=thr, log=True, warn=False)
        u = np.concatenate([got[r]["log_" + name][0] for r in range(world)])
        for r in range(world):
            _, v, errs, niter, st = got[r]["log_" + name]
            assert niter == lg["niter"], name
            np.testing.assert_allclose(v, lg["log_v"], rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(u, lg["log_u"], rtol=1e-10, atol=1e-12)
    for name, reg, reg_m, its, thr in UNBAL:
        Mc = M.copy()
        if name == "guard":
            Mc[:3] *= 1e7
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            Gr, lg = sorc.sinkhorn_knopp_unbalanced(a, b, Mc, reg, reg_m, numItermax=its, stopThr=thr, log=True)
        u = np.concatenate([got[r]["unbal_" + name][0] for r in range(world)])
        for r in range(world):
            _, v, errs, niter, st = got[r]["unbal_" + name]
            assert st == {"fixed": 3, "pair": 1, "semi": 1, "guard": 2}[name], name
            assert len(errs) == len(lg["err"]), name
            np.testing.assert_allclose(errs, lg["err"], rtol=1e-6, atol=1e-16)
            np.testing.assert_allclose(np.log(v + 1e-300), lg["logv"], rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(np.log(u + 1e-300), lg["logu"], rtol=1e-10, atol=1e-12)
