"""Host-buffer entry points: inputs live in (pinned) host memory, results come back to the host.

This is the end-to-end path bench.py times: the cost matrix is streamed host->device in row chunks on
a copy stream while the Gibbs-kernel transform (b2ot_kexp / b2ot_negscale, in place) of the previous
chunk runs on the compute stream; the solve then runs device-resident and only the scalings
(u, v), the error history and the status word are copied back.
"""

import numpy as np
import torch

from . import _dense, _lib
from .bregman import POLL_EVERY

__all__ = ["sinkhorn_from_host"]


def _as_host_tensor(x, dtype=None):
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    if dtype is not None and x.dtype != dtype:
        x = x.to(dtype)
    return x


def sinkhorn_from_host(a, b, M, reg, method="sinkhorn", numItermax=1000, stopThr=1e-9, device=None, group=None,
                       n_total=None, chunk_rows=None, warmstart=None, exchange="peer"):
    """Solve from host buffers.  a: (n_local,), b: (m,), M: (n_local, m) CPU tensors / numpy arrays (pinned memory
    makes the upload asynchronous).  Returns a dict of HOST tensors {"u", "v", "err", "niter", "status"} plus
    {"h2d_bytes", "d2h_bytes"}.  With `group`, M/a are this rank's row block (see pot_b200.sharded)."""
    mode = {"sinkhorn": "knopp", "sinkhorn_log": "log"}[method.lower()]
    M = _as_host_tensor(M)
    dt = M.dtype
    a, b = _as_host_tensor(a, dt), _as_host_tensor(b, dt)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    n, m = M.shape
    if n_total is None:
        n_total = n
        if group is not None:
            from .sharded import _total_rows
            n_total = _total_rows(n, dev, group)
    lib = _lib.load()
    code = _dense.dtype_code(dt)
    fn = lib.b2ot_kexp if mode == "knopp" else lib.b2ot_negscale
    with torch.cuda.device(dev):
        K = torch.empty((n, m), dtype=dt, device=dev)
        if chunk_rows is None:
            chunk_rows = max(1, min(n, (256 << 20) // max(1, m * M.element_size())))
        compute = torch.cuda.current_stream(dev)
        copy = torch.cuda.Stream(device=dev)
        copy.wait_stream(compute)
        events = []
        for r0 in range(0, n, chunk_rows):
            r1 = min(n, r0 + chunk_rows)
            with torch.cuda.stream(copy):
                K[r0:r1].copy_(M[r0:r1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy)
            compute.wait_event(ev)
            events.append(ev)
            _lib.check(fn(K[r0:r1].data_ptr(), K[r0:r1].data_ptr(), r1 - r0, m, K.stride(0), K.stride(0), float(reg),
                          code, _dense.stream_ptr(dev)), "cost transform")
        a_d = a.to(dev, non_blocking=True)
        b_d = b.to(dev, non_blocking=True)
        if mode == "knopp":
            if warmstart is None:
                u0 = torch.full((n,), 1.0 / n_total, dtype=dt, device=dev)
                v0 = torch.full((m,), 1.0 / m, dtype=dt, device=dev)
            else:
                u0 = torch.exp(_as_host_tensor(warmstart[0], dt).to(dev))
                v0 = torch.exp(_as_host_tensor(warmstart[1], dt).to(dev))
            av, logb = a_d, None
        else:
            if warmstart is None:
                u0 = torch.zeros(n, dtype=dt, device=dev)
                v0 = torch.zeros(m, dtype=dt, device=dev)
            else:
                u0 = _as_host_tensor(warmstart[0], dt).to(dev)
                v0 = _as_host_tensor(warmstart[1], dt).to(dev)
            av, logb = torch.log(a_d), torch.log(b_d)
        solver = _dense.DenseSolver(mode, K, av, b_d, logb, max_iter=int(numItermax), check_every=10,
                                    stop_thr=float(stopThr), group=group, exchange=exchange)
        status, niter, _ = solver.solve(u0, v0, poll_every=POLL_EVERY)
        u, v, errs = solver.export()
        solver.close()
        u_h, v_h = u.cpu(), v.cpu()  # synchronous device->host reads of the result
    esz = M.element_size()
    return {"u": u_h, "v": v_h, "err": errs, "niter": niter, "status": status,
            "h2d_bytes": esz * (n * m + n + m), "d2h_bytes": esz * (n + m) + 8 * len(errs) + 80}
