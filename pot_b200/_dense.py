"""Host-side wrapper of the dense Sinkhorn solver handle (b2ot_dense_* in include/b2ot.h).

Torch is used for device memory, streams and (row-sharded runs) the all-reduce plumbing;
every arithmetic step runs in libb2ot.so.
"""

import ctypes as C

import torch

from . import _lib

_DT = {torch.float32: _lib.F32, torch.float64: _lib.F64}


def dtype_code(t):
    try:
        return _DT[t]
    except KeyError:
        raise TypeError(f"pot_b200 supports float32 and float64 tensors, got {t}") from None


def stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def cost_transform(M, reg, op):
    """K = exp(M/(-reg)) (op='kexp', _sinkhorn.py:621) or Mr = -M/reg (op='negscale', :860)."""
    lib = _lib.load()
    out = torch.empty((M.shape[0], M.shape[1]), dtype=M.dtype, device=M.device)
    fn = lib.b2ot_kexp if op == "kexp" else lib.b2ot_negscale
    with torch.cuda.device(M.device):
        _lib.check(fn(M.data_ptr(), out.data_ptr(), M.shape[0], M.shape[1], M.stride(0), out.stride(0), float(reg),
                      dtype_code(M.dtype), stream_ptr(M.device)), "cost transform")
    return out


class DenseSolver:
    """One dense solve (or one row block of a row-sharded solve) on the current device.

    K     : (n_local, m) device tensor, exp(-M/reg) for mode 'knopp', -M/reg for mode 'log' (unit column stride)
    a     : (n_local,) a for 'knopp', log(a) for 'log'
    b     : (m,)
    logb  : (m,) log(b), 'log' mode only
    group : torch.distributed process group of the row-shard ranks (None = single device)
    exchange : "peer" (default; one-shot NVLink peer-memory exchange inside the library's kernels) or "nccl"
    """

    def __init__(self, mode, K, a, b, logb=None, max_iter=1000, check_every=10, stop_thr=1e-9, variant="auto",
                 group=None, exchange="peer"):
        self.lib = _lib.load()
        assert K.is_cuda and K.dim() == 2 and K.stride(1) == 1
        self.mode = _lib.MODE_KNOPP if mode == "knopp" else _lib.MODE_LOG
        self.K, self.a, self.b, self.logb = K, a, b, logb
        self.n, self.m = K.shape
        self.dtype = K.dtype
        self.device = K.device
        self.max_iter = int(max_iter)
        self.group = group
        code = dtype_code(K.dtype)
        nbytes = self.lib.b2ot_dense_workspace_bytes(self.n, self.m, code, self.max_iter, int(check_every))
        self.ws = torch.empty(nbytes, dtype=torch.uint8, device=K.device)
        var = {"auto": _lib.VARIANT_AUTO, "fused": _lib.VARIANT_FUSED, "split": _lib.VARIANT_SPLIT,
               "resident": _lib.VARIANT_RESIDENT}[variant]
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_dense_create(
                C.byref(h), self.mode, code, self.n, self.m, K.stride(0), K.data_ptr(), a.data_ptr(), b.data_ptr(),
                logb.data_ptr() if logb is not None else None, self.ws.data_ptr(), nbytes, self.max_iter,
                int(check_every), float(stop_thr), var, stream_ptr(self.device)), "b2ot_dense_create")
        self.h = h
        self._cb = None
        self.comm = None
        if group is not None:
            if exchange in ("peer", "peer_ll"):
                # NVLink peer-memory exchange.  "peer": put kernel + sequence flags, the finalize kernel is the take
                # (csrc/comm.cuh).  "peer_ll": slots with room for the tagged-word region - the column sums then travel
                # as {value, sequence} words inside the finalize kernel, no put kernel, no flags (b2ot_dense_set_comm
                # decides by slot size).  Measured equal on 2 and 8 GPUs (profiles/r02ll_exchange_ab.md), so the protocol
                # with two rounds of multi-GPU runs behind it stays the default.
                from .comm import get_comm

                self.comm = get_comm(group, self.device, (self.m + 1) * (24 if exchange == "peer_ll" else 8) + 128)
                _lib.check(self.lib.b2ot_dense_set_comm(self.h, self.comm.h), "b2ot_dense_set_comm")
                if exchange == "peer_ll":
                    _lib.check(self.lib.b2ot_dense_set_comm_ll(self.h, 1), "b2ot_dense_set_comm_ll")
            elif exchange == "nccl":  # baseline: one NCCL all-reduce per iteration through torch.distributed
                self._install_allreduce(group)
            else:
                raise ValueError(f"unknown exchange {exchange!r} (peer | peer_ll | nccl)")

    # ---- baseline exchange: the per-iteration column-sum all-reduce goes through torch.distributed ----
    def _install_allreduce(self, group):
        import torch.distributed as dist

        ws, base = self.ws, self.ws.data_ptr()
        esize = self.K.element_size()
        dtype = self.dtype

        def hook(buf, count, _dtype, op, _stream, _ctx):
            try:
                off = buf - base
                t = ws[off:off + count * esize].view(dtype)
                dist.all_reduce(t, op=dist.ReduceOp.MAX if op == _lib.OP_MAX else dist.ReduceOp.SUM, group=group)
                return 0
            except Exception:  # surfaced as an error code through the C ABI
                import traceback
                traceback.print_exc()
                return 1

        self._cb = _lib.ALLREDUCE_FN(hook)
        _lib.check(self.lib.b2ot_dense_set_allreduce(self.h, self._cb, None), "b2ot_dense_set_allreduce")

    def info(self):
        v = [C.c_int() for _ in range(5)]
        _lib.check(self.lib.b2ot_dense_info(self.h, *[C.byref(x) for x in v]), "b2ot_dense_info")
        names = ("variant", "cluster_size", "n_clusters", "stages", "vecs_per_thread")
        out = dict(zip(names, (x.value for x in v)))
        out["variant"] = {1: "fused", 2: "split", 3: "resident"}.get(out["variant"], "?")
        if out["variant"] == "resident":  # K in the grid's shared memory: CTAs and rows per CTA
            out = {"variant": "resident", "ctas": out["n_clusters"], "rows_per_cta": out["stages"]}
            return out
        if out["vecs_per_thread"] >= 100:  # fp32 log domain: two independent half-CTAs (dense_logpp_kernels.cuh)
            out["vecs_per_thread"] -= 100
            out["schedule"] = "ping-pong"
        return out

    def init(self, u0, v0):
        self._u0, self._v0 = u0, v0  # keep alive until the async copies ran
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_dense_init(self.h, u0.data_ptr(), v0.data_ptr()), "b2ot_dense_init")

    def iterate(self, k):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_dense_iterate(self.h, int(k)), "b2ot_dense_iterate")

    def profile(self, k):
        """k iterations with CUDA events around each sweep launch -> (sweep_ms_avg, iter_ms_avg)."""
        a, b = C.c_float(), C.c_float()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_dense_profile(self.h, int(k), C.byref(a), C.byref(b)), "b2ot_dense_profile")
        return a.value, b.value

    def poll(self):
        st, it = C.c_int(), C.c_int()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_dense_poll(self.h, C.byref(st), C.byref(it)), "b2ot_dense_poll")
        return st.value, it.value

    def solve(self, u0, v0, poll_every=0):
        """Run to completion.  Returns (status, niter, n_err)."""
        self._u0, self._v0 = u0, v0
        st, ni, ne = C.c_int(), C.c_int(), C.c_int()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_dense_solve(self.h, u0.data_ptr(), v0.data_ptr(), int(poll_every), C.byref(st),
                                                 C.byref(ni), C.byref(ne)), "b2ot_dense_solve")
        if st.value == _lib.ST_COMM_TIMEOUT:
            raise _lib.B2otError("row-sharded solve: a peer rank never published its exchange flag (timeout)")
        return st.value, ni.value, ne.value

    def export(self):
        """Returns (u, v, err_list_of_floats)."""
        u = torch.empty(self.n, dtype=self.dtype, device=self.device)
        v = torch.empty(self.m, dtype=self.dtype, device=self.device)
        cap = self.max_iter + 2
        errs = (C.c_double * cap)()
        ne = C.c_int()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_dense_export(self.h, u.data_ptr(), v.data_ptr(), errs, cap, C.byref(ne)),
                       "b2ot_dense_export")
        return u, v, [errs[i] for i in range(ne.value)]

    def close(self):
        if getattr(self, "h", None):
            self.lib.b2ot_dense_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def plan(mode, K, u, v, out=None):
    """KNOPP: u_i K_ij v_j (_sinkhorn.py:685-687); LOG: exp(K_ij + u_i + v_j) (:915-918).  `out`: optional destination."""
    lib = _lib.load()
    if out is None:
        out = torch.empty_like(K, memory_format=torch.contiguous_format)
    with torch.cuda.device(K.device):
        _lib.check(lib.b2ot_dense_plan(_lib.MODE_KNOPP if mode == "knopp" else _lib.MODE_LOG, dtype_code(K.dtype),
                                       K.data_ptr(), u.data_ptr(), v.data_ptr(), out.data_ptr(), K.shape[0],
                                       K.shape[1], K.stride(0), out.stride(0), stream_ptr(K.device)),
                   "b2ot_dense_plan")
    return out


def loss(mode, K, M, u, v):
    """<M, plan> without materialising the plan (_sinkhorn.py:435-438).  Returns a 0-d float64 tensor."""
    lib = _lib.load()
    out = torch.empty((), dtype=torch.float64, device=K.device)
    scratch = torch.empty(lib.b2ot_dense_loss_scratch_bytes(K.shape[0], K.shape[1]), dtype=torch.uint8,
                          device=K.device)
    with torch.cuda.device(K.device):
        _lib.check(lib.b2ot_dense_loss(_lib.MODE_KNOPP if mode == "knopp" else _lib.MODE_LOG, dtype_code(K.dtype),
                                       K.data_ptr(), M.data_ptr(), u.data_ptr(), v.data_ptr(), K.shape[0],
                                       K.shape[1], K.stride(0), M.stride(0), out.data_ptr(), scratch.data_ptr(),
                                       stream_ptr(K.device)), "b2ot_dense_loss")
    return out


def plan_value(mode, u, v, *, K=None, M=None, reg=0.0, a=None, b=None, c=None, terms=(), want_plan=True,
               want_rowsum=False, kmulq=False, out=None):
    """Plan + objective terms + row marginal in one pass (b2ot_dense_plan_value).

    K None: the Gibbs kernel / scaled cost is rebuilt from the cost M with `reg` (M read once, plan written once).
    terms : subset of {"linear", "entropy", "kl"}.
    out   : optional (n, m) destination of the plan (unit column stride), e.g. a Gibbs-kernel buffer rewritten in place.
    Returns (plan or None, values (3,) float64 device tensor [sum M*P, sum P log(P+1e-16), sum P log(P/q+1e-16)],
    rowsum or None).
    """
    lib = _lib.load()
    ref = K if K is not None else M
    n, m = ref.shape
    code = dtype_code(ref.dtype)
    bits = 0
    for t in terms:
        bits |= {"linear": _lib.VALUE_LINEAR, "entropy": _lib.VALUE_ENTROPY, "kl": _lib.VALUE_KL}[t]
    if kmulq:
        bits |= _lib.VALUE_KMULQ
    if out is None:
        out = torch.empty((n, m), dtype=ref.dtype, device=ref.device) if want_plan else None
    else:
        assert out.shape == (n, m) and out.dtype == ref.dtype and out.stride(1) == 1
    rows = torch.empty(n, dtype=ref.dtype, device=ref.device) if want_rowsum else None
    vals = torch.empty(3, dtype=torch.float64, device=ref.device)
    scratch = torch.empty(lib.b2ot_dense_plan_value_scratch_bytes(n, m), dtype=torch.uint8, device=ref.device)

    def ptr(t):
        return t.data_ptr() if t is not None else None

    def ld(t):
        return t.stride(0) if t is not None else 0

    with torch.cuda.device(ref.device):
        _lib.check(lib.b2ot_dense_plan_value(
            _lib.MODE_KNOPP if mode == "knopp" else _lib.MODE_LOG, code, ptr(K), ptr(M), ptr(c), float(reg),
            u.data_ptr(), v.data_ptr(), ptr(a), ptr(b), ptr(out), ptr(rows), n, m, ld(K), ld(M), ld(c), ld(out),
            bits, vals.data_ptr(), scratch.data_ptr(), stream_ptr(ref.device)), "b2ot_dense_plan_value")
    return out, vals, rows


def multi_hist_steps(knopp, max_iter, check_every):
    """Steps to enqueue for `max_iter` iterations.  KNOPP: the error of iterate k comes out of step k + 1 (its column
    pass), so one more (probe) step runs when the last iterate is a checked one."""
    return max_iter + (1 if knopp and (max_iter - 1) % check_every == 0 else 0)


def multi_hist_select(status, at, max_iter):
    """Device state (stats[12], stats[13]) -> (buffer index of the result, reference `ii`, solver status).

    Step k reads buffer k & 1 and writes the other one.  Status 1 (converged at `at`): KNOPP notices it in step at + 1,
    which READ that iterate; UNBALANCED in step `at`, which WROTE it - buffer (at + 1) & 1 either way.  Status 2 (guard
    in step `at`): the iterate that step started from - unless it was KNOPP's probe step (at == max_iter), whose own
    update is not part of the solve."""
    if status == 1:
        return (at + 1) & 1, at, _lib.ST_CONVERGED
    if status == 2 and at < max_iter:
        return at & 1, at, _lib.ST_NUMERR
    return max_iter & 1, max_iter - 1, _lib.ST_MAXITER


def multi_hist_solve(mode, K, a, B, u0, v0, max_iter, stop_thr, check_every=10, fi=(1.0, 1.0), poll_every=50):
    """H target histograms on ONE kernel matrix, decisions on the device (b2ot_multi_scaling_step).

    mode "knopp": sinkhorn_knopp with a 2-D b (ot/bregman/_sinkhorn.py:626-663: v-then-u, joint guard, joint error
    every `check_every` iterations); "unbalanced": sinkhorn_knopp_unbalanced with a 2-D b
    (ot/unbalanced/_sinkhorn.py:749-787: u-then-v with exponents `fi`, joint guard, joint error every iteration).
    K (n, m), a (n,), B (m, H), u0 (n, H), v0 (m, H).  Steps are enqueued `poll_every` at a time; once the status word
    is set the remaining steps are no-ops.  Returns (u, v, errs, niter, status) with the reference's `ii` at the end of
    its loop as niter and status in {_lib.ST_CONVERGED, ST_NUMERR, ST_MAXITER}; (u, v) is the iterate the reference
    returns (the previous one after a guard stop)."""
    lib = _lib.load()
    n, m = K.shape
    H = B.shape[1]
    dt, dev = K.dtype, K.device
    knopp = mode == "knopp"
    code = _lib.MULTI_KNOPP if knopp else _lib.MULTI_UNBALANCED
    max_iter, check_every = int(max_iter), int(check_every)
    us = [u0.to(dt).contiguous().clone(), torch.empty((n, H), dtype=dt, device=dev)]
    vs = [v0.to(dt).contiguous().clone(), torch.empty((m, H), dtype=dt, device=dev)]
    err_cap = (max_iter // check_every + 2) if knopp else max_iter + 1
    with torch.cuda.device(dev):
        stats = torch.zeros(16, dtype=torch.float64, device=dev)
        err_hist = torch.zeros(err_cap, dtype=torch.float64, device=dev)
        scratch = torch.empty(lib.b2ot_multi_scratch_bytes(n, m, H), dtype=torch.uint8, device=dev)
        stream = stream_ptr(dev)
        steps = multi_hist_steps(knopp, max_iter, check_every)
        k, st = 0, None
        while k < steps:
            for _ in range(min(int(poll_every), steps - k)):
                s = k & 1
                _lib.check(lib.b2ot_multi_scaling_step(
                    dtype_code(dt), code, K.data_ptr(), n, m, K.stride(0), H, a.data_ptr(), B.data_ptr(),
                    us[s].data_ptr(), vs[s].data_ptr(), us[s ^ 1].data_ptr(), vs[s ^ 1].data_ptr(), k, check_every,
                    float(stop_thr), float(fi[0]), float(fi[1]), stats.data_ptr(), err_hist.data_ptr(), err_cap,
                    scratch.data_ptr(), stream), "b2ot_multi_scaling_step")
                k += 1
            st = stats.tolist()  # synchronises
            if st[12] != 0.0:
                break
        errs = err_hist[:min(int(st[14]), err_cap)].tolist()
    sel, niter, status = multi_hist_select(int(st[12]), int(st[13]), max_iter)
    return us[sel], vs[sel], errs, niter, status


def colsum(P):
    """Column marginal P^T 1 (b2ot_colsum), deterministic."""
    lib = _lib.load()
    n, m = P.shape
    out = torch.empty(m, dtype=P.dtype, device=P.device)
    scratch = torch.empty(lib.b2ot_colsum_scratch_bytes(n, m), dtype=torch.uint8, device=P.device)
    with torch.cuda.device(P.device):
        _lib.check(lib.b2ot_colsum(dtype_code(P.dtype), P.data_ptr(), n, m, P.stride(0), out.data_ptr(),
                                   scratch.data_ptr(), stream_ptr(P.device)), "b2ot_colsum")
    return out
