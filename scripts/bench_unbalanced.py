"""Unbalanced scaling-domain Sinkhorn at BASELINE config 3's shape (n = m = 65536 fp32, K = 16 GiB, one B200).

Same fused sweep kernel as the balanced headline (row exponent fi_1 applied in the row update), finalize_unbal_kernel
for v = (b / K^T u)^fi_2 + guard + inf-norm error every iteration.  Algorithmic bytes per iteration = 4 n m + vectors
(K read ONCE; the reference's two GEMVs read it twice).  CUDA events on the launch stream (b2ot_dense_profile).
Prints one JSON line.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pot_b200 import _dense, _lib  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    dev = "cuda:0"
    peak = 6550.0
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    g = torch.Generator(device=dev).manual_seed(0)
    xs = torch.randn(n, 2, device=dev, generator=g)
    xt = torch.randn(n, 2, device=dev, generator=g) + 1.0
    M = torch.cdist(xs, xt).pow_(2)
    M /= M.max()
    lib = _lib.load()
    _lib.check(lib.b2ot_kexp(M.data_ptr(), M.data_ptr(), n, n, M.stride(0), M.stride(0), 0.05, _lib.F32,
                             _dense.stream_ptr(M.device)), "b2ot_kexp")
    K = M
    a = torch.full((n,), 1.3 / n, device=dev)
    b = torch.full((n,), 1.0 / n, device=dev)
    out = {}
    cases = (("balanced (fi = 1)", None), ("unbalanced reg_m = (1, 0.5), reg = 0.05", (1 / 1.05, 0.5 / 0.55)))
    if len(sys.argv) > 3 and sys.argv[3] == "balanced_only":
        cases = cases[:1]
    for name, fi in cases:
        s = _dense.DenseSolver("knopp", K, a, b, None, max_iter=100000, check_every=1 if fi else 10, stop_thr=0.0)
        if fi:
            _lib.check(lib.b2ot_dense_set_unbalanced(s.h, *fi), "b2ot_dense_set_unbalanced")
            s.init(torch.ones_like(a), torch.ones_like(b))
        else:
            s.init(a.clone(), b.clone())
        s.iterate(5)
        torch.cuda.synchronize()
        sweep_ms, iter_ms = s.profile(iters)
        st, it = s.poll()
        nbytes = 4.0 * n * n + 4.0 * (2 * n + 3 * n)
        out[name] = {"sweep_ms": round(sweep_ms, 4), "iter_ms": round(iter_ms, 4), "it_per_s": round(1e3 / iter_ms, 1),
                     "sweep_GBps": round(nbytes / sweep_ms / 1e6, 1), "frac_of_measured_hbm":
                     round(nbytes / sweep_ms / 1e6 / peak, 3), "status": st, "iter": it}
        s.close()
    print(json.dumps({"workload": f"dense fp32 n=m={n} scaling-domain sweep, {iters} iterations", "peak_GBps": peak,
                      "lib": os.path.basename(_lib.LIB_PATH),
                      **out}))


if __name__ == "__main__":
    main()
