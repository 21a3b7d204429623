"""Errors of the online-cost solver against the fp64 oracle on coincident clouds (X is Y) and on distinct clouds, by reg,
variant and operand kind.  usage: [B2OT_LIB=...] python scripts/diag_coincident.py"""
import os, sys, warnings
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot
from oracle import sinkhorn_oracle as orc

n, d = 900, 64
rng = np.random.RandomState(5)
xs = (rng.randn(n, d) / np.sqrt(d)).astype(np.float32)
xs[1::50] = xs[::50][: len(xs[1::50])]
xt = (rng.randn(n, d) / np.sqrt(d) + 0.1).astype(np.float32)
warnings.simplefilter("ignore")
for same in (True, False):
    for reg in (0.5, 0.1, 0.02):
        Y = xs if same else xt
        fr, gr = orc.empirical_sinkhorn_lazy(xs.astype(np.float64), Y.astype(np.float64), reg, numIterMax=20, stopThr=-1,
                                             batchSize=256)
        for variant, operand in (("tc", "f16"), ("tc", "tf32"), ("simt", "f16")):
            X = torch.as_tensor(xs, device="cuda")
            Yd = X if same else torch.as_tensor(Y, device="cuda")
            f, g = ot.empirical_sinkhorn(X, Yd, reg, numIterMax=20, stopThr=-1, isLazy=True, b2ot_variant=variant,
                                         b2ot_operand=operand)
            ef = np.abs(f.cpu().numpy() - fr).max()
            eg = np.abs(g.cpu().numpy() - gr).max()
            print(f"same={same} reg={reg} {variant}/{operand}: abs err f {ef:.3e} g {eg:.3e}  (|f|max {np.abs(fr).max():.2f}, "
                  f"|g|max {np.abs(gr).max():.2f})", flush=True)
