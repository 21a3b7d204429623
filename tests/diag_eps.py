import os, sys, warnings
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot
from pot_b200 import stabilized
from oracle import solvers_oracle as sorc
warnings.simplefilter("ignore")
z = np.load("tests/golden/golden_stabilized.npz")
M, a, b, reg = z["eps_f64/M"], z["eps_f64/a"], z["eps_f64/b"], float(z["eps_f64/reg"])
Md, ad, bd = (torch.as_tensor(x, device="cuda:0") for x in (M, a, b))
al, be = np.zeros_like(a), np.zeros_like(b)
ald, bed = torch.zeros_like(ad), torch.zeros_like(bd)
for ii in range(40):
    regi = (1e4 - reg) * np.exp(-ii) + reg
    G, lg = sorc.sinkhorn_stabilized(a, b, M, regi, numItermax=100, stopThr=1e-9, warmstart=(al, be), print_period=20, tau=1e3, log=True)
    # GPU from the ORACLE's warm start (isolates one stage) and from its own chain
    Gd, lgd = ot.bregman.sinkhorn_stabilized(ad, bd, Md, regi, numItermax=100, stopThr=1e-9, warmstart=(torch.as_tensor(al, device="cuda:0"), torch.as_tensor(be, device="cuda:0")), log=True)
    Gc, lgc = ot.bregman.sinkhorn_stabilized(ad, bd, Md, regi, numItermax=100, stopThr=1e-9, warmstart=(ald, bed), log=True)
    d1 = np.abs(Gd.cpu().numpy() - G).max() / G.max()
    d2 = np.abs(Gc.cpu().numpy() - G).max() / G.max()
    print(ii, f"reg={regi:.3e}", "n_iter", lg["n_iter"], lgd["n_iter"], lgc["n_iter"], "absorb", lg["n_absorb"],
          "err", [f"{e:.3e}" for e in lg["err"][-2:]], [f"{float(e):.3e}" for e in lgd["err"][-2:]], f"dstage={d1:.2e} dchain={d2:.2e}")
    al, be = lg["alpha"], lg["beta"]
    ald, bed = lgc["alpha"], lgc["beta"]
