"""Tiny driver for ncu / timing experiments: a few sweeps of one configuration.

    python scripts/prof_case.py --mode knopp --n 16384 --m 16384 --iters 5 [--dtype f32] [--variant auto]
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scripts.gpu_diag import perf_case  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="knopp")
ap.add_argument("--n", type=int, default=16384)
ap.add_argument("--m", type=int, default=16384)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--dtype", default="f32")
ap.add_argument("--variant", default="auto")
ap.add_argument("--label", default="")
a = ap.parse_args()
perf_case(a.mode, a.n, a.m, np.float32 if a.dtype == "f32" else np.float64, a.iters, variant=a.variant,
          label=a.label + " ")
