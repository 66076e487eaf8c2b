#!/usr/bin/env python
"""diagnostic: incremental refit (tb200_factor_append + tb200_alpha_lml) vs full refactorisation at n ~ 4000"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import turbo_b200 as tb  # noqa: E402
from turbo_b200 import _lib  # noqa: E402

n0 = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
X, y, theta = bench.synthetic_problem(n0 + 40, bench.DIM)
spec = tb.KernelSpec(_lib.KIND_MATERN52, bench.DIM, True)
mp = dict(kernel=(spec, theta), normalize_y=True)
for incremental in (True, False):
    sur = tb.GpuGPSurrogate(model_params=mp, training_iterations=0, incremental=incremental)
    sur.construct_model(0, X[:n0], y[:n0])
    torch.cuda.synchronize()
    ts = []
    for t in range(1, 21):
        t0 = time.perf_counter()
        m, info = sur.construct_model(t, X[:n0 + t], y[:n0 + t])
        torch.cuda.synchronize()
        ts.append(1e3 * (time.perf_counter() - t0))
    print("incremental=%s: construct_model per trial (one new point, n=%d..%d): median %.2f ms, min %.2f ms, lml %.6f"
          % (incremental, n0 + 1, n0 + 20, np.median(ts), min(ts), m.get_log_likelihood()))
