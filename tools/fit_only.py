#!/usr/bin/env python
"""diagnostic: batched LML+gradient only (n=4096, d=10, R restarts) -- for ncu launch lists of the fit path.
    python tools/fit_only.py [R] [reps]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import turbo_b200 as tb  # noqa: E402
from turbo_b200 import _lib, device  # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
X, y, theta = bench.synthetic_problem()
yn = (y - y.mean()) / y.std()
spec = tb.KernelSpec(_lib.KIND_MATERN52, bench.DIM, True)
rng = np.random.default_rng(2)
thetas = np.clip(theta + rng.normal(0, 0.5, (R, theta.size)), np.log(1e-2) * 2, np.log(1e2))
X_dev, yn_dev = device.to_device(X), device.to_device(yn)
for it in range(reps):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    lml, grad, info = device.lml_batched(X_dev, yn_dev, bench.N_TRAIN, spec, thetas, 1e-10, True)
    torch.cuda.synchronize()
    print("rep %d: %.2f ms  lml[0]=%.6f finite=%s" % (it, 1e3 * (time.perf_counter() - t0), lml[0], np.isfinite(lml).all()))
