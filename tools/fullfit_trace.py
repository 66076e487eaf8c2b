#!/usr/bin/env python
"""diagnostic: batch size and wall time of every tb200_lml_batched call inside one whole fit (R restarts, <= 30 its)"""
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
X, y, theta = bench.synthetic_problem()
spec = tb.KernelSpec(_lib.KIND_MATERN52, bench.DIM, True)
T = np.clip(theta + np.random.default_rng(2).normal(0, 0.5, (R, theta.size)), np.log(1e-2) * 2, np.log(1e2))
log = []
orig = device.lml_batched


def traced(X_dev, yn_dev, n, spec_, thetas, jitter, with_grad=True):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = orig(X_dev, yn_dev, n, spec_, thetas, jitter, with_grad)
    torch.cuda.synchronize()
    log.append((len(np.atleast_2d(thetas)), 1e3 * (time.perf_counter() - t0), int((out[2] != 0).sum())))
    return out


device.lml_batched = traced
sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=R, max_iter=30, seed=0)
t0 = time.perf_counter()
m, info = sur.construct_model(1, X, y, start_thetas=T)
torch.cuda.synchronize()
print("whole fit %.1f ms, %d calls, sum of call times %.1f ms, best -lml %.4f" % (
    1e3 * (time.perf_counter() - t0), len(log), sum(l[1] for l in log), info['best_neg_lml']))
print("iterations per restart:", info['optimiser_iterations'])
for i, (k, ms, bad) in enumerate(log):
    print("call %3d: %3d members %7.1f ms  (%.2f ms/member)%s" % (i, k, ms, ms / k, "  failed=%d" % bad if bad else ""))
