#!/usr/bin/env python
"""diagnostic: cost of the too-close exclusion (tb200_min_sqdist inside tb200_acquire_i8_excl) at the headline shape"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import turbo_b200 as tb  # noqa: E402
from turbo_b200 import _lib, device  # noqa: E402

n, d, m = bench.N_TRAIN, bench.DIM, 1 << 21
X, y, theta = bench.synthetic_problem(n, d)
spec = tb.KernelSpec(_lib.KIND_MATERN52, d, True)
sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0)
model = sur.construct_model(0, X, y)[0]
acq = tb.GpuEI(0.01).construct_function(0, model, "min", float(y.min()))[0]
lo, hi = [0.0] * d, [1.0] * d


def timed(f, reps=5):
    f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        f()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


Xq = device.to_device(np.random.default_rng(0).random((m, d)))
X_dev = device.to_device(X)
out = torch.empty(m, dtype=torch.float64, device=Xq.device)
lib = _lib.load()
t = timed(lambda: _lib.check(lib.tb200_min_sqdist(_lib.ptr(Xq), m, _lib.ptr(X_dev), n, d, _lib.ptr(out), None,
                                                  _lib.stream_ptr())))
ops = 3.0 * m * n * d
print("tb200_min_sqdist m=%d n=%d d=%d: %.3f ms (%.2f TFLOP/s FP64 add/mul, 3 n d per candidate)" % (m, n, d, t, ops / t / 1e9))
for ex in (None, (X, 1e-10)):
    t = timed(lambda: acq.argmax_over(generator="uniform", m=m, lo=lo, hi=hi, seed=1, exclude=ex), reps=3)
    print("argmax_over m=%d exclude=%s: %.2f ms per pass (%.3g cand/s)" % (m, "off" if ex is None else "tol 1e-10", t, m / t * 1e3))
