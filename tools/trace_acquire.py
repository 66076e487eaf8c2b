#!/usr/bin/env python
"""diagnostic: per-kernel event timeline of one tb200_acquire_i8 call (two internal streams).
    python tools/trace_acquire.py [m] -> prints the records of the first 12 batches and per-kind averages"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import turbo_b200 as tb  # noqa: E402
from turbo_b200 import _lib, device  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 17
lib = _lib.load()
X, y, theta = bench.synthetic_problem()
spec = tb.KernelSpec(_lib.KIND_MATERN52, bench.DIM, True)
sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0)
model, _ = sur.construct_model(0, X, y)
acq, _ = tb.GpuEI(0.01).construct_function(0, model, "min", float(y.min()))
a = device.Acquirer(model.factor, acq._spec_c(), 0, engine=None)
gen = _lib.cand_spec(_lib.CAND_UNIFORM, 7, m, [0.0] * bench.DIM, [1.0] * bench.DIM)
Xc = device.generate_candidates(gen, bench.DIM, 0, m)
for _ in range(2):
    a.reset(); a.run(Xc, m, 0, None); a.best()
torch.cuda.synchronize()
lib.tb200_trace_enable(1)
a.reset(); a.run(Xc, m, 0, None); a.best()
torch.cuda.synchronize()
cap = 4096
t = (C.c_double * cap)(); tags = (C.c_int * cap)()
k = lib.tb200_trace_read(t, tags, cap)
lib.tb200_trace_enable(0)
rec = [(t[i], tags[i] // 16, tags[i] % 16) for i in range(k)]
names = {0: "K* start", 4: "K* end", 1: "var start", 2: "var end", 3: "epi end"}
for r in sorted(rec[:60]):
    print("%9.3f ms  slot %d  %s" % (r[0], r[1], names[r[2]]))
by = {kind: np.array([r[0] for r in rec if r[2] == kind]) for kind in names}
nb = len(by[3])
print("batches %d  total %.3f ms  -> %.1f us/batch" % (nb, by[3][-1] - by[0][0], 1e3 * (by[3][-1] - by[0][0]) / nb))
print("avg K* kernel %.1f us, var kernel (start->end events) %.1f us" % (
    1e3 * np.mean(by[4] - by[0]), 1e3 * np.mean(by[2] - by[1])))
print("var-kernel end-to-end spacing: median %.1f us" % (1e3 * np.median(np.diff(np.sort(by[2])))))
