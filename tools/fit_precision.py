#!/usr/bin/env python
"""diagnostic: accuracy and time of the batched LML+gradient under the per-stage slice-product switches
(TB200_DEBUG=fit_engine=dmma,fit_gmax=..,fit_gmax_trtri=..,fit_gmax_kinv=..; csrc/common.cuh), n=4096, d=10, R=64.
    TB200_DEBUG=.. python tools/fit_precision.py run <tag>     # writes gpurun_out/fitprec_<tag>.npz
    python tools/fit_precision.py compare <ref_tag> <tag> [<tag> ..]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "gpurun_out")


def run(tag, R=64):
    import torch
    import bench
    import turbo_b200 as tb
    from turbo_b200 import _lib, device
    X, y, theta = bench.synthetic_problem()
    yn = (y - y.mean()) / y.std()
    spec = tb.KernelSpec(_lib.KIND_MATERN52, bench.DIM, True)
    rng = np.random.default_rng(2)
    thetas = np.clip(theta + rng.normal(0, 0.5, (R, theta.size)), np.log(1e-2) * 2, np.log(1e2))
    # a few hard members: tiny noise (ill-conditioned K) and long length scales
    thetas[1, -1] = np.log(1e-8)
    thetas[2, 1:-1] = np.log(3.0)
    thetas[3, -1], thetas[3, 1:-1] = np.log(1e-6), np.log(2.0)
    X_dev, yn_dev = device.to_device(X), device.to_device(yn)
    ts = []
    for it in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        lml, grad, info = device.lml_batched(X_dev, yn_dev, bench.N_TRAIN, spec, thetas, 1e-10, True)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    os.makedirs(OUT, exist_ok=True)
    np.savez(os.path.join(OUT, "fitprec_%s.npz" % tag), lml=lml, grad=grad, info=info, ms=min(ts[1:]))
    print("%s: %.2f ms per evaluation, info!=0: %d" % (tag, min(ts[1:]), int((info != 0).sum())))


def compare(ref, tags):
    a = np.load(os.path.join(OUT, "fitprec_%s.npz" % ref))
    ok = a["info"] == 0
    print("reference %s: %.2f ms; %d of %d members factorised" % (ref, a["ms"], ok.sum(), ok.size))
    for t in tags:
        b = np.load(os.path.join(OUT, "fitprec_%s.npz" % t))
        both = ok & (b["info"] == 0)
        dl = np.abs(a["lml"][both] - b["lml"][both]) / np.abs(a["lml"][both])
        gs = np.abs(a["grad"][both]).max(axis=1, keepdims=True)
        dg = (np.abs(a["grad"][both] - b["grad"][both]) / gs).max(axis=1)
        print("%-12s %.2f ms  same info: %s  LML rel diff max %.2e median %.2e   grad diff / max|grad| max %.2e median %.2e"
              % (t, b["ms"], np.array_equal(a["info"] != 0, b["info"] != 0), dl.max(), np.median(dl), dg.max(),
                 np.median(dg)))


if __name__ == "__main__":
    if sys.argv[1] == "run":
        run(sys.argv[2])
    else:
        compare(sys.argv[2], sys.argv[3:])
