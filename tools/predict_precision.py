#!/usr/bin/env python
"""diagnostic: the predict-side twin of profiles/r1_fit_precision.md.

For each case and each variance engine -- dmma (FP64 DMMA GEMM), i8l / i8 / i8x (INT8 tensor cores with 21 / 28 / 34
byte-slice products) -- the predictive variance, EI and argmax over a candidate set are compared with the ORACLE
(oracle/gp_oracle.py, checker only):

    max |var - var_oracle| in units of eps * n * prior_var   (the parity floor of the tests is 256 of these)
    the same restricted to candidates where 1e-8 relative does not already hold
    max EI error relative to max |EI|,  argmax agreement

    python tools/predict_precision.py [--cases headline,cfg2,illcond,n8200,cfg5] > gpurun_out/predict_precision.md
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = {
    # name: (n, d, kind, sf2, ls, noise, problem)
    "headline": (4096, 10, 3, 1.0, 0.5, 1e-4, "sines"),
    "cfg2": (1024, 6, 3, 1.0, 0.5, 1e-6, "hartmann6"),
    "illcond": (2048, 3, 0, 1.0, 0.5, 1e-8, "sines"),
    "n8200": (8200, 8, 3, 1.0, 0.5, 1e-4, "sines"),
    "cfg5": (16384, 20, 0, 1.0, 1.0, 1e-3, "sines"),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="headline,cfg2,illcond,n8200")
    ap.add_argument("--m", type=int, default=2048)
    args = ap.parse_args()
    import torch
    import bench
    import turbo_b200 as tb
    from turbo_b200 import device
    from oracle import gp_oracle as O
    eps = np.finfo(np.float64).eps
    print("| case | n | engine | max var err / (eps n prior) | ... where 1e-8 rel fails | max var rel err | EI max err / max EI | argmax = oracle | ms / 2048 cand |")
    print("|---|---|---|---|---|---|---|---|---|")
    for name in args.cases.split(","):
        n, d, kind, sf2, ls, noise, problem = CASES[name]
        X, y, theta = bench.synthetic_problem(n, d, 0, problem, sf2, ls, noise)
        rng = np.random.default_rng(1)
        Xc = rng.random((args.m, d))
        Xc[:256] = X[rng.choice(n, 256, replace=False)]               # on training points: var -> noise
        Xc[256:512] = X[rng.choice(n, 256, replace=False)] + 1e-3 * rng.standard_normal((256, d))  # very close to them
        t0 = time.perf_counter()
        og = O.FittedGP(O.KernelSpec(kind, d, True), theta, X, y)
        mu_o, sd_o = og.predict(Xc)
        t_or = time.perf_counter() - t0
        var_o = sd_o ** 2
        ref = O.acquisition(O.ACQ_EI, mu_o, sd_o, -1.0, float(y.min()), 0.01)
        k_o = O.argmax_first(ref)
        prior = (np.exp(theta[0]) + np.exp(theta[-1])) * og.y_std ** 2
        unit = eps * n * prior
        spec = tb.KernelSpec(kind, d, True)
        for engine in ("dmma", "i8l", "i8", "i8x", "auto"):
            sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0,
                                    variance_engine=engine)
            model, _ = sur.construct_model(0, X, y)
            acq = tb.GpuEI(0.01).construct_function(0, model, "min", float(y.min()))[0]
            mu, sd = model.predict(Xc, return_std_dev=True)
            vals = acq(Xc)
            _v, idx, _x = acq.argmax_over(candidates=Xc)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a = device.Acquirer(model.factor, acq._spec_c(), engine=engine)
            if engine == "auto":
                engine = "auto -> gmax %d (probe %.2f units)" % (a.gmax, model.factor.gmax_probe)
            Xd = device.to_device(Xc)
            a.run(Xd, args.m, 0, None)
            torch.cuda.synchronize()
            e0.record()
            a.run(Xd, args.m, 0, None)
            e1.record()
            torch.cuda.synchronize()
            dv = np.abs(sd ** 2 - var_o)
            need = dv > 1e-8 * var_o
            print("| %s | %d | %s | %.2f | %.2f | %.2e | %.2e | %s | %.3f |" % (
                name, n, engine, dv.max() / unit, (dv[need] / unit).max() if need.any() else 0.0,
                (dv / np.maximum(var_o, 1e-300)).max(), np.abs(vals - ref).max() / max(np.abs(ref).max(), 1e-300),
                "yes" if idx == k_o else "NO (%d vs %d)" % (idx, k_o), e0.elapsed_time(e1) * 2048.0 / args.m))
            del model, a
        print("<!-- %s: oracle %.1f s, max |mu| err vs oracle %.2e -->" % (name, t_or, np.abs(mu - mu_o).max()))
        sys.stdout.flush()


if __name__ == "__main__":
    main()
