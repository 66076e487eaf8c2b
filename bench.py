#!/usr/bin/env python
"""bench.py -- EI candidates/sec (n_train=4096, d=10) + GP fit ms, per BASELINE.json.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config headline|cfg2|cfg4|cfg5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over one batch of synthetic candidates: predict + acquisition + argmax
over this rank's candidates (no data-path collective, one 16-byte all-gather per acquisition pass) on a GP
at fixed hyper-parameters.  --config selects the BASELINE.json configuration:

  headline (default) n=4096, d=10, ARD Matern-5/2 + white; EI ('min', xi=0.01) over 2^21 uniform candidates PER GPU
                     (weak scaling) + the `fit` block: batched LML+gradient over 64 restarts and a whole fit (configs[2])
  cfg2   configs[1]  Hartmann-6, n=1024, d=6; EI over 2^20 uniform candidates (fixed total, 1 B200)
  cfg4   configs[3]  n=4096, d=10; UCB beta=2 ('max') AND -LCB ('min') over a FIXED 2^24-point Latin hypercube sharded over
                     the N GPUs (strong scaling); a step is the two passes
  cfg5   configs[4]  n=16384, d=20, ARD RBF; one factorisation timed + EI over 2^26 candidates sharded over 8 GPUs
                     (2^23 per GPU at any N: 22 s per step on one GPU)

  value     : candidates/s, candidates already resident in HBM (device-timed, max over ranks)
  e2e       : same metric through the plugin call acq.argmax_over(host candidates): pinned host
              buffer -> chunked H2D inside the timed region -> (value, index) read back
  roofline  : the variance contraction V = L^-1 k* (n^2 flop per candidate) on the INT8 tensor cores (tcgen05.mma
              kind::i8, exact byte-slice products): algorithmic INT8 ops / CUDA-event duration against the INT8 tensor
              peak MEASURED on this device by tb200_i8_peak (2 x the driver's bf16 figures kept as cross-checks);
              --engine dmma: FP64 DMMA GEMM against the FP64 tensor peak measured by tb200_fp64_peak
  parity_check : AFTER the timed region, the first --parity-n of THIS step's own candidates (regenerated on the host from
              the same (seed, global index) rule) go through the oracle; mu / variance / acquisition / argmax of the
              timed configuration (same batch size, K* kernel, slots and engine) are compared with it
  sharded_equals_unsharded (N > 1): a 2^18 generated set evaluated sharded over the N ranks, unsharded on every rank,
              and sharded with the index ranges in REVERSE rank order (so the winner also lives on a remote rank)
  cpu_baseline : the reference's own objects (turbo's SciKitGPSurrogate.ModelInstance + EI/UCB FunctionInstance over
              scikit-learn; kind "reference") when mbway/turbo is importable (baseline/_ref), else the oracle port,
              on those same candidates; rank 0, N=1 only
  fit       : GP fit ms -- one batched LML+gradient evaluation over R restarts (R split over ranks)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# rank 0 prints ONE JSON line on stdout.  Libraries write there too (NCCL prints "NCCL version ..." at its default
# debug level, from C, whatever NCCL_DEBUG_FILE says), so file descriptor 1 is pointed at stderr for the whole run and
# the JSON line goes to the saved descriptor of the real stdout.
sys.stdout.flush()
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_TRAIN, DIM = 4096, 10
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel at the headline shape, from the
# committed `ncu --set full` captures under profiles/ (None until captured for that engine / shape)
# profiles/r1_gemm_ncu_full.md, r1_i8v2_ncu_full.md, r2_variance_ncu_full.md (persistent kernel, 21 products)
TRAFFIC = {"dmma": 197.2e6, "i8g8": 163.9e6, "i8g7": 117.9e6}
METRIC = "EI candidates/sec (n=4096,d=10)"
KIND_RBF, KIND_MATERN52 = 0, 3

CONFIGS = {
    "headline": dict(n=4096, d=10, kind=KIND_MATERN52, problem="sines", sf2=1.0, ls=0.5, noise=1e-4,
                     acqs=[("EI", "min", 0.01)], generator="uniform", seed=7, m_per_gpu=1 << 21, scaling="weak",
                     metric=METRIC, baseline_config="BASELINE.json headline; configs[3] shape"),
    "cfg2": dict(n=1024, d=6, kind=KIND_MATERN52, problem="hartmann6", sf2=1.0, ls=0.5, noise=1e-6,
                 acqs=[("EI", "min", 0.01)], generator="uniform", seed=11, m_total=1 << 20, scaling="strong",
                 metric="EI candidates/sec (Hartmann-6, n=1024,d=6, 2^20 candidates)", baseline_config="configs[1]"),
    "cfg4": dict(n=4096, d=10, kind=KIND_MATERN52, problem="sines", sf2=1.0, ls=0.5, noise=1e-4,
                 acqs=[("UCB", "max", 2.0), ("UCB", "min", 2.0)], generator="lhs", seed=13, m_total=1 << 24,
                 scaling="strong", metric="UCB/LCB candidates/sec (n=4096,d=10, 2^24 LHS candidates sharded)",
                 baseline_config="configs[3]"),
    "cfg5": dict(n=16384, d=20, kind=KIND_RBF, problem="sines", sf2=1.0, ls=1.0, noise=1e-3,
                 acqs=[("EI", "min", 0.01)], generator="uniform", seed=17, m_per_gpu=1 << 23, scaling="weak",
                 metric="EI candidates/sec (n=16384,d=20 ARD RBF, 2^26 candidates on 8 GPUs = 2^23 per GPU)",
                 baseline_config="configs[4]"),
}


def synthetic_problem(n=N_TRAIN, d=DIM, seed=0, problem="sines", sf2=1.0, ls=0.5, noise=1e-4):
    """SURVEY 8d data: cfg3/4/5 y = sum sin(2 pi x_j) + 0.01 N(0,1); cfg2 Hartmann-6.  theta = fixed hyper-parameters"""
    rng = np.random.default_rng(seed)
    X = rng.random((n, d))
    if problem == "hartmann6":
        from oracle.gp_oracle import hartmann6  # the test function only (data generation, not the measured path)
        y = hartmann6(X)
    else:
        y = np.sin(2 * np.pi * X).sum(axis=1) + 0.01 * rng.standard_normal(n)
    theta = np.log(np.r_[sf2, np.full(d, ls), noise])
    return X, y, theta


def resolve_config(args):
    cfg = dict(CONFIGS[args.config])
    if args.n:
        cfg["n"] = args.n
    if args.d:
        cfg["d"] = args.d
    if args.m:
        if cfg["scaling"] == "weak":
            cfg["m_per_gpu"] = args.m
        else:
            cfg["m_total"] = args.m
    if args.n or args.d or args.m:
        cfg["metric"] += " [shape overridden: n=%d d=%d]" % (cfg["n"], cfg["d"])
    return cfg


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.lines, self.proc = str(gpu_index), [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", self.gpu, "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True).start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def host_threads():
    """lift the BLAS pools to every host core (torchrun exports OMP_NUM_THREADS=1 to its ranks) -> threads in use"""
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    try:
        from threadpoolctl import threadpool_info, threadpool_limits
        threadpool_limits(limits=cores)
        return max([p.get("num_threads", 1) for p in threadpool_info()] + [1])
    except Exception:
        return 1


def step_candidates_host(cfg, m_total, count, first=0):
    """rows [first, first+count) of the step's candidate set, regenerated on the host from the (seed, global index) rule
    the device generator uses (bit-identical: tests/test_gpu_path.py::test_device_generator_matches_host_mirror)"""
    from turbo_b200 import _lib
    mode = _lib.CAND_UNIFORM if cfg["generator"] == "uniform" else _lib.CAND_LHS
    d = cfg["d"]
    return _lib.host_candidates(mode, cfg["seed"], m_total, [0.0] * d, [1.0] * d, np.arange(first, first + count))


class CpuArm:
    """The reference's CPU implementation of the path at the config's fixed hyper-parameters.
    kind 'reference': turbo's own SciKitGPSurrogate.ModelInstance wrapping a scikit-learn GaussianProcessRegressor
    (kernel fixed, optimizer=None; surrogates.py:328-338) + turbo's EI/UCB FunctionInstance
    (acquisition_functions.py:147-158, :336-358) + the argsort selection of auxiliary_optimisers.py:60-66.
    kind 'port': the oracle restatement of the same lines (only when mbway/turbo is not importable)."""

    def __init__(self, cfg, X, y, theta):
        self.cfg, self.threads = cfg, host_threads()
        d = cfg["d"]
        self.kind = "port"
        self.incumbent = float(y.min())
        t0 = time.perf_counter()
        try:
            from turbo_b200 import _reference_api
            if not _reference_api.HAVE_TURBO:
                raise ImportError("mbway/turbo is not importable")
            import sklearn.gaussian_process as sk_gp
            import turbo.modules as tm
            K = sk_gp.kernels
            sf2, ls, noise = float(np.exp(theta[0])), np.exp(theta[1:1 + d]), float(np.exp(theta[-1]))
            base = K.RBF(length_scale=ls) if cfg["kind"] == KIND_RBF else K.Matern(length_scale=ls, nu=2.5)
            kern = K.ConstantKernel(sf2) * base + K.WhiteKernel(noise)
            gp = sk_gp.GaussianProcessRegressor(kernel=kern, optimizer=None, normalize_y=True).fit(X, y)
            model = tm.SciKitGPSurrogate.ModelInstance(gp)
            self.funcs = []
            for name, ext, param in cfg["acqs"]:
                if name == "EI":
                    f, _ = tm.EI(xi=param).construct_function(0, model, ext, self.incumbent)
                elif name == "PI":
                    f, _ = tm.PI(xi=param).construct_function(0, model, ext, self.incumbent)
                else:
                    f, _ = tm.UCB(beta=param).construct_function(0, model, ext)
                self.funcs.append(f)
            self.kind = "reference"
            self.what = ("turbo.modules.SciKitGPSurrogate.ModelInstance over scikit-learn GaussianProcessRegressor "
                         "(optimizer=None) + turbo.modules.%s FunctionInstance, unmodified reference from baseline/_ref"
                         % "/".join(sorted({a[0] for a in cfg["acqs"]})))
        except Exception as e:  # noqa: BLE001 - the port is the documented second choice
            from oracle import gp_oracle as O
            gp = O.FittedGP(O.KernelSpec(cfg["kind"], d, True), theta, X, y)
            codes = {"EI": O.ACQ_EI, "PI": O.ACQ_PI, "UCB": O.ACQ_UCB}

            def make(name, ext, param):
                sign = 1.0 if ext == "max" else -1.0

                def f(Xc):
                    mu, sd = gp.predict(Xc)
                    return O.acquisition(codes[name], mu, sd, sign, self.incumbent, param)
                return f
            self.funcs = [make(*a) for a in cfg["acqs"]]
            self.what = "oracle/gp_oracle.py, the NumPy/SciPy port of the reference's sklearn path (%s)" % e
        self.fit_s = time.perf_counter() - t0

    def rate(self, Xc, reps=1, chunk=4096):
        """every acquisition of the config over the rows of Xc in chunks (sklearn materialises K* (m,n) and V (n,m)),
        argsort selection as auxiliary_optimisers.py:60-66; best of `reps` -> candidate evaluations / s"""
        best, top = float("inf"), None
        for _ in range(reps):
            t0 = time.perf_counter()
            top = []
            for f in self.funcs:
                bv, bi = -np.inf, -1
                for off in range(0, Xc.shape[0], chunk):
                    v = np.asarray(f(Xc[off:off + chunk])).reshape(-1)
                    k = int(np.argsort(-v, kind="stable")[0])
                    if v[k] > bv:
                        bv, bi = float(v[k]), off + k
                top.append((bv, bi))
            best = min(best, time.perf_counter() - t0)
        return Xc.shape[0] * len(self.funcs) / best, top


def cpu_fit_ms(n=N_TRAIN, d=DIM):
    """one log-marginal-likelihood + gradient evaluation of the reference's CPU path on the host cores (its sklearn
    fit evaluates its restarts one after another: sklearn:_gpr.py:312-333, :584-651). -> (ms, lml, kind)"""
    X, y, theta = synthetic_problem(n, d)
    y_n = (y - y.mean()) / y.std()
    try:
        import sklearn.gaussian_process as sk_gp
        K = sk_gp.kernels
        kern = K.ConstantKernel(1.0) * K.Matern(length_scale=np.full(d, 0.5), nu=2.5) + K.WhiteKernel(1e-4)
        gp = sk_gp.GaussianProcessRegressor(kernel=kern, optimizer=None, normalize_y=False).fit(X, y_n)
        t0 = time.perf_counter()
        lml, _ = gp.log_marginal_likelihood(theta, eval_gradient=True)
        return 1e3 * (time.perf_counter() - t0), float(lml), "reference (scikit-learn log_marginal_likelihood)"
    except ImportError:
        from oracle import gp_oracle as O
        spec = O.KernelSpec(O.KIND_MATERN52, d, True)
        t0 = time.perf_counter()
        lml, _ = O.log_marginal_likelihood(spec, theta, X, y_n, eval_gradient=True)
        return 1e3 * (time.perf_counter() - t0), float(lml), "port (oracle/gp_oracle.py)"


def workload_text(cfg):
    acq = " + ".join("%s(%s,%s)" % (a[0] if a[1] == "max" or a[0] != "UCB" else "-LCB", a[1], a[2]) for a in cfg["acqs"])
    return ("GP predict + %s + argmax, n_train=%d, d=%d, ARD %s + white, fixed theta, %s candidates (%s)"
            % (acq, cfg["n"], cfg["d"], "RBF" if cfg["kind"] == KIND_RBF else "Matern-5/2", cfg["generator"],
               cfg["baseline_config"]))


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores, rank 0 only; each step
    is a bounded sample (--cpu-sample candidates) of the config's candidate set."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cfg = resolve_config(args)
    X, y, theta = synthetic_problem(cfg["n"], cfg["d"], 0, cfg["problem"], cfg["sf2"], cfg["ls"], cfg["noise"])
    arm = CpuArm(cfg, X, y, theta)
    sample = args.cpu_sample
    m_total = cfg.get("m_total") or cfg["m_per_gpu"] * max(1, args.gpus)
    Xc = step_candidates_host(cfg, m_total, sample)
    for _ in range(args.warmup):
        arm.rate(Xc[:min(sample, 2048)])
    vals, t_steps = [], []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        rate, _top = arm.rate(Xc)
        t_steps.append(time.perf_counter() - t0)
        vals.append(rate)
    value = float(np.mean(vals))
    line = {"impl": "reference", "metric": cfg["metric"], "value": value, "unit": "candidates/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(t_steps)),
            "higher_is_better": True, "scaling": cfg["scaling"], "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_text(cfg), "candidates_per_step": sample},
            "cpu_baseline": {"value": value, "unit": "candidates/s", "cores": arm.threads, "kind": arm.kind,
                             "sample": "the first %d candidates of the step's set per step, in chunks of 4096; %s"
                                       % (sample, arm.what), "model_fit_s": arm.fit_s},
            "e2e": {"value": value, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if args.cpu_fit and args.config == "headline":
        c_ms, _, kind = cpu_fit_ms()
        line["fit"] = {"cpu_lml_grad_ms_one_restart": c_ms, "cpu_cores": arm.threads, "restarts": args.restarts,
                       "cpu_kind": kind,
                       "cpu_lml_grad_ms_all_restarts_extrapolated": c_ms * args.restarts,
                       "cpu_note": "ONE LML+gradient evaluation (n=4096, d=10) timed once outside the steps; the "
                                   "reference evaluates its restarts sequentially (extrapolated, not measured)"}
    emit(line)


def parity_check(cfg, model, acq_objs, a_list, m_total, count, X, y, theta):
    """oracle vs the timed configuration on the step's own first `count` candidates.  -> dict (asserted by the caller)"""
    import torch
    from oracle import gp_oracle as O
    from turbo_b200 import _lib, device
    d, n = cfg["d"], cfg["n"]
    Xc = step_candidates_host(cfg, m_total, count)
    og = O.FittedGP(O.KernelSpec(cfg["kind"], d, True), theta, X, y)
    mu_o, sd_o = og.predict(Xc)
    var_o = sd_o ** 2
    prior = (np.exp(theta[0]) + np.exp(theta[-1])) * og.y_std ** 2
    floor = 64 * np.finfo(np.float64).eps * n * prior
    mode = _lib.CAND_UNIFORM if cfg["generator"] == "uniform" else _lib.CAND_LHS
    gen = _lib.cand_spec(mode, cfg["seed"], m_total, [0.0] * d, [1.0] * d)
    dev = device.device()
    out = {"n_checked": int(count), "floor": "64*eps*n*prior_var = %.3e" % floor, "acq": []}
    mu_rel = var_rel = var_abs = acq_rel = 0.0
    argmax_equal = var_ok = True
    for (name, ext, param), a in zip(cfg["acqs"], a_list):
        mu = torch.empty(count, dtype=torch.float64, device=dev)
        sg = torch.empty(count, dtype=torch.float64, device=dev)
        av = torch.empty(count, dtype=torch.float64, device=dev)
        a.reset()
        a.run(None, count, 0, gen, mu=mu, sigma=sg, acq=av)  # generated on the device exactly as in the timed steps
        bv, bi = a.best()
        mu, var, av = mu.cpu().numpy(), sg.cpu().numpy() ** 2, av.cpu().numpy()
        code = {"EI": O.ACQ_EI, "PI": O.ACQ_PI, "UCB": O.ACQ_UCB}[name]
        ref = O.acquisition(code, mu_o, sd_o, 1.0 if ext == "max" else -1.0, float(y.min()), param)
        k = O.argmax_first(ref)
        dv = np.abs(var - var_o)
        ok = (dv <= 1e-8 * var_o) | (dv <= floor)
        mu_rel = max(mu_rel, float(np.abs(mu - mu_o).max() / np.abs(mu_o).max()))
        var_rel = max(var_rel, float((dv / np.maximum(var_o, 1e-300))[~(dv <= floor)].max(initial=0.0)))
        var_abs = max(var_abs, float(dv.max() / floor))
        acq_rel = max(acq_rel, float(np.abs(av - ref).max() / np.abs(ref).max()))
        argmax_equal = argmax_equal and (bi == k)
        var_ok = var_ok and bool(ok.all())
        out["acq"].append({"name": name, "extremum": ext, "argmax_gpu": int(bi), "argmax_oracle": int(k),
                           "max_gpu": bv, "max_oracle": float(ref[k])})
    out.update({"mu_max_rel": mu_rel, "var_max_rel": var_rel, "var_max_abs_over_floor": var_abs,
                "acq_max_rel": acq_rel, "argmax_equal": bool(argmax_equal), "var_within_tolerance": bool(var_ok),
                "var_max_rel_note": "largest relative variance error among candidates whose absolute error exceeds the "
                                    "floor (0 = every candidate is inside the floor)"})
    return out


def run_ours(args):
    import torch
    import torch.distributed as td
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device; there is no CPU fallback (see --impl reference)"
    torch.cuda.set_device(local)
    group = None
    if world > 1:
        td.init_process_group("nccl", device_id=torch.device("cuda", local))
        group = "world"
    import turbo_b200 as tb
    from turbo_b200 import _lib, device, dist
    import ctypes as C
    lib = _lib.load()

    cfg = resolve_config(args)
    n, d = cfg["n"], cfg["d"]
    if cfg["scaling"] == "weak":
        m_total = cfg["m_per_gpu"] * world
    else:
        m_total = cfg["m_total"]
    lo_i, hi_i = dist.shard_range(m_total, rank, world)
    m = hi_i - lo_i
    X, y, theta = synthetic_problem(n, d, 0, cfg["problem"], cfg["sf2"], cfg["ls"], cfg["noise"])
    spec = tb.KernelSpec(cfg["kind"], d, True)
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0)
    t0 = time.perf_counter()
    model, _ = sur.construct_model(0, X, y)
    torch.cuda.synchronize()
    factor_ms_cold = 1e3 * (time.perf_counter() - t0)
    acq_objs = []
    for name, ext, param in cfg["acqs"]:
        if name == "EI":
            f, _ = tb.GpuEI(param).construct_function(0, model, ext, float(y.min()))
        elif name == "PI":
            f, _ = tb.GpuPI(param).construct_function(0, model, ext, float(y.min()))
        else:
            f, _ = tb.GpuUCB(param).construct_function(0, model, ext)
        acq_objs.append(f)
    a_list = [device.Acquirer(model.factor, f._spec_c(), args.batch, engine=args.engine) for f in acq_objs]
    engine, batch = a_list[0].engine, a_list[0].batch

    # synthetic candidates: this rank's contiguous slice of the global set (> 126 MB L2 for every config but cfg2)
    mode = _lib.CAND_UNIFORM if cfg["generator"] == "uniform" else _lib.CAND_LHS
    gen = _lib.cand_spec(mode, cfg["seed"], m_total, [0.0] * d, [1.0] * d)
    Xc_dev = device.generate_candidates(gen, d, lo_i, m)
    Xc_host = torch.empty((m, d), dtype=torch.float64).pin_memory()
    Xc_host.copy_(Xc_dev)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            td.barrier()
        torch.cuda.synchronize()

    def step_resident():
        res = []
        for a in a_list:
            a.reset()
            a.run(Xc_dev, m, lo_i, None)
            v, i = a.best()
            res.append(dist.allreduce_argmax(v, i, group))
        return res

    def step_e2e():
        # the plugin call a user makes: host candidates in, (value, index, x) out
        res = []
        for f in acq_objs:
            if world > 1:
                a2 = device.Acquirer(model.factor, f._spec_c(), args.batch, engine=args.engine)
                a2.run_host(Xc_host, lo_i)
                v, i = a2.best()
                res.append(dist.allreduce_argmax(v, i, group))
            else:
                v, i, _x = f.argmax_over(candidates=Xc_host, batch=args.batch, engine=args.engine)
                res.append((v, i))
        return res

    def timed(fn, steps, warmup, sample_clocks=False, time_gemm=False):
        for _ in range(warmup):
            res = fn()
        barrier()
        sampler = ClockSampler(local) if sample_clocks else None
        if sampler:
            sampler.start()
        if time_gemm:
            lib.tb200_timing_enable(1)
        l0 = device.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            res = fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = device.launch_count() - l0
        gemm = None
        if time_gemm:
            g_ms, g_n = C.c_double(), C.c_int64()
            _lib.check(lib.tb200_timing_read(C.byref(g_ms), C.byref(g_n)))
            lib.tb200_timing_enable(0)
            gemm = (g_ms.value, g_n.value)
        clocks = sampler.stop() if sampler else None
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            td.all_reduce(t, op=td.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches, gemm, clocks, res

    npass = len(a_list)
    ms, launches, _, clocks, res = timed(step_resident, args.steps, args.warmup, sample_clocks=True)
    value = npass * m_total * args.steps / (ms * 1e-3)
    # roofline pass: the same step with CUDA events around every variance-GEMM launch.  The library runs
    # this pass on ONE stream (no overlap between consecutive batches) so each launch is timed alone.
    ser_steps = max(1, min(2, args.steps))
    ms_serial, _, gemm, _, res_serial = timed(step_resident, ser_steps, 1, time_gemm=True)
    assert [r[1] for r in res_serial] == [r[1] for r in res], "multi-stream and single-stream passes disagree on the argmax"
    ms_e2e, _, _, _, res_e2e = timed(step_e2e, args.steps, max(1, args.warmup - 2))
    e2e = npass * m_total * args.steps / (ms_e2e * 1e-3)
    assert [r[1] for r in res_e2e] == [r[1] for r in res], "e2e and resident paths disagree on the argmax"

    # roofline of the dominant kernel (the variance contraction V = L^-1 k*, n^2 flop per candidate)
    dmma_tf, dfma_tf = device.fp64_peak(30.0)
    g_ms, g_n = gemm
    avg_ms = g_ms / max(g_n, 1)
    fp64_flop = float(n) * n * batch  # full batches dominate; the last partial batch is counted full
    fp64_equiv = fp64_flop / (avg_ms * 1e-3) / 1e12
    timed_in = ("a separate single-stream pass of the same step (%.1f ms/step); the headline pass overlaps consecutive "
                "batches on two internal streams" % (ms_serial / ser_steps))
    traffic = None
    if (n, d) == (N_TRAIN, DIM) and batch == 2048:
        traffic = TRAFFIC.get("dmma" if engine == "dmma" else "i8g%d" % a_list[0].gmax)
    if engine == "dmma":
        roofline = {"bound": "tensor", "achieved": fp64_equiv, "peak": dmma_tf, "unit": "TFLOP/s",
                    "frac": fp64_equiv / dmma_tf if dmma_tf else None, "traffic": traffic,
                    "kernel": "gemm_f64_kernel<false,true> (V = Linv * Kstar^T, DMMA.8x8x4, column sum-of-squares epilogue)",
                    "algorithmic_flop_per_launch": fp64_flop, "launches_timed": g_n, "avg_launch_ms": avg_ms,
                    "gemm_share_of_step": g_ms / ms_serial, "timed_in": timed_in,
                    "peak_source": "tb200_fp64_peak DMMA microbenchmark on this device (MEASURED_PEAKS.json has no "
                                   "FP64 figure); DFMA vector peak %.2f TFLOP/s" % dfma_tf}
    else:
        gmax = a_list[0].gmax
        pairs = {7: 21, 8: 28, 9: 34}[gmax]
        NP = (n + 127) // 128 * 128
        int8_ops = 2.0 * pairs * 0.5 * NP * (NP + 64) * batch  # 2 ops per INT8 MAC, lower-triangular k-range
        achieved = int8_ops / (avg_ms * 1e-3) / 1e12
        i8_burst = device.i8_peak(30.0)           # pseudo-random digits, ~30 ms per repetition
        i8_sustained = device.i8_peak(1500.0)     # the same for ~1.5 s per repetition
        i8_zero = device.i8_peak(30.0, random_digits=False)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        bf16, bf16_s = float(peaks.get("bf16_tflops", 0) or 0), float(peaks.get("bf16_tflops_sustained", 0) or 0)
        roofline = {"bound": "tensor", "achieved": achieved, "peak": i8_burst, "unit": "TFLOP/s (INT8 tensor ops)",
                    "frac": achieved / i8_burst if i8_burst else None, "traffic": traffic,
                    "kernel": "i8::colsumsq_i8_kernel<%d> (tcgen05.mma kind::i8, %d byte-slice products of V = Linv * "
                              "Kstar^T accumulated exactly in INT32 TMEM, Linv slices stacked along N, FP64 "
                              "sum-of-squares epilogue)" % (gmax, pairs),
                    "algorithmic_flop_per_launch": int8_ops, "launches_timed": g_n, "avg_launch_ms": avg_ms,
                    "gemm_share_of_step": g_ms / ms_serial, "timed_in": timed_in,
                    "fp64_equivalent_tflops": fp64_equiv, "fp64_dmma_peak_tflops": dmma_tf,
                    "peak_source": "tb200_i8_peak on this device: tcgen05.mma kind::i8 M128xN256xK32 on resident "
                                   "shared-memory operands holding pseudo-random digits, one CTA per SM, ~30 ms burst, "
                                   "best of 5 (of measured)",
                    "peak_sustained": i8_sustained, "frac_of_sustained": achieved / i8_sustained if i8_sustained else None,
                    "peak_sustained_source": "the same loop for ~1.5 s per repetition (power-capped clocks)",
                    "peak_zero_operands": i8_zero, "frac_of_zero_operand_peak": achieved / i8_zero if i8_zero else None,
                    "peak_zero_operands_note": "the same loop on all-zero operands: nothing toggles, the power limit is "
                                               "not reached, the SMs run at the maximum clock (the instruction-rate ceiling)",
                    "cross_check_2x_bf16_burst": 2.0 * bf16 or None, "cross_check_2x_bf16_sustained": 2.0 * bf16_s or None,
                    "step_level_achieved": (int8_ops * (m / batch) * npass * args.steps / (ms * 1e-3) / 1e12)}

    parity = None
    if args.parity_n > 0 and rank == 0:
        parity = parity_check(cfg, model, acq_objs, a_list, m_total, min(args.parity_n, m), X, y, theta)
    sharded = None
    if world > 1:
        ms_chk = 1 << 18
        gchk = _lib.cand_spec(mode, cfg["seed"] + 1, ms_chk, [0.0] * d, [1.0] * d)
        a = a_list[0]
        outs = {}
        for tag, r_eff in (("sharded", rank), ("reversed", world - 1 - rank)):
            lo_c, hi_c = dist.shard_range(ms_chk, r_eff, world)
            a.reset()
            a.run(None, hi_c - lo_c, lo_c, gchk)
            v, i = a.best()
            outs[tag] = dist.allreduce_argmax(v, i, group)
            outs[tag + "_owner"] = [r for r in range(world)
                                    if dist.shard_range(ms_chk, r if tag == "sharded" else world - 1 - r, world)[0]
                                    <= outs[tag][1] < dist.shard_range(ms_chk, r if tag == "sharded" else world - 1 - r, world)[1]][0]
        a.reset()
        a.run(None, ms_chk, 0, gchk)
        outs["unsharded"] = a.best()
        same = outs["sharded"] == outs["unsharded"] == outs["reversed"]
        flag = torch.tensor([1 if same else 0], dtype=torch.int64, device="cuda")
        td.all_reduce(flag, op=td.ReduceOp.MIN)
        x_row = _lib.host_candidates(mode, cfg["seed"] + 1, ms_chk, [0.0] * d, [1.0] * d, [outs["unsharded"][1]])
        sharded = {"sharded_equals_unsharded": bool(flag.item() == 1), "candidates": ms_chk,
                   "value": outs["unsharded"][0], "index": outs["unsharded"][1], "x": x_row[0].tolist(),
                   "winner_rank_in_rank_order": outs["sharded_owner"], "winner_rank_in_reverse_order": outs["reversed_owner"]}
        assert sharded["sharded_equals_unsharded"], outs

    fit = None
    if args.config == "headline" and args.fit:
        fit = fit_block(args, cfg, model, spec, theta, X, y, rank, world, group, barrier)
    elif args.config == "cfg5":
        X_dev, yn_dev = device.to_device(model.X), device.to_device(model.y_n)
        samples = []
        for _ in range(3):
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            f2 = device.Factor(X_dev, yn_dev, n, spec, theta, model.jitter)
            e1.record()
            torch.cuda.synchronize()
            samples.append(e0.elapsed_time(e1))
            del f2
        fit = {"factor_ms": min(samples[1:]), "factor_ms_samples": [round(s, 3) for s in samples],
               "what": "Gram + Cholesky + inverse factor + alpha + LML of one n=%d matrix (tb200_factor), replicated "
                       "per GPU ('replicas only', SURVEY 8e)" % n,
               "cholesky_flop": float(n) ** 3 / 3.0}

    cpu = None
    if rank == 0 and world == 1 and args.cpu_sample > 0:
        arm = CpuArm(cfg, X, y, theta)
        cs = min(args.cpu_sample, m)
        rate, top = arm.rate(step_candidates_host(cfg, m_total, cs), reps=2)
        cpu = {"value": rate, "unit": "candidates/s", "cores": arm.threads, "kind": arm.kind,
               "sample": "the first %d candidates of the step's own set, in chunks of 4096, best of 2; %s"
                         % (cs, arm.what), "model_fit_s": arm.fit_s}
        if fit is not None and args.cpu_fit and args.config == "headline":
            c_ms, _, kind = cpu_fit_ms(n, d)
            fit.update({"cpu_lml_grad_ms_one_restart": c_ms, "cpu_cores": arm.threads, "cpu_kind": kind,
                        "cpu_lml_grad_ms_all_restarts_extrapolated": c_ms * args.restarts,
                        "cpu_note": "ONE LML+gradient evaluation timed on the host cores; the reference evaluates its "
                                    "restarts sequentially, so R of them are R times that (extrapolated, not measured)"})
    if rank == 0:
        line = {"metric": cfg["metric"], "value": value, "unit": "candidates/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": cfg["scaling"], "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_text(cfg), "name": args.config,
                           "candidates_total_per_pass": m_total, "candidates_per_gpu_per_pass": m,
                           "acquisition_passes_per_step": npass, "candidate_batch": batch, "variance_engine": engine,
                           "slice_products": None if engine == "dmma" else {7: 21, 8: 28, 9: 34}[a_list[0].gmax],
                           "slice_cutoff_probe": None if engine != "i8a" else
                           "gmax %d chosen for this model: 21- and 34-product variances differ by %.2f units of "
                           "eps*n*prior_var on 256 probe points (threshold %.0f, parity floor 64)"
                           % (a_list[0].gmax, model.factor.gmax_probe, model.factor.GMAX7_UNITS),
                           "l2": "candidate set %.0f MB per GPU vs 126 MB L2" % (m * d * 8 / 1e6),
                           "parallelism": "candidates sharded x%d, model replicated" % world},
                "e2e": {"value": e2e, "unit": "candidates/s", "h2d_bytes_per_step": npass * m * d * 8 * world,
                        "d2h_bytes_per_step": npass * 16 * world, "ms_per_step": ms_e2e / args.steps},
                "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
                "parity_check": parity, "sharded_check": sharded, "fit": fit,
                "argmax": [{"value": r[0], "index": r[1]} for r in res], "factor_ms_cold": factor_ms_cold}
        emit(line)
        if parity is not None:
            assert parity["mu_max_rel"] <= 1e-8 and parity["acq_max_rel"] <= 1e-8 and parity["argmax_equal"] \
                and parity["var_within_tolerance"], "parity_check failed: %r" % (parity,)
    if world > 1:
        td.destroy_process_group()


def fit_block(args, cfg, model, spec, theta, X, y, rank, world, group, barrier):
    """GP fit ms (configs[2]): one batched LML+gradient evaluation over R restarts (split across ranks) + a whole fit"""
    import torch
    import torch.distributed as td
    import turbo_b200 as tb
    from turbo_b200 import device
    n, d = cfg["n"], cfg["d"]
    R = args.restarts
    Rl = len(range(rank, R, world))
    rng = np.random.default_rng(2)
    thetas = np.clip(theta + rng.normal(0, 0.5, (R, theta.size)), np.log(1e-2) * 2, np.log(1e2))[rank::world]
    X_dev, yn_dev = device.to_device(model.X), device.to_device(model.y_n)
    fit_ms = []
    lml = None
    for it in range(6):  # one warm-up, then the best of five (the first samples after the acquisition loop run under its power cap)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        lml, grad, info = device.lml_batched(X_dev, yn_dev, n, spec, thetas, model.jitter, True)
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            td.all_reduce(t, op=td.ReduceOp.MAX)
        fit_ms.append(float(t.item()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    f2 = device.Factor(X_dev, yn_dev, n, spec, theta, model.jitter)
    e1.record()
    torch.cuda.synchronize()
    del f2
    flop = float(n) ** 3 + float(n) ** 2 * (3 * d + 42)
    fit = {"lml_grad_batched_ms": min(fit_ms[1:]), "lml_grad_batched_ms_samples": [round(v, 3) for v in fit_ms[1:]],
           "restarts": R, "restarts_per_gpu": Rl,
           "engine": "i8 (potrf column updates, trtri, K^-1 on tcgen05 kind::i8, 28/21/21 slice products) when the batch is "
                     "large enough for the left-looking schedule, else FP64 DMMA; TB200_DEBUG=fit_engine=dmma forces DMMA",
           "flop_per_evaluation": flop, "tflops": Rl * flop / (min(fit_ms[1:]) * 1e-3) / 1e12,
           "final_factor_ms": e0.elapsed_time(e1), "all_finite": bool(np.isfinite(lml).all())}
    # a whole hyper-parameter fit: R restarts from the same starting points, at most 30 lock-step L-BFGS-B iterations
    # (SURVEY 8d cfg3), restarts sharded over the ranks, best restart selected with one all-gather
    if args.full_fit:
        all_thetas = np.clip(theta + np.random.default_rng(2).normal(0, 0.5, (R, theta.size)), np.log(1e-2) * 2,
                             np.log(1e2))
        sur2 = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=R,
                                 max_iter=30, seed=0, group=group)
        barrier()
        t0 = time.perf_counter()
        _m2, finfo = sur2.construct_model(1, X, y, start_thetas=all_thetas)
        barrier()
        full_ms = 1e3 * (time.perf_counter() - t0)
        if world > 1:
            t = torch.tensor([full_ms], dtype=torch.float64, device="cuda")
            td.all_reduce(t, op=td.ReduceOp.MAX)
            full_ms = float(t.item())
        fit.update({"full_fit_ms": full_ms, "full_fit_max_iter": 30, "full_fit_lml_evaluations": finfo["lml_evaluations"],
                    "full_fit_best_neg_lml": finfo["best_neg_lml"],
                    "full_fit_ms_per_evaluation": full_ms / max(1, finfo["lml_evaluations"]),
                    "full_fit_note": "wall clock incl. host L-BFGS-B logic and the final factorisation; evaluations "
                                     "cover only the still-active restarts"})
        del _m2
    return fit


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="headline", choices=sorted(CONFIGS))
    ap.add_argument("--m", type=int, default=0, help="override: candidates per GPU (weak configs) / in total (strong)")
    ap.add_argument("--n", type=int, default=0, help="override the config's n_train")
    ap.add_argument("--d", type=int, default=0, help="override the config's dimension")
    ap.add_argument("--batch", type=int, default=0, help="candidate batch per kernel pass (0 = library default)")
    ap.add_argument("--restarts", type=int, default=64)
    ap.add_argument("--cpu-sample", type=int, default=1 << 14)
    ap.add_argument("--parity-n", type=int, default=1 << 14, help="candidates of the step checked against the oracle (0 = off)")
    ap.add_argument("--cpu-fit", type=int, default=1, help="also time one LML+gradient evaluation of the CPU path")
    ap.add_argument("--engine", default=None, help="variance engine: auto (default) | dmma | i8a | i8l | i8 | i8x")
    ap.add_argument("--fit", type=int, default=1, help="headline: also time the batched fit")
    ap.add_argument("--full-fit", type=int, default=1, help="also time one whole fit (R restarts, <= 30 iterations)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
