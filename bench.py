#!/usr/bin/env python
"""bench.py -- EI candidates/sec (n_train=4096, d=10) + GP fit ms, per BASELINE.json.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over one batch of synthetic candidates: predict + EI + argmax
over `--m` candidates PER GPU (weak scaling; no data-path collective, one 16-byte all-gather per step)
on a GP with n_train=4096, d=10, ARD Matern-5/2 + white noise at fixed hyper-parameters.

  value     : candidates/s, candidates already resident in HBM (device-timed, max over ranks)
  e2e       : same metric through the plugin call acq.argmax_over(host candidates): pinned host
              buffer -> chunked H2D inside the timed region -> (value, index) read back
  roofline  : the variance contraction V = L^-1 k* (n^2 flop per candidate).  Default engine: INT8 tensor cores
              (tcgen05.mma kind::i8, exact byte-slice products) -- algorithmic INT8 ops / CUDA-event duration against
              2 x the measured bf16 peak; --engine dmma: FP64 DMMA GEMM against the FP64 tensor peak measured on this
              device by tb200_fp64_peak
  cpu_baseline : the oracle port (NumPy/SciPy, same LAPACK/BLAS calls as the reference's sklearn path)
              on a bounded sample of the same workload, rank 0, N=1 only
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
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
# `ncu --set full` captures under profiles/ (None until captured for that engine)
TRAFFIC = {"dmma": 197.2e6, "i8": 163.9e6, "i8x": None}  # profiles/r1_gemm_ncu_full.md, r1_i8v2_ncu_full.md
METRIC = "EI candidates/sec (n=4096,d=10)"


def synthetic_problem(n=N_TRAIN, d=DIM, seed=0):
    """SURVEY 8d cfg3/cfg4 data: y = sum sin(2 pi x_j) + 0.01 N(0,1); fixed ARD Matern-5/2 theta"""
    rng = np.random.default_rng(seed)
    X = rng.random((n, d))
    y = np.sin(2 * np.pi * X).sum(axis=1) + 0.01 * rng.standard_normal(n)
    theta = np.log(np.r_[1.0, np.full(d, 0.5), 1e-4])
    return X, y, theta


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


def cpu_reference_rate(sample, reps, n=N_TRAIN, d=DIM):
    """the oracle port timed on the host cores: predict + EI + argmax over `sample` candidates in
    chunks of 4096 (sklearn materialises K* and V the same way); best of `reps`. -> (cand/s, threads)"""
    from oracle import gp_oracle as O
    # torchrun exports OMP_NUM_THREADS=1 to every rank: lift the BLAS pools back to all host cores so the
    # CPU figure is the same with and without torchrun
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    try:
        from threadpoolctl import threadpool_info, threadpool_limits
        threadpool_limits(limits=cores)
        threads = max([p.get("num_threads", 1) for p in threadpool_info()] + [1])
    except Exception:
        threads = 1
    X, y, theta = synthetic_problem(n, d)
    gp = O.FittedGP(O.KernelSpec(O.KIND_MATERN52, d, True), theta, X, y)
    Xc = np.random.default_rng(1).random((sample, d))
    inc = float(y.min())
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        top = (-np.inf, -1)
        for off in range(0, sample, 4096):
            mu, sd = gp.predict(Xc[off:off + 4096])
            ei = O.acquisition(O.ACQ_EI, mu, sd, -1.0, inc, 0.01)
            k = O.argmax_first(ei)
            if ei[k] > top[0]:
                top = (float(ei[k]), off + k)
        best = min(best, time.perf_counter() - t0)
    return sample / best, threads


def cpu_fit_ms(n=N_TRAIN, d=DIM):
    """one log-marginal-likelihood + gradient evaluation of the oracle port on the host cores (the reference's sklearn
    fit evaluates its restarts one after another: sklearn:_gpr.py:312-333, :584-651). -> (ms, lml)"""
    from oracle import gp_oracle as O
    X, y, theta = synthetic_problem(n, d)
    y_n = (y - y.mean()) / y.std()
    spec = O.KernelSpec(O.KIND_MATERN52, d, True)
    t0 = time.perf_counter()
    lml, _ = O.log_marginal_likelihood(spec, theta, X, y_n, eval_gradient=True)
    return 1e3 * (time.perf_counter() - t0), float(lml)


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores.
    The reference is pure Python over scikit-learn (nothing compiles into oracle/_ref), so this times
    the oracle port; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sample = args.cpu_sample
    vals = []
    for _ in range(args.warmup):
        cpu_reference_rate(min(sample, 4096), 1)
    t_steps = []
    threads = 1
    for _ in range(args.steps):
        t0 = time.perf_counter()
        rate, threads = cpu_reference_rate(sample, 1)
        t_steps.append(time.perf_counter() - t0)
        vals.append(rate)
    value = float(np.mean(vals))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "candidates/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(t_steps)),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "GP predict + EI + argmax, n_train=4096, d=10, ARD Matern-5/2 + white, fixed theta",
                       "candidates_per_step": sample},
            "cpu_baseline": {"value": value, "unit": "candidates/s", "cores": threads, "kind": "port",
                             "sample": "%d candidates per step in chunks of 4096, oracle/gp_oracle.py "
                                       "(NumPy/SciPy; same LAPACK/BLAS calls as the reference's sklearn path)" % sample},
            "e2e": {"value": value, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if args.cpu_fit:
        c_ms, _ = cpu_fit_ms()
        line["fit"] = {"cpu_lml_grad_ms_one_restart": c_ms, "cpu_cores": threads, "restarts": args.restarts,
                       "cpu_lml_grad_ms_all_restarts_extrapolated": c_ms * args.restarts,
                       "cpu_note": "ONE LML+gradient evaluation (n=4096, d=10) timed once outside the steps; the "
                                   "reference evaluates its restarts sequentially (extrapolated, not measured)"}
    emit(line)


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

    n, d, m = args.n, args.d, args.m
    X, y, theta = synthetic_problem(n, d)
    spec = tb.KernelSpec(_lib.KIND_MATERN52, d, True)
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0)
    t0 = time.perf_counter()
    model, _ = sur.construct_model(0, X, y)
    torch.cuda.synchronize()
    factor_ms_cold = 1e3 * (time.perf_counter() - t0)
    acq, _ = tb.GpuEI(0.01).construct_function(0, model, "min", float(y.min()))
    a = device.Acquirer(model.factor, acq._spec_c(), args.batch, engine=args.engine)
    engine = a.engine

    # synthetic candidates: this rank's slice of a global uniform set (335 MB at m=2^22 > 126 MB L2)
    gen = _lib.cand_spec(_lib.CAND_UNIFORM, 7, m * world, [0.0] * d, [1.0] * d)
    Xc_dev = device.generate_candidates(gen, d, rank * m, m)
    Xc_host = torch.empty((m, d), dtype=torch.float64).pin_memory()
    Xc_host.copy_(Xc_dev)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            td.barrier()
        torch.cuda.synchronize()

    def step_resident():
        a.reset()
        a.run(Xc_dev, m, rank * m, None)
        v, i = a.best()
        return dist.allreduce_argmax(v, i, group)

    def step_e2e():
        # the plugin call a user makes: host candidates in, (value, index, x) out
        if world > 1:
            a2 = device.Acquirer(model.factor, acq._spec_c(), args.batch, engine=args.engine)
            copy_stream, main = torch.cuda.Stream(), torch.cuda.current_stream()
            for off in range(0, m, 1 << 18):
                cnt = min(1 << 18, m - off)
                with torch.cuda.stream(copy_stream):
                    buf = Xc_host[off:off + cnt].to(device.device(), non_blocking=True)
                    ev = torch.cuda.Event(); ev.record(copy_stream)
                main.wait_event(ev)
                a2.run(buf, cnt, rank * m + off, None)
                buf.record_stream(main)
            v, i = a2.best()
            return dist.allreduce_argmax(v, i, group)
        v, i, _x = acq.argmax_over(candidates=Xc_host, batch=args.batch, engine=args.engine)
        return v, i

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

    ms, launches, _, clocks, res = timed(step_resident, args.steps, args.warmup, sample_clocks=True)
    value = m * world * args.steps / (ms * 1e-3)
    # roofline pass: the same step with CUDA events around every variance-GEMM launch.  The library runs
    # this pass on ONE stream (no overlap between consecutive batches) so each launch is timed alone.
    ms_serial, _, gemm, _, res_serial = timed(step_resident, max(1, min(2, args.steps)), 1, time_gemm=True)
    assert res_serial[1] == res[1], "two-stream and single-stream passes disagree on the argmax"
    ms_e2e, _, _, _, res_e2e = timed(step_e2e, args.steps, max(1, args.warmup - 2))
    e2e = m * world * args.steps / (ms_e2e * 1e-3)
    assert res_e2e[1] == res[1], "e2e and resident paths disagree on the argmax"

    # roofline of the dominant kernel (the variance contraction V = L^-1 k*, n^2 flop per candidate)
    dmma_tf, dfma_tf = device.fp64_peak(30.0)
    batch = a.batch
    g_ms, g_n = gemm
    avg_ms = g_ms / max(g_n, 1)
    fp64_flop = float(n) * n * batch  # full batches dominate; the last partial batch is counted full
    fp64_equiv = fp64_flop / (avg_ms * 1e-3) / 1e12
    timed_in = ("a separate single-stream pass of the same step (%.1f ms/step); the headline pass overlaps consecutive "
                "batches on two internal streams" % (ms_serial / max(1, min(2, args.steps))))
    if engine == "dmma":
        roofline = {"bound": "tensor", "achieved": fp64_equiv, "peak": dmma_tf, "unit": "TFLOP/s",
                    "frac": fp64_equiv / dmma_tf if dmma_tf else None, "traffic": TRAFFIC.get("dmma"),
                    "kernel": "gemm_f64_kernel<false,true> (V = Linv * Kstar^T, DMMA.8x8x4, column sum-of-squares epilogue)",
                    "algorithmic_flop_per_launch": fp64_flop, "launches_timed": g_n, "avg_launch_ms": avg_ms,
                    "gemm_share_of_step": g_ms / ms_serial, "timed_in": timed_in,
                    "peak_source": "tb200_fp64_peak DMMA microbenchmark on this device (MEASURED_PEAKS.json has no "
                                   "FP64 figure); DFMA vector peak %.2f TFLOP/s" % dfma_tf}
    else:
        pairs = 28 if engine == "i8" else 34
        NP = (n + 127) // 128 * 128
        int8_ops = 2.0 * pairs * 0.5 * NP * (NP + 64) * batch  # 2 ops per INT8 MAC, lower-triangular k-range
        achieved = int8_ops / (avg_ms * 1e-3) / 1e12
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        bf16 = float(peaks.get("bf16_tflops", 0) or 0)
        peak = 2.0 * bf16 if bf16 else 4500.0
        roofline = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s (INT8 tensor ops)",
                    "frac": achieved / peak, "traffic": TRAFFIC.get(engine),
                    "kernel": "i8::colsumsq_i8_kernel<%d> (tcgen05.mma kind::i8, %d byte-slice products of V = Linv * "
                              "Kstar^T accumulated exactly in INT32 TMEM, Linv slices stacked along N: %d MMAs per K=32 "
                              "step, FP64 sum-of-squares epilogue)"
                              % (8 if engine == "i8" else 9, pairs, 10 if engine == "i8" else 11),
                    "algorithmic_flop_per_launch": int8_ops, "launches_timed": g_n, "avg_launch_ms": avg_ms,
                    "gemm_share_of_step": g_ms / ms_serial, "timed_in": timed_in,
                    "fp64_equivalent_tflops": fp64_equiv, "fp64_dmma_peak_tflops": dmma_tf,
                    "peak_source": ("2 x the measured dense bf16 burst peak in MEASURED_PEAKS.json (%.1f TFLOP/s): INT8 "
                                    "tcgen05 MMAs retire K=32 per instruction where bf16 retires K=16 (no INT8 figure "
                                    "is measured by the driver)" % bf16) if bf16 else
                                   "nominal 4.5 PFLOP/s dense INT8 (B200_PROFILING.md fallback; MEASURED_PEAKS.json absent)"}

    # GP fit ms: one batched LML+gradient evaluation over R restarts (split across ranks)
    R = args.restarts
    Rl = len(range(rank, R, world))
    rng = np.random.default_rng(2)
    thetas = np.clip(theta + rng.normal(0, 0.5, (R, theta.size)), np.log(1e-2) * 2, np.log(1e2))[rank::world]
    X_dev, yn_dev = device.to_device(model.X), device.to_device(model.y_n)
    fit_ms = []
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
    fit = {"lml_grad_batched_ms": min(fit_ms[1:]), "lml_grad_batched_ms_samples": [round(v, 3) for v in fit_ms[1:]], "restarts": R, "restarts_per_gpu": Rl,
           "engine": "i8 (potrf column updates, trtri, K^-1 on tcgen05 kind::i8, 28/28/21 slice products) when the batch is "
                     "large enough for the left-looking schedule, else FP64 DMMA; TB200_FIT_ENGINE=dmma forces DMMA",
           "flop_per_evaluation": float(n) ** 3 + float(n) ** 2 * (3 * d + 42),
           "tflops": Rl * (float(n) ** 3 + float(n) ** 2 * (3 * d + 42)) / (min(fit_ms[1:]) * 1e-3) / 1e12,
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
                    "full_fit_note": "wall clock incl. host L-BFGS-B logic and the final factorisation; evaluations "
                                     "cover only the still-active restarts"})
        del _m2

    cpu = None
    if rank == 0 and world == 1 and args.cpu_sample > 0:
        rate, threads = cpu_reference_rate(args.cpu_sample, 2)
        cpu = {"value": rate, "unit": "candidates/s", "cores": threads, "kind": "port",
               "sample": "%d of the step's candidates in chunks of 4096, best of 2; oracle/gp_oracle.py "
                         "(NumPy/SciPy port of the reference's sklearn path)" % args.cpu_sample}
        if args.cpu_fit:
            c_ms, _ = cpu_fit_ms(n, d)
            fit.update({"cpu_lml_grad_ms_one_restart": c_ms, "cpu_cores": threads,
                        "cpu_lml_grad_ms_all_restarts_extrapolated": c_ms * R,
                        "cpu_note": "ONE LML+gradient evaluation of the oracle port timed on the host cores; the "
                                    "reference evaluates its restarts sequentially, so R of them are R times that "
                                    "(extrapolated, not measured)"})
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "candidates/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "GP predict + EI + argmax, n_train=%d, d=%d, ARD Matern-5/2 + white, fixed "
                                       "theta (BASELINE.json headline; configs[3] shape)" % (n, d),
                           "candidates_per_gpu_per_step": m, "candidate_batch": batch, "variance_engine": engine,
                           "l2": "candidate set %.0f MB per GPU > 126 MB L2" % (m * d * 8 / 1e6),
                           "parallelism": "candidates sharded x%d, model replicated" % world},
                "e2e": {"value": e2e, "unit": "candidates/s", "h2d_bytes_per_step": m * d * 8 * world,
                        "d2h_bytes_per_step": 16 * world, "ms_per_step": ms_e2e / args.steps},
                "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "fit": fit,
                "argmax": {"value": res[0], "index": res[1]}, "factor_ms_cold": factor_ms_cold}
        emit(line)
    if world > 1:
        td.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--m", type=int, default=1 << 21, help="candidates per GPU per step")
    ap.add_argument("--n", type=int, default=N_TRAIN)
    ap.add_argument("--d", type=int, default=DIM)
    ap.add_argument("--batch", type=int, default=0, help="candidate batch per kernel pass (0 = library default)")
    ap.add_argument("--restarts", type=int, default=64)
    ap.add_argument("--cpu-sample", type=int, default=1 << 14)
    ap.add_argument("--cpu-fit", type=int, default=1, help="also time one LML+gradient evaluation of the CPU port")
    ap.add_argument("--engine", default=None, help="variance engine: auto (default) | dmma | i8 | i8x")
    ap.add_argument("--full-fit", type=int, default=1, help="also time one whole fit (R restarts, <= 30 iterations)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
