"""Thin wrappers that call the C ABI on PyTorch device buffers (the buffer/stream layer only).

Every function raises if CUDA or libturbo_b200.so is unavailable; there is no CPU path.
"""
import ctypes as C
import math

import numpy as np
import torch

from . import _lib


def _require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("turbo_b200 needs a CUDA device (sm_100a); there is no CPU fallback")


def device():
    _require_cuda()
    return torch.device("cuda", torch.cuda.current_device())


def to_device(a):
    """host float64 array -> device tensor (pinned staging for large arrays)"""
    a = np.ascontiguousarray(a, dtype=np.float64)
    t = torch.from_numpy(a)
    if a.nbytes >= (1 << 20):
        t = t.pin_memory()
    return t.to(device(), non_blocking=True)


def _workspace(nbytes):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device())


def padded_n(n):
    return _lib.load().tb200_padded_n(int(n))


def lml_batched(X_dev, yn_dev, n, spec, thetas, jitter, with_grad=True):
    """log-marginal likelihood (+ gradient) of R theta rows in one device call (several when the R working sets --
    about 23 n^2 bytes each -- do not fit in the free device memory: 64 restarts at n = 16384 would need 390 GB).
    -> (lml (R,), grad (R,p) or None, info (R,)) as NumPy."""
    lib = _lib.load()
    thetas = np.ascontiguousarray(np.atleast_2d(thetas), dtype=np.float64)
    R, p = thetas.shape
    assert p == spec.p
    chunk = R
    need = lib.tb200_fit_workspace_bytes(n, spec.d, R, int(with_grad))
    # the driver is asked for the free memory only when the working set is a large part of the device: cudaMemGetInfo
    # on the call path of every evaluation showed up as 2x outliers in the 8-process timing of the fit
    if need > 0.2 * torch.cuda.get_device_properties(torch.cuda.current_device()).total_memory:
        free_bytes = torch.cuda.mem_get_info()[0] + torch.cuda.memory_reserved() - torch.cuda.memory_allocated()
        while chunk > 1 and lib.tb200_fit_workspace_bytes(n, spec.d, chunk, int(with_grad)) > 0.8 * free_bytes:
            chunk = (chunk + 1) // 2
    if chunk < R:
        parts = [lml_batched(X_dev, yn_dev, n, spec, thetas[o:o + chunk], jitter, with_grad) for o in range(0, R, chunk)]
        return (np.concatenate([q[0] for q in parts]),
                np.concatenate([q[1] for q in parts]) if with_grad else None,
                np.concatenate([q[2] for q in parts]))
    dev = device()
    th = torch.from_numpy(thetas).to(dev)
    lml = torch.empty(R, dtype=torch.float64, device=dev)
    grad = torch.empty((R, p), dtype=torch.float64, device=dev) if with_grad else None
    info = torch.empty(R, dtype=torch.int32, device=dev)
    nbytes = lib.tb200_fit_workspace_bytes(n, spec.d, R, int(with_grad))
    work = _workspace(nbytes)
    cs = spec.to_c()
    _lib.check(lib.tb200_lml_batched(_lib.ptr(X_dev), _lib.ptr(yn_dev), n, C.byref(cs), _lib.ptr(th), R,
                                     float(jitter), int(with_grad), _lib.ptr(lml), _lib.ptr(grad), _lib.ptr(info),
                                     _lib.ptr(work), nbytes, _lib.stream_ptr()))
    out = (lml.cpu().numpy(), grad.cpu().numpy() if with_grad else None, info.cpu().numpy())
    del work
    return out


class Factor:
    """device-resident state predict/acquire need: XsT, Linv, alpha (+ theta), see tb200_factor"""

    def __init__(self, X_dev, yn_dev, n, spec, theta, jitter):
        lib = _lib.load()
        dev = device()
        self.n, self.spec = int(n), spec
        NP = padded_n(n)
        self.NP = NP
        self.theta = torch.from_numpy(np.ascontiguousarray(theta, dtype=np.float64)).to(dev)
        self.XsT = torch.empty((spec.d, NP), dtype=torch.float64, device=dev)
        self.Linv = torch.empty((NP, NP), dtype=torch.float64, device=dev)
        self.alpha = torch.empty(NP, dtype=torch.float64, device=dev)
        self.Ldiag = torch.empty(NP, dtype=torch.float64, device=dev)
        lml = torch.empty(1, dtype=torch.float64, device=dev)
        info = torch.empty(1, dtype=torch.int32, device=dev)
        nbytes = lib.tb200_fit_workspace_bytes(n, spec.d, 1, 0)
        work = _workspace(nbytes)
        cs = spec.to_c()
        _lib.check(lib.tb200_factor(_lib.ptr(X_dev), _lib.ptr(yn_dev), n, C.byref(cs), _lib.ptr(self.theta),
                                    float(jitter), _lib.ptr(self.XsT), _lib.ptr(self.Linv), _lib.ptr(self.alpha),
                                    _lib.ptr(self.Ldiag), _lib.ptr(lml), _lib.ptr(info), _lib.ptr(work), nbytes,
                                    _lib.stream_ptr()))
        self.info = int(info.item())
        self.lml = float(lml.item())
        del work

        self.X_dev = X_dev  # unscaled training inputs (n, d): the probe of auto_gmax() is built from them
        self.linvS = self.rowscale = None
        self.gmax_auto = self.gmax_probe = None

    def prepare_i8(self):
        """byte-slice the inverse factor for the INT8 tensor-core variance engine (once per model)"""
        if self.linvS is None:
            lib = _lib.load()
            dev = device()
            self.linvS = torch.empty(int(lib.tb200_i8_factor_bytes(self.n)), dtype=torch.uint8, device=dev)
            self.rowscale = torch.empty(2 * self.NP, dtype=torch.float64, device=dev)
            cs = self.spec.to_c()
            _lib.check(lib.tb200_i8_prepare(_lib.ptr(self.Linv), self.n, C.byref(cs), _lib.ptr(self.theta),
                                            _lib.ptr(self.linvS), _lib.ptr(self.rowscale), _lib.stream_ptr()))
        return self

    # slice-pair cut-off of the INT8 variance engine, chosen PER MODEL from a measurement on the model itself.
    # 21 products (gmax 7, six digit planes) truncate every term of V = Linv k* at ~2^-46 of (row scale x sf2) instead of
    # ~2^-54 (gmax 8).  profiles/r2_predict_precision.md: against the oracle that is 2.7-5.2 units of eps*n*prior_var at
    # the BASELINE shapes (the parity floor is 64 of them, the same constant the oracle itself needs against the
    # reference), EI within 3e-11, argmax identical -- and 35 units on an ill-conditioned factor (noise 1e-8).  So the
    # cheaper cut-off is used only where THIS factor shows it is safe: 256 probe points -- 128 training points (where
    # k** - v.v cancels completely and |v| is largest) and 128 midpoints of neighbouring training points -- are
    # evaluated with 21 and with 34 products; gmax 7 is taken when the two variances differ by at most GMAX7_UNITS
    # (= 1/8 of the floor), otherwise gmax 8.
    GMAX7_UNITS = 8.0

    def auto_gmax(self):
        if self.gmax_auto is None:
            self.prepare_i8()
            lib, dev = _lib.load(), self.X_dev.device
            n, d = self.n, self.spec.d
            P = min(n, 128)
            probe = torch.empty((2 * P, d), dtype=torch.float64, device=dev)
            _lib.check(lib.tb200_i8_probe_points(_lib.ptr(self.X_dev), n, d, P, _lib.ptr(probe), _lib.stream_ptr()))
            th = self.theta.cpu().numpy()
            prior = (math.exp(th[0]) if self.spec.has_const else 1.0) + (math.exp(th[-1]) if self.spec.has_white else 0.0)
            sg = {}
            for eng in ("i8l", "i8x"):
                sg[eng] = torch.empty(2 * P, dtype=torch.float64, device=dev)
                a = Acquirer(self, acq_spec(_lib.ACQ_UCB, 1.0, 0.0, 0.0, 0.0, 1.0), batch=256, engine=eng)
                a.run(probe, 2 * P, 0, None, sigma=sg[eng])
            worst = torch.empty(1, dtype=torch.float64, device=dev)
            _lib.check(lib.tb200_max_sqdiff(_lib.ptr(sg["i8l"]), _lib.ptr(sg["i8x"]), 2 * P, _lib.ptr(worst),
                                            _lib.stream_ptr()))
            unit = np.finfo(np.float64).eps * n * prior
            self.gmax_probe = float(worst.item()) / unit
            self.gmax_auto = 7 if self.gmax_probe <= self.GMAX7_UNITS else 8
        return self.gmax_auto

    def append(self, X_new, yn_dev, n_total, jitter):
        """incremental refit at fixed theta (tb200_factor_append): append the rows of X_new (k,d) one by
        one, then recompute alpha and the LML for the n_total (re-normalised) targets in yn_dev.  Returns False --
        with the factor unchanged in every way that matters -- if the padding is exhausted or a pivot fails."""
        lib = _lib.load()
        X_new = np.ascontiguousarray(np.atleast_2d(X_new), dtype=np.float64)
        k = X_new.shape[0]
        if self.n + k != n_total or n_total > self.NP or self.info != 0:
            return False
        dev = device()
        xd = torch.from_numpy(X_new).to(dev)
        info = torch.zeros(1, dtype=torch.int32, device=dev)
        nbytes = lib.tb200_append_workspace_bytes(n_total)
        work = _workspace(nbytes)
        cs = self.spec.to_c()
        n0 = self.n
        for i in range(k):
            _lib.check(lib.tb200_factor_append(_lib.ptr(xd[i:]), _lib.ptr(self.XsT), _lib.ptr(self.Linv),
                                               _lib.ptr(self.Ldiag), n0 + i, C.byref(cs), _lib.ptr(self.theta),
                                               float(jitter), _lib.ptr(info), _lib.ptr(work), nbytes, _lib.stream_ptr()))
            if int(info.item()) != 0:
                # rows n0..n0+i-1 were appended: put the identity padding back so the factor is the old one again
                for r in range(n0, n0 + i):
                    self.Linv[r].zero_()
                    self.Linv[r, r] = 1.0
                    self.XsT[:, r] = 0.0
                return False
        self.n = n_total
        self.X_dev = torch.cat([self.X_dev[:n0], xd])
        self.gmax_auto = self.gmax_probe = None
        lml = torch.empty(1, dtype=torch.float64, device=dev)
        _lib.check(lib.tb200_alpha_lml(_lib.ptr(yn_dev), n_total, _lib.ptr(self.Linv), _lib.ptr(self.Ldiag),
                                       _lib.ptr(self.alpha), _lib.ptr(lml), _lib.ptr(work), nbytes, _lib.stream_ptr()))
        self.lml = float(lml.item())
        self.linvS = self.rowscale = None  # the byte-slice planes are rebuilt on the next acquisition (~0.1 ms)
        del work
        return True

    def nbytes(self):
        ts = [self.XsT, self.Linv, self.alpha, self.Ldiag] + ([self.linvS, self.rowscale] if self.linvS is not None else [])
        return sum(t.numel() * t.element_size() for t in ts)


I8_GMAX = {"i8l": 7, "i8": 8, "i8x": 9}  # slice-pair cut-off of the INT8 variance engine: 21 / 28 / 34 products


def resolve_engine(engine, NP):
    """'auto' -> the INT8 tensor-core engine for n <= I8_MAX_N (INT32 accumulation is exact per pass of 8192 training
    points; longer rows take several passes) with the cut-off Factor.auto_gmax() picks for the model ('i8a');
    'i8l' / 'i8' / 'i8x' fix it at 21 / 28 / 34 products.  'dmma' is the FP64 DMMA GEMM (small-n path and cross-check)."""
    engine = (engine or "auto").lower()
    if engine == "auto":
        engine = "i8a" if NP <= _lib.I8_MAX_N else "dmma"
    if engine in I8_GMAX or engine == "i8a":
        if NP > _lib.I8_MAX_N:
            raise ValueError("the i8 variance engine needs n <= %d; use engine='dmma'" % _lib.I8_MAX_N)
        return engine
    if engine != "dmma":
        raise ValueError("unknown variance engine %r (auto | dmma | i8a | i8l | i8 | i8x)" % engine)
    return engine


class Acquirer:
    """runs tb200_acquire / tb200_acquire_i8 over device candidates or a generated set, keeping the running
    best on device.  engine: 'dmma' (FP64 tensor-core GEMM), 'i8' (INT8 tensor cores, 28 slice products),
    'i8x' (34 slice products), 'i8l' (21 slice products), 'auto' (i8 when n <= I8_MAX_N = 32768)."""

    def __init__(self, factor, acq_c, batch=0, engine=None, exclusion=None):
        """exclusion: None or (X_dev (n,d) device tensor of UNSCALED points, tol): candidates whose squared distance to
        a row of X is <= tol are never selected (tb200_exclusion; turbo/utils.py:111-130)"""
        lib = _lib.load()
        self.lib, self.f, self.acq_c = lib, factor, acq_c
        self.excl_c, self._excl_X = None, None
        if exclusion is not None and exclusion[1] >= 0 and exclusion[0].shape[0] > 0:
            self._excl_X = exclusion[0].contiguous()
            assert self._excl_X.shape[1] == factor.spec.d, 'different number of attributes'
            self.excl_c = _lib.ExclusionC(_lib.ptr(self._excl_X), int(self._excl_X.shape[0]), float(exclusion[1]), None)
        self.engine = resolve_engine(engine, factor.NP)
        self.batch = int(batch) if batch else int(lib.tb200_default_batch(factor.n))
        if self.engine == "dmma":
            self.nbytes = lib.tb200_acquire_workspace_bytes(factor.n, factor.spec.d, self.batch)
        else:
            factor.prepare_i8()
            self.gmax = factor.auto_gmax() if self.engine == "i8a" else I8_GMAX[self.engine]
            self.nbytes = lib.tb200_acquire_i8_workspace_bytes(factor.n, factor.spec.d, self.batch)
        self.work = _workspace(self.nbytes)
        dev = device()
        self.best_val = torch.empty(1, dtype=torch.float64, device=dev)
        self.best_idx = torch.empty(1, dtype=torch.int64, device=dev)
        self.spec_c = factor.spec.to_c()
        self.reset()

    def reset(self):
        _lib.check(self.lib.tb200_best_reset(_lib.ptr(self.best_val), _lib.ptr(self.best_idx), _lib.stream_ptr()))

    def run(self, Xc_dev, m, index_base=0, cand_c=None, mu=None, sigma=None, acq=None):
        """candidates [index_base, index_base+m): Xc_dev (m,d) device tensor, or generated from cand_c"""
        f = self.f
        ex = C.byref(self.excl_c) if self.excl_c is not None else None
        if self.engine != "dmma":
            _lib.check(self.lib.tb200_acquire_i8_excl(
                ex, _lib.ptr(Xc_dev), int(m), int(index_base), C.byref(cand_c) if cand_c is not None else None,
                _lib.ptr(f.XsT), _lib.ptr(f.linvS), _lib.ptr(f.rowscale), self.gmax,
                _lib.ptr(f.alpha), f.n, C.byref(self.spec_c), _lib.ptr(f.theta), C.byref(self.acq_c), _lib.ptr(mu),
                _lib.ptr(sigma), _lib.ptr(acq), _lib.ptr(self.best_val), _lib.ptr(self.best_idx), self.batch,
                _lib.ptr(self.work), self.nbytes, _lib.stream_ptr()))
            return
        _lib.check(self.lib.tb200_acquire_excl(
            ex, _lib.ptr(Xc_dev), int(m), int(index_base), C.byref(cand_c) if cand_c is not None else None,
            _lib.ptr(f.XsT), _lib.ptr(f.Linv), _lib.ptr(f.alpha), f.n, C.byref(self.spec_c), _lib.ptr(f.theta),
            C.byref(self.acq_c), _lib.ptr(mu), _lib.ptr(sigma), _lib.ptr(acq), _lib.ptr(self.best_val),
            _lib.ptr(self.best_idx), self.batch, _lib.ptr(self.work), self.nbytes, _lib.stream_ptr()))

    def run_host(self, host, index_base=0, chunk=1 << 18, after_chunk=None):
        """candidates from a (pinned) host tensor (m,d), streamed in chunks: the H2D copy of chunk i+1 overlaps the
        kernels of chunk i.  TWO device buffers are allocated once and reused under events -- a fresh allocation per
        chunk went through the caching allocator with `record_stream`, whose deferred frees made the end-to-end time
        depend on the allocator's state (6.8e6 candidates/s in one run, 4.9e6 in the next).  after_chunk(): optional
        callback after each chunk's kernels are queued (the polish stage collects the local bests there)."""
        m, d = host.shape
        if m == 0:
            return
        dev = device()
        rows = min(int(chunk), m)
        nchunks = (m + rows - 1) // rows
        bufs = [torch.empty((rows, d), dtype=torch.float64, device=dev) for _ in range(min(2, nchunks))]
        used = [None] * len(bufs)  # event after the kernels that read the buffer
        copy_stream, main = torch.cuda.Stream(), torch.cuda.current_stream()
        copy_stream.wait_stream(main)  # the buffers were allocated in main's stream order
        for ci in range(nchunks):
            off = ci * rows
            cnt = min(rows, m - off)
            b = bufs[ci % len(bufs)]
            with torch.cuda.stream(copy_stream):
                if used[ci % len(bufs)] is not None:
                    copy_stream.wait_event(used[ci % len(bufs)])
                b[:cnt].copy_(host[off:off + cnt], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy_stream)
            main.wait_event(ev)
            self.run(b, cnt, index_base + off, None)
            if after_chunk is not None:
                after_chunk()
            done = torch.cuda.Event()
            done.record(main)
            used[ci % len(bufs)] = done
        # the last user of every buffer is a kernel on `main`: freeing them in main's stream order is safe

    def best(self):
        """(value, global index); index -1 when nothing was selectable (all NaN / empty)"""
        return float(self.best_val.item()), int(self.best_idx.item())

    def local_bests(self):
        """(values, global indices) of the per-CTA running bests of the LAST run() call (tb200_acquire_local_bests):
        distinct good candidates, index >= 0 only, best first (ties: lowest index)."""
        cnt = int(self.lib.tb200_acquire_local_best_count(self.batch, self.f.n))
        dev = device()
        v = torch.empty(cnt, dtype=torch.float64, device=dev)
        i = torch.empty(cnt, dtype=torch.int64, device=dev)
        _lib.check(self.lib.tb200_acquire_local_bests(_lib.ptr(self.work), self.f.n, self.f.spec.d, self.batch,
                                                      int(self.engine != "dmma"), _lib.ptr(v), _lib.ptr(i),
                                                      _lib.stream_ptr()))
        v, i = v.cpu().numpy(), i.cpu().numpy()
        keep = (i >= 0) & ~np.isnan(v)
        v, i = v[keep], i[keep]
        order = np.lexsort((i, -v))
        return v[order], i[order]


def min_sqdist(Xq, X, want_arg=False):
    """smallest squared Euclidean distance from each row of Xq (m,d) to the rows of X (n,d), on the device
    (tb200_min_sqdist; batched turbo/utils.py:111-130).  Host or device inputs; returns host arrays."""
    lib = _lib.load()
    Xq_dev = Xq if torch.is_tensor(Xq) else to_device(np.ascontiguousarray(Xq, dtype=np.float64))
    X_dev = X if torch.is_tensor(X) else to_device(np.ascontiguousarray(X, dtype=np.float64))
    assert Xq_dev.dim() == 2 and X_dev.dim() == 2, 'must be 2D arrays'
    assert Xq_dev.shape[1] == X_dev.shape[1], 'different number of attributes'
    m, d = Xq_dev.shape
    dev = device()
    out = torch.empty(m, dtype=torch.float64, device=dev)
    arg = torch.empty(m, dtype=torch.int64, device=dev) if want_arg else None
    _lib.check(lib.tb200_min_sqdist(_lib.ptr(Xq_dev), m, _lib.ptr(X_dev), int(X_dev.shape[0]), d, _lib.ptr(out),
                                    _lib.ptr(arg), _lib.stream_ptr()))
    return (out.cpu().numpy(), arg.cpu().numpy()) if want_arg else out.cpu().numpy()


def acquisition_value_and_grad(factor, acq_c, X):
    """acquisition value (P,) and gradient (P,d) at host points X through tb200_acquire_grad (128 points per call)"""
    lib = _lib.load()
    X = np.ascontiguousarray(np.atleast_2d(X), dtype=np.float64)
    P, d = X.shape
    assert d == factor.spec.d
    dev = device()
    Xd = torch.from_numpy(X).to(dev)
    val = torch.empty(P, dtype=torch.float64, device=dev)
    grad = torch.empty((P, d), dtype=torch.float64, device=dev)
    nbytes = lib.tb200_acquire_grad_workspace_bytes(factor.n, d)
    work = _workspace(nbytes)
    cs = factor.spec.to_c()
    for off in range(0, P, 128):
        cnt = min(128, P - off)
        _lib.check(lib.tb200_acquire_grad(_lib.ptr(Xd[off:]), cnt, _lib.ptr(factor.XsT), _lib.ptr(factor.Linv),
                                          _lib.ptr(factor.alpha), factor.n, C.byref(cs), _lib.ptr(factor.theta),
                                          C.byref(acq_c), _lib.ptr(val[off:]), _lib.ptr(grad[off:]), None, None,
                                          _lib.ptr(work), nbytes, _lib.stream_ptr()))
    out = val.cpu().numpy(), grad.cpu().numpy()
    del work
    return out


def acq_spec(acq, sign, incumbent, param, y_mean, y_std):
    return _lib.AcqSpecC(int(acq), float(sign), float(incumbent), float(param), float(y_mean), float(y_std))


def generate_candidates(cand_c, d, index_base, m):
    """device tensor (m,d) of generated candidates (tests / winner-row recovery)"""
    out = torch.empty((int(m), int(d)), dtype=torch.float64, device=device())
    _lib.check(_lib.load().tb200_generate_candidates(C.byref(cand_c), int(d), int(index_base), int(m),
                                                     _lib.ptr(out), _lib.stream_ptr()))
    return out


def fp64_peak(ms=50.0):
    """measured (DMMA TFLOP/s, DFMA TFLOP/s) of the current device"""
    _require_cuda()
    a, b = C.c_double(), C.c_double()
    _lib.check(_lib.load().tb200_fp64_peak(float(ms), C.byref(a), C.byref(b)))
    return a.value, b.value


def i8_peak(ms=30.0, random_digits=True):
    """measured INT8 tensor-core peak (TOP/s) of the current device (tb200_i8_peak): with pseudo-random digits (the
    power the real kernels draw) or with zero operands (the ceiling at the maximum clock)"""
    _require_cuda()
    a = C.c_double()
    _lib.check(_lib.load().tb200_i8_peak(float(ms), int(bool(random_digits)), C.byref(a)))
    return a.value


def shutdown():
    """release the library's per-device streams and events (tb200_shutdown); they are re-created on demand"""
    _lib.check(_lib.load().tb200_shutdown())


def dgemm(ta, tb, alpha, A, B, beta, Cm):
    """C = alpha*op(A)op(B) + beta*C on the library's DMMA GEMM (row-major torch tensors)"""
    M, N = Cm.shape
    K = A.shape[0] if ta else A.shape[1]
    _lib.check(_lib.load().tb200_dgemm(int(ta), int(tb), M, N, K, float(alpha), _lib.ptr(A), A.stride(0),
                                       _lib.ptr(B), B.stride(0), float(beta), _lib.ptr(Cm), Cm.stride(0),
                                       _lib.stream_ptr()))
    return Cm


def launch_count():
    return int(_lib.load().tb200_launch_count())
