"""Lock-step batched L-BFGS-B over R restarts (SciPy L-BFGS-B instances answered by one batched device call).

scikit-learn runs SciPy's L-BFGS-B once per restart, sequentially
(sklearn:_gpr.py:312-333, :658-674 behind turbo/modules/surrogates.py:313-318), and the reference's
auxiliary optimiser does the same per starting point (turbo/modules/auxiliary_optimisers.py:84-93).
Here every restart IS a SciPy L-BFGS-B instance (same algorithm, line search and stopping rules as
the reference: projected-gradient tolerance 1e-5, relative decrease factr*eps = 2.2e-9), but the R
instances advance together: each runs in its own thread and blocks inside its objective call until
the coordinator has collected the requests of all live instances and answered them with ONE batched
device call (tb200_lml_batched / tb200_acquire_grad over the members that asked).  The number of
device calls is therefore max_r(nfev_r), not sum_r(nfev_r), and the iterates of every restart are
exactly SciPy's.

Pure host logic on (R, p) NumPy arrays: p <= 66 numbers per restart, no device math here.  The
device is only ever touched from the calling thread.
"""
import threading

import numpy as np
import scipy.optimize


def _single_threaded_blas():
    """L-BFGS-B's own linear algebra is on (history x history) matrices; a multi-threaded BLAS spends ~0.5 ms per call
    synchronising its pool on them (measured: 1.23 s -> 0.07 s for 64 restarts x 30 iterations)"""
    try:
        from threadpoolctl import threadpool_limits
        return threadpool_limits(limits=1)
    except ImportError:
        import contextlib
        return contextlib.nullcontext()


class _Aborted(Exception):
    """raised inside a worker's objective when the coordinator of the threads driver has failed"""


class _Instance:
    """one SciPy L-BFGS-B state driven through its reverse-communication routine `setulb`, exactly as
    scipy.optimize._lbfgsb_py._minimize_lbfgsb does (same workspace sizes, same task codes)"""

    def __init__(self, setulb, x0, lo, hi, m, factr, pgtol, maxls, max_iter, int_dtype):
        n = x0.size
        self.setulb, self.m, self.factr, self.pgtol, self.maxls, self.max_iter = setulb, m, factr, pgtol, maxls, max_iter
        self.x = np.array(x0, dtype=np.float64)
        self.f = np.array(0.0, dtype=np.float64)
        self.g = np.zeros(n, dtype=np.float64)
        self.lo, self.hi = np.array(lo, dtype=np.float64), np.array(hi, dtype=np.float64)
        self.nbd = np.full(n, 2, dtype=int_dtype)  # both bounds finite
        self.wa = np.zeros(2 * m * n + 5 * n + 11 * m * m + 8 * m, np.float64)
        self.iwa = np.zeros(3 * n, dtype=int_dtype)
        self.task = np.zeros(2, dtype=int_dtype)
        self.ln_task = np.zeros(2, dtype=int_dtype)
        self.lsave = np.zeros(4, dtype=int_dtype)
        self.isave = np.zeros(44, dtype=int_dtype)
        self.dsave = np.zeros(29, dtype=np.float64)
        self.n_iter = 0
        self.done = False

    def advance(self):
        """run until the routine wants f and g at self.x (returns True) or stops (returns False)"""
        while True:
            self.setulb(self.m, self.x, self.lo, self.hi, self.nbd, self.f, self.g, self.factr, self.pgtol, self.wa,
                        self.iwa, self.task, self.lsave, self.isave, self.dsave, self.maxls, self.ln_task)
            if self.task[0] == 3:
                return True
            if self.task[0] == 1:  # new iteration
                self.n_iter += 1
                if self.n_iter >= self.max_iter:
                    self.task[0], self.task[1] = 5, 504
            else:
                self.done = True
                return False


_RC_AVAILABLE = None


def _setulb_and_int_dtype():
    from scipy.optimize import _lbfgsb
    try:
        from scipy.optimize._lbfgsb_py import HAS_ILP64
    except ImportError:
        HAS_ILP64 = False
    return _lbfgsb.setulb, (np.int64 if HAS_ILP64 else np.int32)


def _reverse_communication_available():
    """True when SciPy's private `setulb` exists and accepts the argument list `_Instance` passes (probed once on a
    one-dimensional quadratic: the routine must come back asking for f and g)."""
    global _RC_AVAILABLE
    if _RC_AVAILABLE is None:
        try:
            setulb, int_dtype = _setulb_and_int_dtype()
            probe = _Instance(setulb, np.array([0.5]), np.array([0.0]), np.array([1.0]), 3, 1e7, 1e-5, 20, 5, int_dtype)
            _RC_AVAILABLE = bool(probe.advance())
        except (ImportError, AttributeError, TypeError, ValueError):
            _RC_AVAILABLE = False
    return _RC_AVAILABLE


def _minimize_reverse_communication(fun, x0, lo, hi, f0, g0, live, max_iter, history, pgtol, ftol, max_ls):
    setulb, int_dtype = _setulb_and_int_dtype()
    R, p = x0.shape
    factr = ftol / np.finfo(float).eps
    inst = {r: _Instance(setulb, x0[r], lo, hi, int(history), factr, pgtol, int(max_ls), int(max_iter), int_dtype)
            for r in live}
    x_out, f_out, n_iter = x0.copy(), f0.copy(), np.zeros(R, dtype=int)
    n_eval = 1
    first = True
    while True:
        asking = [r for r in live if not inst[r].done and inst[r].advance()]
        if not asking:
            break
        X = np.stack([inst[r].x for r in asking])
        if first and all(np.array_equal(inst[r].x, x0[r]) for r in asking):
            f, g = f0[asking], g0[asking]  # the routine starts at x0, which the caller already evaluated
        else:
            f, g = fun(X, np.array(asking, dtype=int))
            n_eval += 1
            f, g = np.asarray(f, dtype=np.float64), np.asarray(g, dtype=np.float64)
        first = False
        for k, r in enumerate(asking):
            if np.isfinite(f[k]) and np.all(np.isfinite(g[k])):
                inst[r].f, inst[r].g = np.array(float(f[k])), np.array(g[k], dtype=np.float64)
            else:
                inst[r].f, inst[r].g = np.array(1e8 * max(1.0, abs(f0[r]))), np.zeros(p)
    for r in live:
        fr = float(inst[r].f)
        if np.isfinite(fr) and fr <= f0[r]:
            x_out[r], f_out[r] = np.clip(inst[r].x, lo, hi), fr
        n_iter[r] = inst[r].n_iter
    return x_out, f_out, n_iter, n_eval


def minimize_batched(fun, x0, bounds, max_iter=200, history=10, pgtol=1e-5, ftol=2.220446049250313e-09,
                     max_ls=20, driver="auto"):
    """minimise R independent functions f_r(x_r) under the same box.

    driver: 'rc' drives SciPy's compiled L-BFGS-B routine (`scipy.optimize._lbfgsb.setulb`, the reverse-communication
    core `scipy.optimize.minimize` itself loops over) for all R instances from ONE thread; 'threads' wraps R public
    `scipy.optimize.minimize` calls in threads (same iterates, but the GIL hand-offs between 64 threads cost ~20 ms per
    round); 'auto' = 'rc' when the private routine has the expected signature, else 'threads'.

    fun(X (k,p), members (k,) int) -> (f (k,), g (k,p)) evaluates the listed members only;
    non-finite f means "infeasible here" (a failed Cholesky).  SciPy's line search does not survive a
    +inf trial value (it returns the start and reports convergence, which is what happens to a
    scikit-learn restart that steps into a non-PD region); such trials are answered with a large
    finite penalty and a zero gradient instead, so the line search backs off and the restart goes on.
    Returns (x (R,p), f (R,), n_iter (R,), n_eval) where n_eval is the
    number of batched calls of ``fun``.
    """
    bounds = np.asarray(bounds, dtype=np.float64)
    lo, hi = bounds[:, 0], bounds[:, 1]
    x0 = np.clip(np.array(x0, dtype=np.float64, ndmin=2), lo, hi)
    R, p = x0.shape
    f0, g0 = fun(x0, np.arange(R))
    f0, g0 = np.array(f0, dtype=np.float64), np.array(g0, dtype=np.float64)
    n_eval = 1
    x_out, f_out, n_iter = x0.copy(), f0.copy(), np.zeros(R, dtype=int)
    live = [r for r in range(R) if np.isfinite(f0[r]) and np.all(np.isfinite(g0[r]))]
    if not live:
        return x_out, f_out, n_iter, n_eval  # every start infeasible: reported, not crashed
    import os
    driver = os.environ.get("TB200_LBFGS_DRIVER", driver)  # diagnostics: threads | rc | rc-nolimit
    if driver in ("auto", "rc", "rc-nolimit"):
        # the guard covers the PROBE of SciPy's private routine only: an error raised by `fun` (the device call) or by
        # the optimisation itself propagates instead of silently re-running the fit on the threads driver
        if _reverse_communication_available():
            import contextlib
            with (contextlib.nullcontext() if driver == "rc-nolimit" else _single_threaded_blas()):
                return _minimize_reverse_communication(fun, x0, lo, hi, f0, g0, live, max_iter, history, pgtol, ftol,
                                                       max_ls)
        if driver == "rc":
            raise RuntimeError("scipy.optimize._lbfgsb.setulb is missing or has an unexpected signature")

    cond = threading.Condition()  # only the coordinator waits on it; each worker sleeps on its own event
    requests, answers, finished = {}, {}, set()
    wake = {r: threading.Event() for r in live}
    errors = []
    aborted = []

    def worker(r):
        first = [True]

        def objective(x):
            if first[0]:  # SciPy starts by evaluating x0, which the first batched call already did
                first[0] = False
                if np.array_equal(x, x0[r]):
                    return f0[r], g0[r].copy()
            with cond:
                requests[r] = np.array(x, dtype=np.float64)
                cond.notify()
            wake[r].wait()
            wake[r].clear()
            if aborted:  # the coordinator failed (an error from `fun`): leave the optimisation, free the thread
                raise _Aborted()
            return answers.pop(r)

        try:
            res = scipy.optimize.minimize(objective, x0[r], jac=True, method="L-BFGS-B",
                                          bounds=list(zip(lo, hi)),
                                          options=dict(maxiter=int(max_iter), maxcor=int(history), ftol=ftol,
                                                       gtol=pgtol, maxls=int(max_ls)))
            if np.isfinite(res.fun) and res.fun <= f0[r]:
                x_out[r], f_out[r] = np.clip(res.x, lo, hi), float(res.fun)
            n_iter[r] = int(res.nit)
        except _Aborted:
            pass
        except Exception as e:  # never leave the coordinator waiting
            errors.append(e)
        finally:
            with cond:
                finished.add(r)
                cond.notify()

    threads = [threading.Thread(target=worker, args=(r,), daemon=True) for r in live]
    for t in threads:
        t.start()
    try:
        while True:
            with cond:
                while len(requests) + len(finished) < len(live):
                    cond.wait()
                if not requests:
                    break
                members = np.array(sorted(requests), dtype=int)
                X = np.stack([requests[r] for r in members])
                requests.clear()
            f, g = fun(X, members)  # the one device call of this round, from the calling thread
            n_eval += 1
            f, g = np.asarray(f, dtype=np.float64), np.asarray(g, dtype=np.float64)
            for k, r in enumerate(members):
                if np.isfinite(f[k]) and np.all(np.isfinite(g[k])):
                    answers[int(r)] = (float(f[k]), g[k].copy())
                else:
                    answers[int(r)] = (1e8 * max(1.0, abs(f0[r])), np.zeros(p))
                wake[int(r)].set()
    except BaseException:
        # an error from `fun`: every worker blocked in its objective is released with the abort flag set, so no thread
        # is left waiting on its event forever
        aborted.append(True)
        for ev in wake.values():
            ev.set()
        for t in threads:
            t.join()
        raise
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    return x_out, f_out, n_iter, n_eval
