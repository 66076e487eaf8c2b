"""Lock-step batched L-BFGS-B over R restarts.

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


def minimize_batched(fun, x0, bounds, max_iter=200, history=10, pgtol=1e-5, ftol=2.220446049250313e-09,
                     max_ls=20):
    """minimise R independent functions f_r(x_r) under the same box.

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

    cond = threading.Condition()
    requests, answers, finished = {}, {}, set()
    errors = []

    def worker(r):
        first = [True]

        def objective(x):
            if first[0]:  # SciPy starts by evaluating x0, which the first batched call already did
                first[0] = False
                if np.array_equal(x, x0[r]):
                    return f0[r], g0[r].copy()
            with cond:
                requests[r] = np.array(x, dtype=np.float64)
                cond.notify_all()
                while r not in answers:
                    cond.wait()
                f, g = answers.pop(r)
            return f, g

        try:
            res = scipy.optimize.minimize(objective, x0[r], jac=True, method="L-BFGS-B",
                                          bounds=list(zip(lo, hi)),
                                          options=dict(maxiter=int(max_iter), maxcor=int(history), ftol=ftol,
                                                       gtol=pgtol, maxls=int(max_ls)))
            if np.isfinite(res.fun) and res.fun <= f0[r]:
                x_out[r], f_out[r] = np.clip(res.x, lo, hi), float(res.fun)
            n_iter[r] = int(res.nit)
        except Exception as e:  # never leave the coordinator waiting
            errors.append(e)
        finally:
            with cond:
                finished.add(r)
                cond.notify_all()

    threads = [threading.Thread(target=worker, args=(r,), daemon=True) for r in live]
    for t in threads:
        t.start()
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
        with cond:
            for k, r in enumerate(members):
                if np.isfinite(f[k]) and np.all(np.isfinite(g[k])):
                    answers[int(r)] = (float(f[k]), g[k].copy())
                else:
                    answers[int(r)] = (1e8 * max(1.0, abs(f0[r])), np.zeros(p))
            cond.notify_all()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    return x_out, f_out, n_iter, n_eval
