"""Lock-step batched L-BFGS-B (projected L-BFGS with box bounds) over R restarts.

scikit-learn runs SciPy's L-BFGS-B once per restart, sequentially
(sklearn:_gpr.py:312-333, :658-674 behind turbo/modules/surrogates.py:313-318).  Here all R
restarts advance together so that every objective evaluation is ONE batched device call
(tb200_lml_batched over the still-active members).  The iterates are not SciPy's iterates --
parity with the reference is defined at fixed theta (BASELINE.json) -- but the stopping rules are
SciPy's defaults: projected-gradient tolerance 1e-5 and relative decrease factr*eps = 2.2e-9.

Pure host logic on (R, p) NumPy arrays: p <= 66 numbers per restart, no device math here.
"""
import numpy as np


def _two_loop(g, S, Y):
    """H*g by the L-BFGS two-loop recursion for one member (lists of s, y pairs, newest last)."""
    q = g.copy()
    al = []
    for s, y in zip(reversed(S), reversed(Y)):
        a = (s @ q) / (y @ s)
        al.append(a)
        q -= a * y
    if S:
        q *= (S[-1] @ Y[-1]) / (Y[-1] @ Y[-1])
    for (s, y), a in zip(zip(S, Y), reversed(al)):
        b = (y @ q) / (y @ s)
        q += (a - b) * s
    return q


def minimize_batched(fun, x0, bounds, max_iter=200, history=10, pgtol=1e-5, ftol=2.220446049250313e-09,
                     max_ls=20, c1=1e-4):
    """minimise R independent functions f_r(x_r) under the same box.

    fun(X (k,p), members (k,) int) -> (f (k,), g (k,p)) evaluates the listed members only;
    non-finite f means "infeasible here" (the line search backs off).
    Returns (x (R,p), f (R,), n_iter (R,), n_eval).
    """
    x = np.clip(np.array(x0, dtype=np.float64), bounds[:, 0], bounds[:, 1])
    R, p = x.shape
    lo, hi = bounds[:, 0], bounds[:, 1]
    f, g = fun(x, np.arange(R))
    f, g = np.array(f, dtype=np.float64), np.array(g, dtype=np.float64)
    n_eval = 1
    S = [[] for _ in range(R)]
    Y = [[] for _ in range(R)]
    active = np.isfinite(f)
    n_iter = np.zeros(R, dtype=int)

    def proj_grad(xr, gr):
        return xr - np.clip(xr - gr, lo, hi)

    for _ in range(max_iter):
        for r in np.flatnonzero(active):
            if np.max(np.abs(proj_grad(x[r], g[r]))) <= pgtol:
                active[r] = False
        idx = np.flatnonzero(active)
        if idx.size == 0:
            break
        # search directions
        D = np.zeros((R, p))
        for r in idx:
            at_bound = ((x[r] <= lo) & (g[r] > 0)) | ((x[r] >= hi) & (g[r] < 0))
            gf = np.where(at_bound, 0.0, g[r])
            d = -_two_loop(gf, S[r], Y[r])
            d[at_bound] = 0.0
            if not np.all(np.isfinite(d)) or d @ g[r] >= 0:
                S[r], Y[r] = [], []
                d = -gf
            if not S[r]:
                nrm = np.linalg.norm(d)
                if nrm > 1.0:
                    d = d / nrm  # first step (or after a reset): at most unit length, as L-BFGS-B does
            D[r] = d
        # projected backtracking line search, all members in lock step
        t = np.ones(R)
        pending = idx.copy()
        x_new, f_new, g_new = x.copy(), f.copy(), g.copy()
        accepted = np.zeros(R, dtype=bool)
        for _ls in range(max_ls):
            if pending.size == 0:
                break
            trial = np.clip(x[pending] + t[pending, None] * D[pending], lo, hi)
            ft, gt = fun(trial, pending)
            n_eval += 1
            still = []
            for k, r in enumerate(pending):
                decrease = g[r] @ (trial[k] - x[r])
                if np.isfinite(ft[k]) and np.all(np.isfinite(gt[k])) and ft[k] <= f[r] + c1 * decrease:
                    x_new[r], f_new[r], g_new[r] = trial[k], ft[k], gt[k]
                    accepted[r] = True
                else:
                    t[r] *= 0.5 if np.isfinite(ft[k]) else 0.1
                    still.append(r)
            pending = np.array(still, dtype=int)
        for r in idx:
            n_iter[r] += 1
            if not accepted[r]:
                if S[r]:  # retry once from steepest descent before giving up
                    S[r], Y[r] = [], []
                else:
                    active[r] = False
                continue
            s, y = x_new[r] - x[r], g_new[r] - g[r]
            if s @ y > 1e-10 * (y @ y):
                S[r].append(s)
                Y[r].append(y)
                if len(S[r]) > history:
                    S[r].pop(0)
                    Y[r].pop(0)
            rel = (f[r] - f_new[r]) / max(abs(f[r]), abs(f_new[r]), 1.0)
            x[r], f[r], g[r] = x_new[r], f_new[r], g_new[r]
            if rel <= ftol:
                active[r] = False
    return x, f, n_iter, n_eval
