"""CPU oracle for the GP-surrogate + acquisition hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, in NumPy/SciPy, the algorithm the reference executes on its
scikit-learn path.  It is the *checker*: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.  Nothing under ``turbo_b200/`` imports it and the
product path has no CPU fallback.

The arithmetic of the reference lives in third-party scikit-learn (unpinned in the
reference: ``requirements.txt:7`` says ``sklearn``; 1.9.0 on this image), SciPy and
NumPy, reached from these reference call sites:

* ``turbo/modules/surrogates.py:294-326`` (``construct_model`` -> ``GaussianProcessRegressor.fit``)
* ``turbo/modules/surrogates.py:332-338`` (``ModelInstance.predict``)
* ``turbo/modules/acquisition_functions.py:147-158, 225-247, 336-358`` (UCB / PI / EI)
* ``turbo/modules/auxiliary_optimisers.py:60-66, 117-129`` (batched evaluate + argsort)
* ``turbo/modules/naive_selectors.py:41-46, 65-83`` (uniform / LHS candidates)

Parity pinning: the reference's own tests hold NO golden vector for this path
(``tests/test_utils.py`` pins only ``remap``), so the oracle is pinned against
outputs of the reference itself: ``oracle/make_golden.py`` imports the reference
from ``/root/reference`` plus scikit-learn, runs ``SciKitGPSurrogate.ModelInstance``
and ``EI/PI/UCB.FunctionInstance`` on seeded inputs and commits the vectors under
``tests/golden/``; ``tests/test_oracle.py`` checks this file against them (and
against scikit-learn live, which exists on both machines).

``sklearn:`` citations mean ``sklearn/gaussian_process/`` of scikit-learn 1.9.0.
"""
import math

import numpy as np
from scipy.linalg import cho_solve, cholesky, solve_triangular
from scipy.spatial.distance import cdist, pdist, squareform
from scipy.special import ndtr

# kernel families (same integer codes as include/turbo_b200.h)
KIND_RBF, KIND_MATERN12, KIND_MATERN32, KIND_MATERN52 = 0, 1, 2, 3
ACQ_EI, ACQ_PI, ACQ_UCB = 0, 1, 2
_NORM_PDF_C = math.sqrt(2 * math.pi)


class KernelSpec:
    """``[const *] (RBF | Matern-nu) [+ White]`` with the hyper-parameters in
    scikit-learn's theta order ``[log const, log length_scale..., log noise]``
    (sklearn:kernels.py:838-890 Sum, :936-990 Product; names probed in SURVEY 8a5)."""

    def __init__(self, kind, d, ard, has_const=True, has_white=True):
        self.kind, self.d, self.ard = int(kind), int(d), bool(ard)
        self.has_const, self.has_white = bool(has_const), bool(has_white)
        self.n_ls = self.d if self.ard else 1
        self.p = int(self.has_const) + self.n_ls + int(self.has_white)

    def unpack(self, theta):
        theta = np.asarray(theta, dtype=np.float64)
        assert theta.shape == (self.p,)
        e = np.exp(theta)
        i = 0
        sf2 = 1.0
        if self.has_const:
            sf2 = e[0]
            i = 1
        ls = e[i:i + self.n_ls]
        noise = e[i + self.n_ls] if self.has_white else 0.0
        return sf2, ls, noise


def _shape(kind, dists, sq):
    """stationary shape g(r); ``dists`` is r (Matern) or r^2 (RBF, ``sq``).
    sklearn:kernels.py:1561-1570 (RBF), :1722-1729 (Matern 1/2, 3/2, 5/2)."""
    if kind == KIND_RBF:
        assert sq
        return np.exp(-0.5 * dists)
    assert not sq
    if kind == KIND_MATERN12:
        return np.exp(-dists)
    if kind == KIND_MATERN32:
        K = dists * math.sqrt(3)
        return (1.0 + K) * np.exp(-K)
    K = dists * math.sqrt(5)
    return (1.0 + K + K ** 2 / 3.0) * np.exp(-K)


def gram_train(spec, theta, X, alpha_jitter=1e-10, eval_gradient=False):
    """K(X,X) + jitter*I and optionally dK/dtheta (n,n,p).
    sklearn:kernels.py:1558-1587 / :1713-1786 (stationary part, diagonal forced to 1),
    :1283-1291 (const gradient = K_const), :967-969 (product rule),
    :1408-1413 (white gradient = noise*I); sklearn:_gpr.py:589 (``K[diag] += alpha``)."""
    sf2, ls, noise = spec.unpack(theta)
    Xs = X / ls
    metric = "sqeuclidean" if spec.kind == KIND_RBF else "euclidean"
    dists = pdist(Xs, metric=metric)
    G = squareform(_shape(spec.kind, dists, spec.kind == KIND_RBF))
    np.fill_diagonal(G, 1)
    K = sf2 * G
    n = X.shape[0]
    if spec.has_white:
        K = K + noise * np.eye(n)
    Kj = K.copy()
    Kj[np.diag_indices_from(Kj)] += alpha_jitter
    if not eval_gradient:
        return Kj
    grads = []
    if spec.has_const:
        grads.append((sf2 * G)[:, :, None])
    if spec.ard:
        D = (X[:, None, :] - X[None, :, :]) ** 2 / (ls ** 2)
    else:
        d2 = dists if spec.kind == KIND_RBF else dists ** 2
        D = squareform(d2)[:, :, None]
    if spec.kind == KIND_RBF:
        Gg = D * G[..., None]
    elif spec.kind == KIND_MATERN12:
        den = np.sqrt(D.sum(axis=2))[:, :, None]
        div = np.zeros_like(D)
        np.divide(D, den, out=div, where=den != 0)
        Gg = G[..., None] * div
    elif spec.kind == KIND_MATERN32:
        Gg = 3 * D * np.exp(-np.sqrt(3 * D.sum(-1)))[..., None]
    else:
        tmp = np.sqrt(5 * D.sum(-1))[..., None]
        Gg = 5.0 / 3.0 * D * (tmp + 1) * np.exp(-tmp)
    grads.append(sf2 * Gg)
    if spec.has_white:
        grads.append((noise * np.eye(n))[:, :, None])
    return Kj, np.concatenate(grads, axis=2)


def cross_cov(spec, theta, Xc, X):
    """K(Xc, X): white noise contributes 0 off the training diagonal
    (sklearn:kernels.py:1418-1419); cdist on scaled points (:1566-1570, :1719-1720)."""
    sf2, ls, _ = spec.unpack(theta)
    metric = "sqeuclidean" if spec.kind == KIND_RBF else "euclidean"
    dists = cdist(Xc / ls, X / ls, metric=metric)
    return sf2 * _shape(spec.kind, dists, spec.kind == KIND_RBF)


def prior_diag(spec, theta):
    """kernel_.diag(X): const*1 + noise (sklearn:kernels.py:1315, :1438)."""
    sf2, _, noise = spec.unpack(theta)
    return sf2 + noise


class FittedGP:
    """GaussianProcessRegressor at FIXED theta (sklearn:_gpr.py:275-280, 349-367)."""

    def __init__(self, spec, theta, X, y, normalize_y=True, alpha_jitter=1e-10):
        self.spec, self.theta = spec, np.asarray(theta, dtype=np.float64).copy()
        self.X = np.ascontiguousarray(X, dtype=np.float64)
        y = np.asarray(y, dtype=np.float64).reshape(-1)
        if normalize_y:
            self.y_mean = float(np.mean(y))
            s = float(np.std(y))
            self.y_std = 1.0 if s < 10 * np.finfo(np.float64).eps else s  # _handle_zeros_in_scale
        else:
            self.y_mean, self.y_std = 0.0, 1.0
        self.y_n = (y - self.y_mean) / self.y_std
        self.alpha_jitter = alpha_jitter
        K = gram_train(spec, self.theta, self.X, alpha_jitter)
        self.L = cholesky(K, lower=True, check_finite=False)  # raises LinAlgError like sklearn
        self.alpha = cho_solve((self.L, True), self.y_n, check_finite=False)

    def predict(self, Xc, return_std=True):
        """sklearn:_gpr.py:446-500 behind turbo/modules/surrogates.py:332-338."""
        Ks = cross_cov(self.spec, self.theta, Xc, self.X)
        mu = self.y_std * (Ks @ self.alpha) + self.y_mean
        if not return_std:
            return mu
        V = solve_triangular(self.L, Ks.T, lower=True, check_finite=False)
        var = np.full(Xc.shape[0], prior_diag(self.spec, self.theta))
        var -= np.einsum("ij,ji->i", V.T, V)
        var[var < 0] = 0.0
        var = var * self.y_std ** 2
        return mu, np.sqrt(var)

    def lml(self, theta=None, eval_gradient=False):
        return log_marginal_likelihood(self.spec, self.theta if theta is None else theta,
                                       self.X, self.y_n, self.alpha_jitter, eval_gradient)


def log_marginal_likelihood(spec, theta, X, y_n, alpha_jitter=1e-10, eval_gradient=False):
    """sklearn:_gpr.py:584-651.  Cholesky failure -> (-inf, zeros) (:590-593)."""
    theta = np.asarray(theta, dtype=np.float64)
    if eval_gradient:
        K, dK = gram_train(spec, theta, X, alpha_jitter, True)
    else:
        K = gram_train(spec, theta, X, alpha_jitter)
    try:
        L = cholesky(K, lower=True, check_finite=False)
    except np.linalg.LinAlgError:
        return (-np.inf, np.zeros_like(theta)) if eval_gradient else -np.inf
    a = cho_solve((L, True), y_n, check_finite=False)
    lml = -0.5 * float(y_n @ a) - float(np.log(np.diag(L)).sum()) - K.shape[0] / 2 * math.log(2 * math.pi)
    if not eval_gradient:
        return lml
    inner = np.outer(a, a) - cho_solve((L, True), np.eye(K.shape[0]), check_finite=False)
    grad = 0.5 * np.einsum("ij,jik->k", inner, dK)
    return lml, grad


def acquisition(acq, mus, sigmas, sign, incumbent=0.0, param=0.0):
    """EI / PI / UCB(-LCB) exactly as turbo/modules/acquisition_functions.py
    :336-358 (EI), :225-247 (PI), :147-158 (UCB; beta=inf -> sigma).
    ``norm.cdf`` is ``ndtr``; ``norm.pdf(x) = exp(-x**2/2)/sqrt(2*pi)``."""
    mus, sigmas = np.asarray(mus), np.asarray(sigmas)
    if acq == ACQ_UCB:
        if math.isinf(param):
            return sigmas
        return sign * mus + param * sigmas
    mask = sigmas != 0
    s = sigmas[mask]
    diff = sign * (mus[mask] - incumbent) - param
    Z = diff / s
    out = np.zeros_like(mus)
    if acq == ACQ_PI:
        out[mask] = ndtr(Z)
    else:
        out[mask] = diff * ndtr(Z) + s * (np.exp(-Z ** 2 / 2.0) / _NORM_PDF_C)
    return out


def argmax_first(values):
    """winner of ``np.argsort(-acq)[0]`` (auxiliary_optimisers.py:61-66) with the
    deterministic tie-break the device uses: lowest index; NaN never wins."""
    v = np.where(np.isnan(values), -np.inf, values)
    return int(np.argmax(v))


def min_sqdist(Xq, X):
    """smallest squared Euclidean distance from each row of Xq to the rows of X, the quantity
    turbo/utils.py:130 compares with the tolerance: ``np.sum((xs - x)**2, axis=1)`` (same expression, row by row,
    so NumPy's own summation order is the reference).  Empty X: +inf (turbo/utils.py:119-121: not close)."""
    Xq, X = np.atleast_2d(Xq), np.atleast_2d(X)
    out = np.full(Xq.shape[0], np.inf)
    arg = np.full(Xq.shape[0], -1, dtype=np.int64)
    if X.shape[0] == 0:
        return out, arg
    for i in range(Xq.shape[0]):
        s = np.sum((X - Xq[i:i + 1]) ** 2, axis=1)
        if not np.all(np.isnan(s)):
            arg[i] = int(np.nanargmin(s))
            out[i] = s[arg[i]]
    return out, arg


def close_to_any(x, xs, tol=1e-5):
    """turbo/utils.py:111-130"""
    assert x.shape[1] == xs.shape[1] and x.shape[0] == 1
    return bool(np.any(np.sum((xs - x) ** 2, axis=1) <= tol))


def uniform_candidates(rng, m, lo, hi):
    """naive_selectors.py:41-46: one uniform column per bound."""
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    return np.hstack([rng.uniform(lo[k], hi[k], size=(m, 1)) for k in range(len(lo))])


def lhs_candidates(rng, m, lo, hi):
    """naive_selectors.py:65-83: stratified per dim, then per-dim permutation."""
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    d = len(lo)
    seq = lo + (hi - lo) * (np.arange(m).reshape(-1, 1) + rng.random((m, d))) / m
    for k in range(d):
        seq[:, k] = rng.permutation(seq[:, k])
    return seq


def hartmann6(X):
    """Hartmann-6 (BASELINE.json config 2), standard constants."""
    a = np.array([1.0, 1.2, 3.0, 3.2])
    A = np.array([[10, 3, 17, 3.5, 1.7, 8], [0.05, 10, 17, 0.1, 8, 14],
                  [3, 3.5, 1.7, 10, 17, 8], [17, 8, 0.05, 10, 0.1, 14]], dtype=float)
    P = 1e-4 * np.array([[1312, 1696, 5569, 124, 8283, 5886], [2329, 4135, 8307, 3736, 1004, 9991],
                         [2348, 1451, 3522, 2883, 3047, 6650], [4047, 8828, 8732, 5743, 1091, 381]], dtype=float)
    e = np.einsum("ij,nij->ni", A, (X[:, None, :] - P[None, :, :]) ** 2)
    return -(a * np.exp(-e)).sum(axis=1)


def branin(x, y):
    """demos/Branin-Hoo.ipynb cell 5 objective (BASELINE.json config 1)."""
    return (y - (5.1 / (4 * np.pi ** 2)) * x ** 2 + 5 * x / np.pi - 6) ** 2 + 10 * (1 - 1 / (8 * np.pi)) * np.cos(x) + 10
