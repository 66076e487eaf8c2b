"""GpuGPSurrogate: drop-in for turbo.modules.SciKitGPSurrogate (turbo/modules/surrogates.py:225-365).

Same constructor arguments, same ``construct_model(trial_num, X, y) -> (model, fitting_info)``
contract, same warm-start / ``training_iterations`` semantics; the arithmetic runs in
libturbo_b200.so on the current CUDA device instead of scikit-learn:

* fit  : R = ``training_iterations`` restarts advance in lock step (lbfgs.py), one batched
         tb200_lml_batched call per objective evaluation; restart 0 starts from the kernel's theta
         (or last trial's, ``param_continuity``), the others log-uniformly inside the bounds exactly
         as sklearn:_gpr.py:321-333 does; the best restart is refactored with tb200_factor.
* ``training_iterations == 0`` keeps theta fixed.  (The reference intends this -- docstring at
  surrogates.py:254 -- but its branch at :306-310 raises UnboundLocalError.)
* multi-GPU: with ``group`` (a torch.distributed process group) the restarts are split across the
  ranks and the best (-LML, theta) is all-gathered (dist.py); every rank then holds the same model.

``ModelInstance`` subclasses the reference's ``SciKitGPSurrogate.ModelInstance`` when turbo is
importable so turbo.plotting's isinstance checks (plotting/surrogates.py:67-75) accept it, and it
pickles through dill (recorder.py:141-146) by dropping all device state.
"""
import copy
import warnings

import numpy as np

from . import _lib, device, kernels, lbfgs
from ._reference_api import SciKitModelInstance, Surrogate


def _normalise_y(y, normalize_y):
    """sklearn:_gpr.py:275-285 (+ _handle_zeros_in_scale)"""
    y = np.asarray(y, dtype=np.float64).reshape(-1)
    if normalize_y:
        mean = float(np.mean(y))
        std = float(np.std(y))
        if std < 10 * np.finfo(np.float64).eps:
            std = 1.0
        return (y - mean) / std, mean, std
    return y.copy(), 0.0, 1.0


def _resolve_kernel(kernel, d):
    if isinstance(kernel, kernels.KernelSpec):
        if kernel.d != d:
            raise ValueError("KernelSpec.d=%d but X has %d columns" % (kernel.d, d))
        return kernel, np.zeros(kernel.p), kernel.default_bounds()
    if isinstance(kernel, tuple):  # (KernelSpec, theta0[, bounds])
        spec = kernel[0]
        theta0 = np.asarray(kernel[1], dtype=np.float64)
        bounds = np.asarray(kernel[2], dtype=np.float64) if len(kernel) > 2 else spec.default_bounds()
        return spec, theta0, bounds
    return kernels.from_sklearn(kernel, d)


class _DeviceModelInstance:
    """surrogates.py:328-365 contract; holds only NumPy state + lazily built device buffers.  The reference base
    class (scikit-learn or GPy flavour) is mixed in by the surrogate that builds it."""

    def __init__(self, spec, theta, X, y_n, y_mean, y_std, jitter, device_inputs=None, variance_engine=None,
                 factor=None):
        self.variance_engine = variance_engine
        self.spec, self.theta = spec, np.asarray(theta, dtype=np.float64).copy()
        self.X, self.y_n = X, y_n
        self.y_mean, self.y_std, self.jitter = float(y_mean), float(y_std), float(jitter)
        self.model = self  # the reference exposes the sklearn estimator here; plotting only calls predict
        self._factor = None
        self._log_likelihood = None
        if factor is not None:  # an incrementally updated factor handed over by the surrogate
            self._factor, self._log_likelihood = factor, factor.lml
        else:
            self._build(device_inputs)

    # --- device state -------------------------------------------------------------------
    def _build(self, device_inputs=None):
        X_dev, yn_dev = device_inputs or (device.to_device(self.X), device.to_device(self.y_n))
        f = device.Factor(X_dev, yn_dev, self.X.shape[0], self.spec, self.theta, self.jitter)
        if f.info != 0:
            # sklearn raises from the final factorisation (sklearn:_gpr.py:351-362)
            raise np.linalg.LinAlgError(
                "The kernel, %r, is not returning a positive definite matrix (pivot %d). Try gradually "
                "increasing the 'alpha' parameter." % (self.spec, f.info))
        self._factor, self._log_likelihood = f, f.lml
        return f

    @property
    def factor(self):
        return self._factor if self._factor is not None else self._build()

    def release(self):
        """drop this model's reference to the O(n^2) device buffers (rebuilt on demand); the Recorder keeps every
        trial's model alive (optimiser.py:343), so the aux-optimiser calls this after choosing a point.  The surrogate
        that built the model keeps the LATEST factor alive on its own (`GpuGPSurrogate._live_factor`) so that the next
        trial can extend it in O(n^2): at most one factor stays resident per surrogate."""
        self._factor = None

    def __getstate__(self):
        s = dict(self.__dict__)
        s['_factor'] = None
        s['model'] = None
        return s

    def __setstate__(self, s):
        self.__dict__.update(s)
        self.model = self

    # --- reference contract ---------------------------------------------------------------
    def predict(self, X, return_std_dev=False, return_std=None):
        """mus (m,) [, sigmas (m,)] as NumPy (surrogates.py:332-338)"""
        if return_std is not None:  # sklearn spelling, for callers that hold `.model`
            return_std_dev = return_std
        import torch
        X = np.ascontiguousarray(np.atleast_2d(X), dtype=np.float64)
        m = X.shape[0]
        assert X.shape[1] == self.spec.d, 'X has the wrong number of columns'
        dev = device.device()
        mu = torch.empty(m, dtype=torch.float64, device=dev)
        sigma = torch.empty(m, dtype=torch.float64, device=dev)
        acq_c = device.acq_spec(_lib.ACQ_UCB, 1.0, 0.0, 0.0, self.y_mean, self.y_std)
        a = device.Acquirer(self.factor, acq_c, engine=getattr(self, 'variance_engine', None))
        a.run(device.to_device(X), m, 0, None, mu=mu, sigma=sigma)
        if return_std_dev:
            return mu.cpu().numpy(), sigma.cpu().numpy()
        return mu.cpu().numpy()

    def get_hyper_params(self):
        return np.exp(self.theta)

    def get_hyper_param_names(self):
        return self.spec.names()

    def get_log_likelihood(self):
        if self._log_likelihood is None:
            self._build()
        return self._log_likelihood

    def log_marginal_likelihood(self, theta=None, eval_gradient=False):
        """GaussianProcessRegressor.log_marginal_likelihood (sklearn:_gpr.py:541-656) on the device"""
        if theta is None and not eval_gradient:
            return self.get_log_likelihood()
        th = self.theta if theta is None else np.asarray(theta, dtype=np.float64)
        lml, grad, info = device.lml_batched(device.to_device(self.X), device.to_device(self.y_n),
                                             self.X.shape[0], self.spec, th[None, :], self.jitter, eval_gradient)
        return (float(lml[0]), grad[0]) if eval_gradient else float(lml[0])


class GpuGPSurrogate(Surrogate):
    default_model_params = {'kernel': 'default', 'normalize_y': True}

    def __init__(self, model_params=None, training_iterations=None, param_continuity=True, max_iter=200,
                 seed=None, group=None, variance_engine=None, incremental=True):
        """model_params: as for SciKitGPSurrogate -- ``kernel`` (a scikit-learn kernel of the supported
        family, a KernelSpec, or (KernelSpec, theta0[, bounds])), ``normalize_y``, ``alpha`` (jitter),
        ``n_restarts_optimizer``, ``random_state``.  The default kernel is the reference's
        ``1.0 * Matern(nu=2.5) + WhiteKernel()``."""
        self.model_params = dict(model_params) if model_params else dict(self.default_model_params)
        unknown = set(self.model_params) - {'kernel', 'normalize_y', 'alpha', 'n_restarts_optimizer',
                                            'random_state', 'optimizer'}
        if unknown:
            raise ValueError("unsupported GaussianProcessRegressor parameters: %s" % sorted(unknown))
        self.training_iterations = training_iterations
        assert training_iterations is None or self.model_params.get('n_restarts_optimizer') is None, \
            'cannot specify n_restarts_optimizer and training_iterations at the same time'
        self.param_continuity = param_continuity
        self.max_iter = max_iter
        self.group = group
        # predictive-variance engine: 'auto' (default) | 'dmma' | 'i8' | 'i8x' (device.resolve_engine)
        self.variance_engine = variance_engine
        # trials that only append rows at unchanged hyper-parameters extend the previous factor in O(n^2) (8f-2)
        self.incremental = bool(incremental)
        self._live = None
        self._live_factor = None
        rs = self.model_params.get('random_state', seed)
        self._rng = rs if isinstance(rs, np.random.RandomState) else np.random.RandomState(rs)
        self._last_model_params = None

    def _get_training_iterations(self, trial_num):
        """surrogates.py:283-292"""
        if self.training_iterations is None:
            iterations = self.model_params.get('n_restarts_optimizer')  # as the reference: used as-is
        elif callable(self.training_iterations):
            iterations = self.training_iterations(trial_num)
        else:
            iterations = self.training_iterations
        assert iterations is not None, 'must specify the number of training iterations'
        assert iterations >= 0, 'invalid number of iterations: {}'.format(iterations)
        return iterations

    def _kernel(self, d):
        k = self.model_params.get('kernel', 'default')
        if isinstance(k, str) and k == 'default':
            spec = kernels.KernelSpec(_lib.KIND_MATERN52, d, False, True, True)
            return spec, np.zeros(spec.p), spec.default_bounds()
        return _resolve_kernel(k, d)

    def construct_model(self, trial_num, X, y, start_thetas=None):
        """surrogates.py:294-326.  ``start_thetas`` (R,p) overrides the restart starting points
        (lets a CPU run and this one share them)."""
        iterations = self._get_training_iterations(trial_num)
        fitting_info = {'iterations': iterations}
        X = np.ascontiguousarray(np.atleast_2d(X), dtype=np.float64)
        n, d = X.shape
        spec, theta0, bounds = self._kernel(d)
        if self.param_continuity and self._last_model_params is not None:
            theta0 = np.log(self._last_model_params.copy())  # theta is log-transformed (:302-304)
        normalize_y = bool(self.model_params.get('normalize_y', False))
        jitter = float(self.model_params.get('alpha', 1e-10))
        y_n, y_mean, y_std = _normalise_y(y, normalize_y)

        X_dev, yn_dev = device.to_device(X), device.to_device(y_n)
        with warnings.catch_warnings(record=True) as ws:
            warnings.simplefilter('always')
            if iterations == 0 or self.model_params.get('optimizer', 'fmin_l_bfgs_b') is None:
                theta = theta0
                fitting_info['fixed'] = theta.copy()
            else:
                theta, fit_stats = self._optimise(X_dev, yn_dev, n, spec, theta0, bounds, iterations, jitter,
                                                  start_thetas)
                fitting_info.update(fit_stats)
            factor = None
            if 'fixed' in fitting_info and self.incremental:
                factor = self._try_incremental(spec, theta, X, yn_dev, jitter)
                if factor is not None:
                    theta = self._live[2].copy()  # the factor's own theta (equal up to an exp/log round trip)
                    fitting_info['incremental'] = True
            model = GpuGPSurrogate.ModelInstance(spec, theta, X, y_n, y_mean, y_std, jitter,
                                                 device_inputs=(X_dev, yn_dev), variance_engine=self.variance_engine,
                                                 factor=factor)
            self._remember(model, spec, theta, jitter)
        if len(ws) > 0:
            fitting_info['warnings'] = [w.message for w in ws]
        if iterations > 0:
            self._last_model_params = np.exp(theta.copy())
        return model, fitting_info

    def _remember(self, model, spec, theta, jitter):
        """the latest model and ITS factor: the model may drop its own reference (`release`, the aux-optimiser's default)
        while the surrogate keeps the factor for the next trial's incremental update"""
        self._live = (model, spec, np.asarray(theta, dtype=np.float64).copy(), float(jitter))
        self._live_factor = model._factor

    def __getstate__(self):
        """the Recorder pickles the optimiser with its modules (recorder.py:117-146): no device state, no process group"""
        s = dict(self.__dict__)
        s['_live'] = s['_live_factor'] = None
        if not (s.get('group') is None or isinstance(s.get('group'), str)):
            s['group'] = None
        return s

    def _try_incremental(self, spec, theta, X, yn_dev, jitter):
        """when a trial only ADDS rows to the previous trial's X and theta, kernel and jitter are unchanged
        (the reference's `training_iterations` callable returning 0, surrogates.py:250-269), the previous factor is
        extended in O(n^2) per new point instead of refactorised.  The factor object moves to the new model; the old
        model rebuilds its own lazily if plotting asks for it.  Returns None when the conditions do not hold."""
        live = getattr(self, '_live', None)
        if live is None:
            return None
        old, ospec, otheta, ojit = live
        f = old._factor if old._factor is not None else getattr(self, '_live_factor', None)
        n_old = old.X.shape[0]
        if (f is None or ospec != spec or ojit != float(jitter) or not np.allclose(otheta, theta, rtol=0.0, atol=1e-13)
                or X.shape[0] <= n_old or X.shape[1] != old.X.shape[1] or not np.array_equal(X[:n_old], old.X)):
            return None
        if f.n != n_old or not f.append(X[n_old:], yn_dev, X.shape[0], jitter):
            return None
        old._factor = None
        return f

    def _optimise(self, X_dev, yn_dev, n, spec, theta0, bounds, iterations, jitter, start_thetas):
        if start_thetas is not None:
            starts = np.atleast_2d(np.asarray(start_thetas, dtype=np.float64))
        else:
            starts = [theta0]
            if iterations > 1:
                if not np.isfinite(bounds).all():
                    raise ValueError("Multiple optimizer restarts requires that all bounds are finite.")
                for _ in range(iterations - 1):  # sklearn:_gpr.py:329-330
                    starts.append(self._rng.uniform(bounds[:, 0], bounds[:, 1]))
            starts = np.array(starts)
        R = starts.shape[0]
        rank, world = 0, 1
        if self.group is not None:
            from . import dist
            rank, world = dist.rank_world(self.group)
        mine = np.arange(R)[rank::world]  # restarts shard naturally

        def fun(T, members):
            lml, grad, info = device.lml_batched(X_dev, yn_dev, n, spec, T, jitter, True)
            f = -lml
            f[info != 0] = np.inf  # sklearn returns -inf LML on a failed Cholesky (:590-593)
            return f, -grad

        if mine.size:
            xs, fs, iters, n_eval = lbfgs.minimize_batched(fun, starts[mine], bounds, max_iter=self.max_iter)
        else:
            xs, fs, iters, n_eval = np.zeros((0, spec.p)), np.zeros(0), np.zeros(0, int), 0
        if self.group is not None:
            from . import dist
            theta, best_f = dist.select_best_restart(xs, fs, spec.p, self.group)
        else:
            if not np.isfinite(fs).any():
                raise np.linalg.LinAlgError("no restart produced a positive-definite kernel matrix")
            best = int(np.argmin(fs))
            theta, best_f = xs[best].copy(), float(fs[best])
        return theta, {'restarts': R, 'lml_evaluations': int(n_eval), 'optimiser_iterations': iters.tolist(),
                       'best_neg_lml': best_f}

    class ModelInstance(_DeviceModelInstance, SciKitModelInstance):
        """isinstance(model, turbo.modules.SciKitGPSurrogate.ModelInstance) holds (plotting/surrogates.py:67-75)"""
