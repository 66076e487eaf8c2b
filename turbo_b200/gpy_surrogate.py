"""GpuGPySurrogate: the reference's DEFAULT surrogate, turbo.modules.GPySurrogate
(turbo/modules/surrogates.py:84-222; selected by turbo/optimiser_presets.py:16), on the device.

GPy itself is a third-party dependency that is absent from this image (and un-pinned in the reference's
requirements.txt), so this mode restates GPy's published ``GPRegression`` model and cannot be pinned against a live
GPy run: PARITY UNPINNED.  What it mirrors, with the GPy/turbo lines it follows:

* model (GPy/models/gp_regression.py, GPy/core/gp.py): ``K = variance * g(r / lengthscale) + noise_variance * I``,
  no jitter; the kernel variance lives inside the kernel and the Gaussian likelihood carries the noise term
  (surrogates.py:91-95).  This is the device path's ``Constant * kernel + White`` family with ``alpha = 0``.
* ``normalizer=True`` (surrogates.py:98): GPy's ``Standardize`` -- ``(y - mean) / std`` with the population std,
  predictions mapped back with ``mean * std + mean`` and ``var * std**2``.
* ``predict`` returns the mean and the variance INCLUDING the noise term (GPy predicts through the likelihood),
  square-rooted by surrogates.py:204-210.
* hyper-parameters are reported untransformed, in GPy's order and with GPy's names:
  ``<kern>.variance, <kern>.lengthscale, Gaussian_noise.variance`` (surrogates.py:212-216).
* ``training_iterations`` = GPy's ``num_restarts`` (surrogates.py:185-186: restarts == iterations, not iterations-1);
  the first restart starts from the current parameters, the others from paramz's ``randomize()``: a standard normal
  draw in the optimiser's space, i.e. ``softplus(N(0,1))`` for GPy's default ``Logexp`` positive constraint; the
  restart with the lowest objective wins (paramz/model.py optimize_restarts).  ``iterations == 0`` keeps the
  parameters fixed and reports them under ``fitting_info['fixed']`` (surrogates.py:176-177).
* ``param_continuity`` (surrogates.py:171-173): last trial's parameter array is the next trial's starting point.

Not mirrored: the optimiser's path.  GPy minimises over softplus-transformed parameters with SciPy's L-BFGS-B and no
bounds; here the same objective is minimised over log-parameters by the lock-step batched L-BFGS-B of ``lbfgs.py``
inside wide bounds (1e-8, 1e8), so the iterates differ even though objective, starts and restart count agree.
``sparse=True`` (SparseGPRegression, inducing points) is outside the hot path and raises.
"""
import warnings

import numpy as np

from . import _lib, device, kernels
from ._reference_api import GPyModelInstance
from .surrogate import GpuGPSurrogate, _DeviceModelInstance, _normalise_y

# GPy kernel class name -> (device kernel kind, GPy's default parameter-name prefix)
_GPY_KERNELS = {
    'rbf': (_lib.KIND_RBF, 'rbf'), 'RBF': (_lib.KIND_RBF, 'rbf'),
    'Matern52': (_lib.KIND_MATERN52, 'Mat52'), 'Mat52': (_lib.KIND_MATERN52, 'Mat52'),
    'Matern32': (_lib.KIND_MATERN32, 'Mat32'), 'Mat32': (_lib.KIND_MATERN32, 'Mat32'),
    'Exponential': (_lib.KIND_MATERN12, 'Exponential'), 'OU': (_lib.KIND_MATERN12, 'Exponential'),
}
_LOG_BOUNDS = np.log([1e-8, 1e8])


def gpy_kernel(name='rbf', ARD=False, variance=1.0, lengthscale=1.0):
    """stand-in for ``GPy.kern.<name>(input_dim, variance, lengthscale, ARD)``: what ``model_params['kernel']`` takes
    when GPy is not installed.  A real GPy kernel object is accepted too (its class name, ARD flag, variance and
    lengthscale are read)."""
    if name not in _GPY_KERNELS:
        raise ValueError("unsupported GPy kernel %r (one of %s)" % (name, sorted(set(_GPY_KERNELS))))
    return dict(name=name, ARD=bool(ARD), variance=float(variance), lengthscale=lengthscale)


def _describe_kernel(k, d):
    """-> (KernelSpec, prefix, raw parameter array [variance, lengthscale.., noise placeholder excluded])"""
    if k is None:  # GPRegression's default: GPy.kern.RBF(input_dim)
        k = gpy_kernel()
    if not isinstance(k, dict):  # a GPy kernel object: read the public attributes GPy documents
        k = dict(name=type(k).__name__, ARD=bool(getattr(k, 'ARD', False)),
                 variance=float(np.asarray(k.variance).reshape(-1)[0]),
                 lengthscale=np.asarray(k.lengthscale, dtype=np.float64).reshape(-1))
        if k['name'] not in _GPY_KERNELS:
            raise ValueError("unsupported GPy kernel %r" % k['name'])
    kind, prefix = _GPY_KERNELS[k['name']]
    spec = kernels.KernelSpec(kind, d, k['ARD'], True, True)
    ls = np.broadcast_to(np.asarray(k['lengthscale'], dtype=np.float64).reshape(-1), (spec.n_ls,)).copy()
    return spec, prefix, np.r_[k['variance'], ls]


class GpuGPySurrogate(GpuGPSurrogate):
    default_model_params = {'normalizer': True}            # surrogates.py:98
    default_optimise_params = {'parallel': True, 'verbose': False}  # surrogates.py:99

    def __init__(self, model_params=None, optimise_params=None, training_iterations=10, param_continuity=True,
                 sparse=False, seed=None, group=None, variance_engine=None, incremental=True):
        """same arguments as GPySurrogate.__init__ (surrogates.py:101-147).  model_params: ``kernel`` (``gpy_kernel(..)``
        or a GPy kernel; default RBF), ``normalizer`` (bool), ``noise_var`` (GPRegression's initial noise variance,
        1.0).  optimise_params: ``num_restarts``, ``max_iters`` (GPy: 1000); ``parallel``/``verbose``/``robust``/
        ``messages`` are accepted and have no meaning here (every restart already advances in one batched call)."""
        if sparse:
            raise NotImplementedError('SparseGPRegression (inducing points) is not part of the device path')
        mp = dict(model_params) if model_params else dict(self.default_model_params)
        unknown = set(mp) - {'kernel', 'normalizer', 'noise_var'}
        if unknown:
            raise ValueError("unsupported GPRegression parameters: %s" % sorted(unknown))
        self.optimise_params = dict(optimise_params) if optimise_params else dict(self.default_optimise_params)
        unknown = set(self.optimise_params) - {'num_restarts', 'max_iters', 'parallel', 'verbose', 'robust', 'messages',
                                               'optimizer'}
        if unknown:
            raise ValueError("unsupported optimize_restarts parameters: %s" % sorted(unknown))
        if self.optimise_params.get('optimizer', 'lbfgsb') not in ('lbfgsb', 'lbfgs', 'bfgs'):
            raise ValueError("only GPy's default L-BFGS-B optimiser is mirrored")
        assert training_iterations is None or self.optimise_params.get('num_restarts') is None, \
            'cannot specify num_restarts and training_iterations at the same time'  # surrogates.py:138-139
        super().__init__(model_params={'kernel': 'default'}, training_iterations=training_iterations,
                         param_continuity=param_continuity, max_iter=int(self.optimise_params.get('max_iters', 1000)),
                         seed=seed, group=group, variance_engine=variance_engine, incremental=incremental)
        self.model_params = mp
        self.sparse = False

    def _get_training_iterations(self, trial_num):
        """surrogates.py:149-158"""
        if self.training_iterations is None:
            iterations = self.optimise_params.get('num_restarts')
        elif callable(self.training_iterations):
            iterations = self.training_iterations(trial_num)
        else:
            iterations = self.training_iterations
        assert iterations is not None, 'must specify the number of training iterations'
        assert iterations >= 0, 'invalid number of iterations: {}'.format(iterations)
        return iterations

    def construct_model(self, trial_num, X, y, start_thetas=None):
        """surrogates.py:160-197"""
        iterations = self._get_training_iterations(trial_num)
        fitting_info = {'iterations': iterations}
        X = np.ascontiguousarray(np.atleast_2d(X), dtype=np.float64)
        n, d = X.shape
        spec, prefix, kparams = _describe_kernel(self.model_params.get('kernel'), d)
        params = np.r_[kparams, float(self.model_params.get('noise_var', 1.0))]
        if self.param_continuity and self._last_model_params is not None:
            params = np.asarray(self._last_model_params, dtype=np.float64).copy()  # model[:] = last (:172-173)
        theta0 = np.log(params)
        bounds = np.tile(_LOG_BOUNDS, (spec.p, 1))
        y_n, y_mean, y_std = _normalise_y(y, bool(self.model_params.get('normalizer', False)))
        jitter = 0.0
        X_dev, yn_dev = device.to_device(X), device.to_device(y_n)
        with warnings.catch_warnings(record=True) as ws:
            warnings.simplefilter('always')
            if iterations == 0:
                theta = theta0
                fitting_info['fixed'] = params.copy()  # :176-177
            else:
                if start_thetas is None:
                    # paramz randomize(): N(0,1) in the optimiser's space = softplus of it in parameter space
                    starts = [theta0]
                    for _ in range(iterations - 1):
                        starts.append(np.log(np.logaddexp(0.0, self._rng.normal(0.0, 1.0, size=spec.p))))
                    start_thetas = np.clip(np.array(starts), bounds[:, 0], bounds[:, 1])
                theta, fit_stats = self._optimise(X_dev, yn_dev, n, spec, theta0, bounds, iterations, jitter,
                                                  start_thetas)
                fitting_info.update(fit_stats)
            factor = None
            if 'fixed' in fitting_info and self.incremental:
                factor = self._try_incremental(spec, theta, X, yn_dev, jitter)
                if factor is not None:
                    theta = self._live[2].copy()
                    fitting_info['incremental'] = True
            model = GpuGPySurrogate.ModelInstance(spec, theta, X, y_n, y_mean, y_std, jitter,
                                                  device_inputs=(X_dev, yn_dev), variance_engine=self.variance_engine,
                                                  factor=factor)
            model.kernel_prefix = prefix
            self._remember(model, spec, theta, jitter)
        if len(ws) > 0:
            fitting_info['warnings'] = [w.message for w in ws]
        if iterations > 0:
            self._last_model_params = np.exp(theta.copy())  # model.param_array[:] (:195)
        return model, fitting_info

    class ModelInstance(_DeviceModelInstance, GPyModelInstance):
        """surrogates.py:199-220 contract on the device state of GpuGPSurrogate.ModelInstance"""
        kernel_prefix = 'rbf'

        def get_hyper_params(self):
            """``model.param_array[:]``: untransformed [variance, lengthscale.., Gaussian_noise.variance]"""
            return np.exp(self.theta)

        def get_hyper_param_names(self):
            """GPy's ``parameter_names()``; an ARD lengthscale is spelt out per dimension so that the list is as long
            as ``get_hyper_params()`` (plotting/surrogates.py:64 asserts that)"""
            k = self.kernel_prefix
            ls = ['%s.lengthscale[%d]' % (k, i) for i in range(self.spec.d)] if self.spec.ard else [k + '.lengthscale']
            return [k + '.variance'] + ls + ['Gaussian_noise.variance']
