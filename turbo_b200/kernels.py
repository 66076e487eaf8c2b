"""Kernel description shared by the host layer and the C ABI (tb200_kernel_spec).

The reference configures its surrogate with a scikit-learn kernel object
(turbo/modules/surrogates.py:231-243: ``1.0 * Matern(nu=2.5) + WhiteKernel()``).  The device
kernels implement the family ``[Constant *] (RBF | Matern nu in {1/2, 3/2, 5/2}) [+ White]`` with
an isotropic or ARD length scale; :func:`from_sklearn` maps a scikit-learn kernel of that form to
a :class:`KernelSpec` plus its log-space theta and bounds (same order as ``kernel.theta``).
"""
import math

import numpy as np

from . import _lib


class KernelSpec:
    def __init__(self, kind, d, ard, has_const=True, has_white=True):
        self.kind, self.d, self.ard = int(kind), int(d), bool(ard)
        self.has_const, self.has_white = bool(has_const), bool(has_white)
        if not 1 <= self.d <= _lib.MAX_D:
            raise ValueError("d must be in 1..%d" % _lib.MAX_D)
        self.n_ls = self.d if self.ard else 1
        self.p = int(self.has_const) + self.n_ls + int(self.has_white)

    def to_c(self):
        return _lib.KernelSpecC(self.kind, self.d, int(self.ard), int(self.has_const), int(self.has_white))

    def _key(self):
        return (self.kind, self.d, self.ard, self.has_const, self.has_white)

    def __eq__(self, other):
        return isinstance(other, KernelSpec) and self._key() == other._key()

    def __ne__(self, other):
        return not self.__eq__(other)

    def __hash__(self):
        return hash(self._key())

    def names(self):
        """same strings as SciKitGPSurrogate.ModelInstance.get_hyper_param_names
        (surrogates.py:350-362) gives for the equivalent scikit-learn kernel."""
        base = "k1__" if self.has_white else ""
        names = []
        if self.has_const:
            names.append(base + "k1__constant_value")
        ls = base + ("k2__" if self.has_const else "") + "length_scale"
        if not self.has_white and not self.has_const:
            ls = "length_scale"
        names.extend(["%s_%d" % (ls, i) for i in range(self.d)] if self.ard else [ls])
        if self.has_white:
            names.append("k2__noise_level")
        return names

    def default_bounds(self):
        """scikit-learn's default hyper-parameter bounds (1e-5, 1e5), log space"""
        return np.tile(np.log([1e-5, 1e5]), (self.p, 1))

    def __getstate__(self):
        return dict(kind=self.kind, d=self.d, ard=self.ard, has_const=self.has_const, has_white=self.has_white)

    def __setstate__(self, s):
        self.__init__(**s)

    def __repr__(self):
        fam = {0: "RBF", 1: "Matern12", 2: "Matern32", 3: "Matern52"}[self.kind]
        return "KernelSpec(%s%s, d=%d, %s%s)" % ("Const*" if self.has_const else "", fam, self.d,
                                                  "ARD" if self.ard else "iso", "+White" if self.has_white else "")


def from_sklearn(kernel, d):
    """-> (KernelSpec, theta0 (p,), log-bounds (p,2)) for a supported scikit-learn kernel."""
    import sklearn.gaussian_process.kernels as K

    white = const = None
    k = kernel
    if isinstance(k, K.Sum):
        if isinstance(k.k2, K.WhiteKernel):
            k, white = k.k1, k.k2
        elif isinstance(k.k1, K.WhiteKernel):
            raise ValueError("write the WhiteKernel as the second term (theta order follows the kernel tree)")
        else:
            raise ValueError("unsupported Sum kernel: %r" % (kernel,))
    if isinstance(k, K.Product):
        if not isinstance(k.k1, K.ConstantKernel):
            raise ValueError("unsupported Product kernel (expected ConstantKernel * stationary): %r" % (kernel,))
        const, k = k.k1, k.k2
    if isinstance(k, K.Matern):  # Matern subclasses RBF: test it first
        kind = {0.5: _lib.KIND_MATERN12, 1.5: _lib.KIND_MATERN32, 2.5: _lib.KIND_MATERN52,
                math.inf: _lib.KIND_RBF}.get(float(k.nu))
        if kind is None:
            raise ValueError("Matern nu must be 0.5, 1.5, 2.5 or inf (got %r)" % (k.nu,))
    elif isinstance(k, K.RBF):
        kind = _lib.KIND_RBF
    else:
        raise ValueError("unsupported stationary kernel: %r" % (k,))
    for part in (const, k, white):
        if part is not None and any(h.fixed for h in part.hyperparameters):
            raise ValueError("fixed hyper-parameters are not supported: %r" % (part,))
    ard = bool(k.anisotropic)
    if ard and len(np.atleast_1d(k.length_scale)) != d:
        raise ValueError("anisotropic length_scale has %d entries for %d dimensions"
                         % (len(np.atleast_1d(k.length_scale)), d))
    spec = KernelSpec(kind, d, ard, const is not None, white is not None)
    theta0 = np.asarray(kernel.theta, dtype=np.float64).copy()
    bounds = np.asarray(kernel.bounds, dtype=np.float64).copy()
    assert theta0.shape == (spec.p,), (theta0.shape, spec.p)
    return spec, theta0, bounds
