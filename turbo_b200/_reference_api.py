"""The reference's plugin base classes, or interface-identical stand-ins.

turbo.plotting / turbo.recorder decide what they can draw with ``isinstance`` checks against
``turbo.modules`` classes (plotting/surrogates.py:67-75, plotting/overview.py:53-61), so when
mbway/turbo is importable our plugins SUBCLASS its classes.  ``turbo`` is looked for on the normal
import path first, then in the directory named by ``TB200_TURBO_PATH``, then in ``baseline/_ref`` next to
this package (the unmodified reference installed by ``baseline/install_reference.py``, which is how the
GPU box gets it).  Only on a machine with none of those are the names bound to minimal stand-ins with the
same constructor arguments and method names, so the host layer still imports.
"""
import importlib
import os
import sys
import warnings


def _import_turbo_modules():
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", SyntaxWarning)  # the reference's docstrings hold '\m' escapes
        return _import_turbo_modules_impl()


def _import_turbo_modules_impl():
    try:
        return importlib.import_module("turbo.modules")
    except Exception:
        pass
    here = os.path.dirname(os.path.abspath(__file__))
    for cand in (os.environ.get("TB200_TURBO_PATH"), os.path.join(os.path.dirname(here), "baseline", "_ref")):
        if cand and os.path.isfile(os.path.join(cand, "turbo", "__init__.py")) and cand not in sys.path:
            sys.path.append(cand)
            for k in [k for k in sys.modules if k == "turbo" or k.startswith("turbo.")]:
                del sys.modules[k]  # a half-imported package from the failed attempt
            try:
                return importlib.import_module("turbo.modules")
            except Exception:
                sys.path.remove(cand)
    return None


_tm = _import_turbo_modules()
if _tm is not None:  # pragma: no cover - depends on the environment
    HAVE_TURBO = True
    Surrogate = _tm.Surrogate
    SciKitModelInstance = _tm.SciKitGPSurrogate.ModelInstance
    GPyModelInstance = _tm.GPySurrogate.ModelInstance  # defined even when GPy itself is missing
    EI, PI, UCB = _tm.EI, _tm.PI, _tm.UCB
else:  # turbo (or one of its imports) is not available
    HAVE_TURBO = False

    class Surrogate:
        """turbo/modules/surrogates.py:22-81"""

        def construct_model(self, trial_num, X, y):
            raise NotImplementedError()

        class ModelInstance:
            def predict(self, X, return_std_dev=False):
                raise NotImplementedError()

            def get_hyper_params(self):
                raise NotImplementedError()

            def get_hyper_param_names(self):
                raise NotImplementedError()

            def get_log_likelihood(self):
                raise NotImplementedError()

    SciKitModelInstance = Surrogate.ModelInstance

    class GPyModelInstance(Surrogate.ModelInstance):
        """turbo/modules/surrogates.py:199-220"""

    class _AcquisitionFunction:
        """turbo/modules/acquisition_functions.py:12-77"""

        class FunctionInstance:
            def __init__(self, model, desired_extremum):
                self.model = model
                assert desired_extremum in ('min', 'max')
                self.desired_extremum = desired_extremum

    class UCB(_AcquisitionFunction):
        """acquisition_functions.py:80-158"""

        def __init__(self, beta):
            self.beta = beta

        def get_type(self):
            return 'optimism'

        class FunctionInstance(_AcquisitionFunction.FunctionInstance):
            def __init__(self, model, desired_extremum, beta):
                super().__init__(model, desired_extremum)
                self.beta = beta
                self.scale_factor = 1 if desired_extremum == 'max' else -1

            def get_name(self):
                return 'UCB' if self.desired_extremum == 'max' else '-LCB'

    class _Improvement(_AcquisitionFunction):
        def __init__(self, xi):
            self.xi = xi

        def get_type(self):
            return 'improvement'

        class FunctionInstance(_AcquisitionFunction.FunctionInstance):
            def __init__(self, model, desired_extremum, incumbent_cost, xi):
                super().__init__(model, desired_extremum)
                self.incumbent_cost = incumbent_cost
                self.xi = xi
                self.scale_factor = 1 if desired_extremum == 'max' else -1

    class PI(_Improvement):
        """acquisition_functions.py:163-247"""

        class FunctionInstance(_Improvement.FunctionInstance):
            def get_name(self):
                return 'PI'

    class EI(_Improvement):
        """acquisition_functions.py:250-358"""

        class FunctionInstance(_Improvement.FunctionInstance):
            def get_name(self):
                return 'EI'
