"""The reference's plugin base classes, or interface-identical stand-ins.

turbo.plotting / turbo.recorder decide what they can draw with ``isinstance`` checks against
``turbo.modules`` classes (plotting/surrogates.py:67-75, plotting/overview.py:53-61), so when
mbway/turbo is importable our plugins SUBCLASS its classes.  On a machine without the reference
(the GPU box) the same names are bound to minimal stand-ins with the same constructor arguments
and method names, so the host layer and its tests run unchanged.
"""
try:  # pragma: no cover - depends on the environment
    import turbo.modules as _tm

    HAVE_TURBO = True
    Surrogate = _tm.Surrogate
    SciKitModelInstance = _tm.SciKitGPSurrogate.ModelInstance
    GPyModelInstance = _tm.GPySurrogate.ModelInstance  # defined even when GPy itself is missing
    EI, PI, UCB = _tm.EI, _tm.PI, _tm.UCB
except Exception:  # turbo (or one of its imports) is not available
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
