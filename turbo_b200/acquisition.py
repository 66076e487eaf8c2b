"""GpuEI / GpuPI / GpuUCB: drop-ins for turbo.modules.EI / PI / UCB
(turbo/modules/acquisition_functions.py:80-358).

The factories subclass the reference's so ``isinstance(acq, tm.EI)`` checks in turbo.plotting keep
working (plotting/overview.py:53-61, plotting/trials.py:373).  Each FunctionInstance offers

* ``__call__(X) -> (m,) ndarray``: the reference contract (values for host points), and
* ``argmax_over(...) -> (value, index, x)``: the fused device path -- predict, acquisition and argmax
  in one pass over a candidate set that is either given (host array, streamed in chunks) or generated
  on the device from (seed, global index), optionally sharded across ranks.
"""
import math

import numpy as np

from . import _lib, device
from ._reference_api import EI, PI, UCB


class _DeviceFunction:
    """mixin: everything that talks to the device for one (model, acquisition) pair"""

    _acq_code = None

    def _acq_param(self):
        raise NotImplementedError()

    def _spec_c(self):
        m = self.model
        return device.acq_spec(self._acq_code, self.scale_factor, getattr(self, 'incumbent_cost', 0.0),
                               self._acq_param(), m.y_mean, m.y_std)

    def __call__(self, X):
        import torch
        X = np.ascontiguousarray(np.atleast_2d(X), dtype=np.float64)
        m = X.shape[0]
        out = torch.empty(m, dtype=torch.float64, device=device.device())
        a = device.Acquirer(self.model.factor, self._spec_c(), engine=getattr(self.model, 'variance_engine', None))
        a.run(device.to_device(X), m, 0, None, acq=out)
        return out.cpu().numpy()

    def value_and_grad(self, X):
        """acquisition values (P,) and analytic gradients (P,d) at host points X (tb200_acquire_grad): what the
        L-BFGS-B stage of RandomAndQuasiNewton needs (auxiliary_optimisers.py:84-93 uses numerical gradients)"""
        return device.acquisition_value_and_grad(self.model.factor, self._spec_c(), X)

    def argmax_over(self, candidates=None, generator=None, m=None, lo=None, hi=None, seed=0, group=None,
                    chunk=1 << 18, batch=0, engine=None, keep_top=0, exclude=None):
        """argmax of the acquisition function over a candidate set.

        candidates: (m,d) host array (every rank passes the same full set when ``group`` is given), or
        generator : 'uniform' | 'lhs' with ``m``, ``lo``, ``hi``, ``seed`` -- generated on the device
                    from the global index, so the result does not depend on how ranks split the range.
        Returns (value, global index, x (1,d)); (-inf, -1, None) if no candidate is selectable.
        Ties keep the lowest index; NaN never wins (auxiliary_optimisers.py:63-66 picks argsort's first).
        exclude = (X (k,d), tol): candidates whose squared Euclidean distance to a row of X is <= tol are never selected
        (turbo/utils.py:111-130 `close_to_any`, applied to every candidate on the device in the same pass instead of
        to the winner afterwards, turbo/optimiser.py:349-352); tol < 0 disables it (fallback.py:51-52).
        keep_top > 0: additionally returns (values, indices, X) of up to keep_top distinct good candidates, best
        first (the winner, then per-CTA running bests merged over chunks and ranks) as a 4th element.
        """
        import torch
        from . import dist
        rank, world = dist.rank_world(group)
        exclusion = None
        if exclude is not None and exclude[1] >= 0 and len(exclude[0]):
            exclusion = (device.to_device(np.ascontiguousarray(np.atleast_2d(exclude[0]), dtype=np.float64)),
                         float(exclude[1]))
        a = device.Acquirer(self.model.factor, self._spec_c(), batch,
                            engine=engine or getattr(self.model, 'variance_engine', None), exclusion=exclusion)
        d = self.model.spec.d
        if candidates is not None:
            if isinstance(candidates, torch.Tensor):  # a (pinned) host tensor is used in place
                assert candidates.device.type == 'cpu' and candidates.dtype == torch.float64
                X = candidates.contiguous().numpy()
            else:
                X = np.ascontiguousarray(np.atleast_2d(candidates), dtype=np.float64)
            m = X.shape[0]
            lo_i, hi_i = dist.shard_range(m, rank, world)
            host = candidates[lo_i:hi_i] if isinstance(candidates, torch.Tensor) else torch.from_numpy(X[lo_i:hi_i])
            if host.numel() * 8 >= (1 << 20) and not host.is_pinned():
                host = host.pin_memory()
            # stream the shard in chunks: H2D of chunk i+1 overlaps the kernels of chunk i
            tops = []
            a.run_host(host, lo_i, chunk, (lambda: tops.append(a.local_bests())) if keep_top else None)
            cand_c = None
        else:
            mode = {'uniform': _lib.CAND_UNIFORM, 'lhs': _lib.CAND_LHS}[generator]
            cand_c = _lib.cand_spec(mode, seed, m, lo, hi)
            lo_i, hi_i = dist.shard_range(m, rank, world)
            tops = []
            if hi_i > lo_i:
                a.run(None, hi_i - lo_i, lo_i, cand_c)
                if keep_top:
                    tops.append(a.local_bests())
        val, idx = a.best()
        val, idx = dist.allreduce_argmax(val, idx, group)
        if idx < 0:
            return (-math.inf, -1, None, None) if keep_top else (-math.inf, -1, None)

        def rows(ids):
            if candidates is not None:
                return X[np.asarray(ids, dtype=np.int64)].copy()
            return _lib.host_candidates(cand_c.mode, seed, m, lo, hi, [int(i) for i in ids])

        x = rows([idx])
        if not keep_top:
            return val, idx, x
        tv = np.concatenate([t[0] for t in tops]) if tops else np.zeros(0)
        ti = np.concatenate([t[1] for t in tops]) if tops else np.zeros(0, dtype=np.int64)
        tv, ti = dist.allgather_top(tv, ti, keep_top, group)
        return val, idx, x, (tv, ti, rows(ti))


class GpuUCB(UCB):
    def construct_function(self, trial_num, model, desired_extremum):
        beta = self.beta(trial_num) if callable(self.beta) else self.beta
        return GpuUCB.FunctionInstance(model, desired_extremum, beta), {'beta': beta}

    class FunctionInstance(_DeviceFunction, UCB.FunctionInstance):
        _acq_code = _lib.ACQ_UCB

        def _acq_param(self):
            return self.beta


class GpuPI(PI):
    def construct_function(self, trial_num, model, desired_extremum, incumbent_cost):
        xi = self.xi(trial_num) if callable(self.xi) else self.xi
        return GpuPI.FunctionInstance(model, desired_extremum, incumbent_cost, xi), {'xi': xi}

    class FunctionInstance(_DeviceFunction, PI.FunctionInstance):
        _acq_code = _lib.ACQ_PI

        def _acq_param(self):
            return self.xi


class GpuEI(EI):
    def construct_function(self, trial_num, model, desired_extremum, incumbent_cost):
        xi = self.xi(trial_num) if callable(self.xi) else self.xi
        return GpuEI.FunctionInstance(model, desired_extremum, incumbent_cost, xi), {'xi': xi}

    class FunctionInstance(_DeviceFunction, EI.FunctionInstance):
        _acq_code = _lib.ACQ_EI

        def _acq_param(self):
            return self.xi
