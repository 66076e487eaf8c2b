"""GpuCandidateArgmax: drop-in for turbo.modules.RandomAndQuasiNewton's random stage
(turbo/modules/auxiliary_optimisers.py:36-66, :117-129).

``aux_optimiser(latent_bounds, acq) -> (x (1,d), {'max_acq': float})`` with the whole
evaluate-and-argmax step fused on the device through ``acq.argmax_over``.  The reference default is
1000 uniform candidates followed by L-BFGS-B polishing; here the candidate set is scaled up by
orders of magnitude (10^6..10^8), which is what the north star asks of this stage, and the polishing
stage (auxiliary_optimisers.py:69-112) is available as ``grad_restarts > 0``: the starts are the best
``start_from_best`` candidates of the random stage plus fresh uniform points, all advanced TOGETHER by
the lock-step batched L-BFGS-B of ``lbfgs.py`` on analytic gradients from ``tb200_acquire_grad`` (the
reference runs SciPy's L-BFGS-B per start on numerical gradients, which is broken on current SciPy:
SURVEY 8c).  The polished point replaces the random-stage winner only if its acquisition value is
higher (auxiliary_optimisers.py:108-110).
"""
import math

import numpy as np


class GpuCandidateArgmax:
    def __init__(self, num_random=1 << 20, candidates='uniform', seed=0, group=None, release_model=True,
                 grad_restarts=0, start_from_best=2, polish_max_iter=100, close_tolerance=-1.0):
        """candidates:
             'uniform' -- generated on the device (Philox keyed by seed, trial counter and global index);
             'lhs'     -- Latin hypercube on the device (naive_selectors.py:65-83 semantics);
             'numpy'   -- drawn on the host exactly as the reference's random_selector does
                          (naive_selectors.py:41-46: one np.random.uniform column per bound), then
                          streamed to the device: reproduces the reference's candidate sequence;
             callable  -- ``f(num_points, latent_bounds) -> (m,d)`` (any reference selector).
        close_tolerance >= 0: candidates whose squared distance to a training point is <= close_tolerance
        (``Fallback.close_tolerance``, turbo/modules/fallback.py:49-54 -> turbo/utils.py:111-130) are skipped by the
        device argmax, so the optimiser's too-close test (turbo/optimiser.py:349-352) never has to throw the Bayesian
        trial away; a polished point that ends up too close is dropped for the random-stage winner.  Default -1:
        off, the reference's behaviour.
        release_model: free the model's O(n^2) device buffers after the argmax (the Recorder keeps
        every trial's model object alive; they are rebuilt lazily if plotting calls predict later)."""
        self.num_random = int(num_random)
        self.candidates = candidates
        self.seed = int(seed)
        self.group = group
        self.release_model = release_model
        self.grad_restarts = int(grad_restarts)
        self.start_from_best = min(int(start_from_best), self.grad_restarts)
        self.polish_max_iter = int(polish_max_iter)
        self.close_tolerance = float(close_tolerance)
        assert self.start_from_best <= self.num_random
        self._calls = 0

    def __call__(self, latent_bounds, acq):
        if not hasattr(acq, 'argmax_over'):
            raise TypeError('GpuCandidateArgmax needs a turbo_b200 acquisition function (GpuEI/GpuPI/GpuUCB); '
                            'got %r -- there is no CPU fallback' % (type(acq),))
        lo = [b[1] for b in latent_bounds.ordered]
        hi = [b[2] for b in latent_bounds.ordered]
        maximisation_info = {}
        keep = self.start_from_best if self.grad_restarts > 0 else 0
        top = None
        seed = (self.seed * 1000003 + self._calls) & (2 ** 63 - 1)  # a fresh set every trial
        exclude = (acq.model.X, self.close_tolerance) if self.close_tolerance >= 0 else None
        if self.candidates in ('uniform', 'lhs'):
            res = acq.argmax_over(generator=self.candidates, m=self.num_random, lo=lo, hi=hi, seed=seed,
                                  group=self.group, keep_top=keep, exclude=exclude)
        else:
            if self.candidates == 'numpy':
                cols = [np.random.uniform(pmin, pmax, size=(self.num_random, 1)) for pmin, pmax in zip(lo, hi)]
                X = np.hstack(cols)
            else:
                X = np.asarray(self.candidates(self.num_random, latent_bounds), dtype=np.float64)
            if self.group is not None:
                # host-drawn sets come from each process's own generator: argmax_over shards ONE set by index, so the
                # set rank 0 drew is the set (every rank still consumes its generator, keeping the streams aligned)
                from . import dist
                X = dist.broadcast_array(X, self.group, src=0)
            res = acq.argmax_over(candidates=X, group=self.group, keep_top=keep, exclude=exclude)
        val, idx, x = res[:3]
        if keep:
            top = res[3]
        self._calls += 1
        polished = False
        if self.grad_restarts > 0 and x is not None:
            x_random = x
            val, x = self._polish(acq, lo, hi, val, idx, x, top, seed, maximisation_info)
            polished = x is not x_random
        if self.release_model and hasattr(acq.model, 'release'):
            acq.model.release()
        if x is None:  # auxiliary_optimisers.py:117-118
            maximisation_info['max_acq'] = -math.inf
            return None, maximisation_info
        x = np.clip(x, lo, hi)  # :120-124
        maximisation_info.update({'max_acq': float(val), 'num_candidates': self.num_random})
        # the index belongs to the random stage: a polished point is not a member of the candidate set
        maximisation_info['random_stage_argmax_index' if polished else 'argmax_index'] = int(idx)
        return x, maximisation_info

    def __getstate__(self):
        """pickled with the optimiser by the Recorder (recorder.py:117-146): a process-group object does not pickle"""
        s = dict(self.__dict__)
        if not (s.get('group') is None or isinstance(s.get('group'), str)):
            s['group'] = None
        return s

    def _polish(self, acq, lo, hi, val, idx, x, top, seed, info):
        """gradient stage of auxiliary_optimisers.py:69-112 on the device: every start advances in lock step"""
        from . import lbfgs
        lo, hi = np.asarray(lo, dtype=np.float64), np.asarray(hi, dtype=np.float64)
        starts = [np.atleast_2d(x)]
        if top is not None and len(top[1]):
            starts.append(top[2][top[1] != idx][:max(self.start_from_best - 1, 0)])
        have = sum(len(s) for s in starts)
        if self.grad_restarts > have:  # the rest are fresh uniform points, identical on every rank
            rng = np.random.RandomState(seed % (2 ** 32))
            starts.append(rng.uniform(lo, hi, size=(self.grad_restarts - have, len(lo))))
        x0 = np.vstack(starts)[:self.grad_restarts]

        def fun(Xq, members):
            v, g = acq.value_and_grad(Xq)
            return -v, -g

        xs, fs, n_iter, n_eval = lbfgs.minimize_batched(fun, x0, np.stack([lo, hi], axis=1),
                                                        max_iter=self.polish_max_iter)
        fs = np.where(np.isfinite(fs), fs, np.inf)
        j = int(np.argmin(fs))
        info.update({'polish_starts': len(x0), 'polish_evaluations': int(n_eval), 'random_stage_max_acq': float(val)})
        if -fs[j] > val:  # :108-110 keep the better of the two stages
            if self.close_tolerance >= 0:
                # same test, same arithmetic as turbo/utils.py:130 on the one polished point
                if np.any(np.sum((acq.model.X - xs[j:j + 1]) ** 2, axis=1) <= self.close_tolerance):
                    info['polish_rejected'] = 'too_close'
                    return val, x
            return float(-fs[j]), xs[j:j + 1].copy()
        return val, x
