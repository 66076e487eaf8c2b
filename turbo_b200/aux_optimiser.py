"""GpuCandidateArgmax: drop-in for turbo.modules.RandomAndQuasiNewton's random stage
(turbo/modules/auxiliary_optimisers.py:36-66, :117-129).

``aux_optimiser(latent_bounds, acq) -> (x (1,d), {'max_acq': float})`` with the whole
evaluate-and-argmax step fused on the device through ``acq.argmax_over``.  The reference default is
1000 uniform candidates followed by L-BFGS-B polishing; here the candidate set is scaled up by
orders of magnitude instead (10^6..10^8), which is what the north star asks of this stage.
"""
import math

import numpy as np


class GpuCandidateArgmax:
    def __init__(self, num_random=1 << 20, candidates='uniform', seed=0, group=None, release_model=True):
        """candidates:
             'uniform' -- generated on the device (Philox keyed by seed, trial counter and global index);
             'lhs'     -- Latin hypercube on the device (naive_selectors.py:65-83 semantics);
             'numpy'   -- drawn on the host exactly as the reference's random_selector does
                          (naive_selectors.py:41-46: one np.random.uniform column per bound), then
                          streamed to the device: reproduces the reference's candidate sequence;
             callable  -- ``f(num_points, latent_bounds) -> (m,d)`` (any reference selector).
        release_model: free the model's O(n^2) device buffers after the argmax (the Recorder keeps
        every trial's model object alive; they are rebuilt lazily if plotting calls predict later)."""
        self.num_random = int(num_random)
        self.candidates = candidates
        self.seed = int(seed)
        self.group = group
        self.release_model = release_model
        self._calls = 0

    def __call__(self, latent_bounds, acq):
        if not hasattr(acq, 'argmax_over'):
            raise TypeError('GpuCandidateArgmax needs a turbo_b200 acquisition function (GpuEI/GpuPI/GpuUCB); '
                            'got %r -- there is no CPU fallback' % (type(acq),))
        lo = [b[1] for b in latent_bounds.ordered]
        hi = [b[2] for b in latent_bounds.ordered]
        maximisation_info = {}
        if self.candidates in ('uniform', 'lhs'):
            seed = (self.seed * 1000003 + self._calls) & (2 ** 63 - 1)  # a fresh set every trial
            val, idx, x = acq.argmax_over(generator=self.candidates, m=self.num_random, lo=lo, hi=hi, seed=seed,
                                          group=self.group)
        else:
            if self.candidates == 'numpy':
                cols = [np.random.uniform(pmin, pmax, size=(self.num_random, 1)) for pmin, pmax in zip(lo, hi)]
                X = np.hstack(cols)
            else:
                X = np.asarray(self.candidates(self.num_random, latent_bounds), dtype=np.float64)
            val, idx, x = acq.argmax_over(candidates=X, group=self.group)
        self._calls += 1
        if self.release_model and hasattr(acq.model, 'release'):
            acq.model.release()
        if x is None:  # auxiliary_optimisers.py:117-118
            maximisation_info['max_acq'] = -math.inf
            return None, maximisation_info
        x = np.clip(x, lo, hi)  # :120-124
        maximisation_info.update({'max_acq': float(val), 'argmax_index': int(idx), 'num_candidates': self.num_random})
        return x, maximisation_info
