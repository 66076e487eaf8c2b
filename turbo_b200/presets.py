"""Optimiser presets that swap the B200 surrogate + acquisition in (turbo/optimiser_presets.py:8-37).

``Optimiser.__setattr__`` only lets the six module slots be assigned (turbo/optimiser.py:98-110);
this assigns exactly those, so the recorder and plotting see a normally configured optimiser."""
from .acquisition import GpuEI
from .aux_optimiser import GpuCandidateArgmax
from .gpy_surrogate import GpuGPySurrogate
from .surrogate import GpuGPSurrogate


def load_gpu_preset(optimiser, name='default', num_random=1 << 20, training_iterations=10, group=None,
                    surrogate='sklearn', close_tolerance=-1.0, candidates='uniform', seed=0, model_params=None):
    """surrogate: 'sklearn' -- the reference's commented-out scikit-learn default (optimiser_presets.py:17-22);
    'gpy' -- the model of the preset's live line ``tm.GPySurrogate()`` (optimiser_presets.py:16; parity unpinned, GPy
    is not installed).  close_tolerance >= 0 lets the device argmax skip candidates the Fallback would reject
    (fallback.py:49-54); pass the Fallback's own tolerance (1e-10 by default there).
    candidates / seed: the random stage of the auxiliary optimiser (``GpuCandidateArgmax``): 'uniform' / 'lhs' generated
    on the device, 'numpy' drawn exactly as the reference's ``random_selector`` does.  model_params: passed to the
    surrogate (e.g. the demo's ``kernel=1.0 * Matern(nu=2.5) + WhiteKernel(), normalize_y=True``)."""
    if name != 'default':
        raise ValueError('unknown preset name: {}'.format(name))
    import turbo.modules as tm  # the host application
    optimiser.latent_space = tm.NoLatentSpace()
    optimiser.pre_phase_select = tm.LHS_selector(num_total=optimiser.pre_phase_trials)
    optimiser.fallback = tm.Fallback(selector=tm.random_selector())
    optimiser.aux_optimiser = GpuCandidateArgmax(num_random=num_random, candidates=candidates, seed=seed, group=group,
                                                 close_tolerance=close_tolerance)
    if surrogate == 'gpy':
        optimiser.surrogate = GpuGPySurrogate(model_params=model_params, training_iterations=training_iterations,
                                              seed=seed, group=group)
    elif surrogate == 'sklearn':
        optimiser.surrogate = GpuGPSurrogate(model_params=model_params, training_iterations=training_iterations,
                                             seed=seed, group=group)
    else:
        raise ValueError("surrogate must be 'sklearn' or 'gpy'")
    optimiser.acquisition = GpuEI(xi=0.01)
