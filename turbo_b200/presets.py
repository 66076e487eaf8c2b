"""Optimiser presets that swap the B200 surrogate + acquisition in (turbo/optimiser_presets.py:8-37).

``Optimiser.__setattr__`` only lets the six module slots be assigned (turbo/optimiser.py:98-110);
this assigns exactly those, so the recorder and plotting see a normally configured optimiser."""
from .acquisition import GpuEI
from .aux_optimiser import GpuCandidateArgmax
from .surrogate import GpuGPSurrogate


def load_gpu_preset(optimiser, name='default', num_random=1 << 20, training_iterations=10, group=None):
    if name != 'default':
        raise ValueError('unknown preset name: {}'.format(name))
    import turbo.modules as tm  # the host application
    optimiser.latent_space = tm.NoLatentSpace()
    optimiser.pre_phase_select = tm.LHS_selector(num_total=optimiser.pre_phase_trials)
    optimiser.fallback = tm.Fallback(selector=tm.random_selector())
    optimiser.aux_optimiser = GpuCandidateArgmax(num_random=num_random, group=group)
    # the reference's commented-out scikit-learn default (optimiser_presets.py:17-22)
    optimiser.surrogate = GpuGPSurrogate(training_iterations=training_iterations, group=group)
    optimiser.acquisition = GpuEI(xi=0.01)
