"""turbo_b200: B200-native GP surrogate + acquisition plugins for mbway/turbo.

Host layer (Python, mirrors turbo.modules' plugin interface) over libturbo_b200.so (hand-written
sm_100a CUDA behind the C ABI in include/turbo_b200.h).  No CPU fallback anywhere.
"""
from . import _lib
from .kernels import KernelSpec, from_sklearn
from .surrogate import GpuGPSurrogate
from .gpy_surrogate import GpuGPySurrogate, gpy_kernel
from .acquisition import GpuEI, GpuPI, GpuUCB
from .aux_optimiser import GpuCandidateArgmax
from .presets import load_gpu_preset

__all__ = ['KernelSpec', 'from_sklearn', 'GpuGPSurrogate', 'GpuGPySurrogate', 'gpy_kernel', 'GpuEI', 'GpuPI', 'GpuUCB', 'GpuCandidateArgmax',
           'load_gpu_preset']
