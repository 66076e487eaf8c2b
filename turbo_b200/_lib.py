"""ctypes binding of libturbo_b200.so (include/turbo_b200.h).

There is NO fallback: if the shared library is missing or a call fails this module raises.
PyTorch is used only as the device-buffer / stream layer; every number is produced by the
hand-written sm_100a kernels behind the C ABI.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TB200_LIB") or os.path.join(_HERE, "libturbo_b200.so")  # TB200_LIB: A/B builds
MAX_D = 64
I8_MAX_N = 32768  # TB200_I8_MAX_N (include/turbo_b200.h)

KIND_RBF, KIND_MATERN12, KIND_MATERN32, KIND_MATERN52 = 0, 1, 2, 3
ACQ_EI, ACQ_PI, ACQ_UCB = 0, 1, 2
CAND_GIVEN, CAND_UNIFORM, CAND_LHS = 0, 1, 2


class KernelSpecC(C.Structure):
    _fields_ = [("kind", C.c_int), ("d", C.c_int), ("ard", C.c_int), ("has_const", C.c_int), ("has_white", C.c_int)]


class AcqSpecC(C.Structure):
    _fields_ = [("acq", C.c_int), ("sign", C.c_double), ("incumbent", C.c_double), ("param", C.c_double),
                ("y_mean", C.c_double), ("y_std", C.c_double)]


class CandSpecC(C.Structure):
    _fields_ = [("mode", C.c_int), ("seed", C.c_uint64), ("m_total", C.c_int64),
                ("lo", C.c_double * MAX_D), ("hi", C.c_double * MAX_D)]


class ExclusionC(C.Structure):
    _fields_ = [("X", C.c_void_p), ("n", C.c_int), ("tol", C.c_double), ("min_out", C.c_void_p)]


_P = C.c_void_p
_SIGNATURES = {
    # name: (restype, argtypes)  -- one entry per declaration in include/turbo_b200.h
    "tb200_version": (C.c_int, []),
    "tb200_last_error": (C.c_char_p, []),
    "tb200_padded_n": (C.c_int, [C.c_int]),
    "tb200_shutdown": (C.c_int, []),
    "tb200_debug_reload": (C.c_int, []),
    "tb200_fit_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "tb200_acquire_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int64]),
    "tb200_default_batch": (C.c_int64, [C.c_int]),
    "tb200_lml_batched": (C.c_int, [_P, _P, C.c_int, C.POINTER(KernelSpecC), _P, C.c_int, C.c_double, C.c_int,
                                    _P, _P, _P, _P, C.c_size_t, _P]),
    "tb200_factor": (C.c_int, [_P, _P, C.c_int, C.POINTER(KernelSpecC), _P, C.c_double, _P, _P, _P, _P, _P, _P,
                               _P, C.c_size_t, _P]),
    "tb200_acquire": (C.c_int, [_P, C.c_int64, C.c_int64, C.POINTER(CandSpecC), _P, _P, _P, C.c_int,
                                C.POINTER(KernelSpecC), _P, C.POINTER(AcqSpecC), _P, _P, _P, _P, _P, C.c_int64,
                                _P, C.c_size_t, _P]),
    "tb200_best_reset": (C.c_int, [_P, _P, _P]),
    "tb200_i8_factor_bytes": (C.c_size_t, [C.c_int]),
    "tb200_i8_prepare": (C.c_int, [_P, C.c_int, C.POINTER(KernelSpecC), _P, _P, _P, _P]),
    "tb200_acquire_i8_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int64]),
    "tb200_acquire_i8": (C.c_int, [_P, C.c_int64, C.c_int64, C.POINTER(CandSpecC), _P, _P, _P, C.c_int, _P, C.c_int,
                                   C.POINTER(KernelSpecC), _P, C.POINTER(AcqSpecC), _P, _P, _P, _P, _P, C.c_int64,
                                   _P, C.c_size_t, _P]),
    "tb200_acquire_excl": (C.c_int, [C.POINTER(ExclusionC), _P, C.c_int64, C.c_int64, C.POINTER(CandSpecC), _P, _P, _P,
                                     C.c_int, C.POINTER(KernelSpecC), _P, C.POINTER(AcqSpecC), _P, _P, _P, _P, _P,
                                     C.c_int64, _P, C.c_size_t, _P]),
    "tb200_acquire_i8_excl": (C.c_int, [C.POINTER(ExclusionC), _P, C.c_int64, C.c_int64, C.POINTER(CandSpecC), _P, _P,
                                        _P, C.c_int, _P, C.c_int, C.POINTER(KernelSpecC), _P, C.POINTER(AcqSpecC), _P,
                                        _P, _P, _P, _P, C.c_int64, _P, C.c_size_t, _P]),
    "tb200_min_sqdist": (C.c_int, [_P, C.c_int64, _P, C.c_int, C.c_int, _P, _P, _P]),
    "tb200_i8_probe_points": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P]),
    "tb200_max_sqdiff": (C.c_int, [_P, _P, C.c_int64, _P, _P]),
    "tb200_generate_candidates": (C.c_int, [C.POINTER(CandSpecC), C.c_int, C.c_int64, C.c_int64, _P, _P]),
    "tb200_fp64_peak": (C.c_int, [C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "tb200_i8_peak": (C.c_int, [C.c_double, C.c_int, C.POINTER(C.c_double)]),
    "tb200_i8_mma_rate": (C.c_int, [C.c_double, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "tb200_i8_mma_list": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int),
                                    C.POINTER(C.c_int)]),
    "tb200_append_workspace_bytes": (C.c_size_t, [C.c_int]),
    "tb200_factor_append": (C.c_int, [_P, _P, _P, _P, C.c_int, C.POINTER(KernelSpecC), _P, C.c_double, _P, _P,
                                      C.c_size_t, _P]),
    "tb200_alpha_lml": (C.c_int, [_P, C.c_int, _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "tb200_alpha_lml_static": (C.c_int, [_P, C.c_int, _P, _P, _P, _P, _P]),
    "tb200_acquire_local_best_count": (C.c_int, [C.c_int64, C.c_int]),
    "tb200_acquire_local_bests": (C.c_int, [_P, C.c_int, C.c_int, C.c_int64, C.c_int, _P, _P, _P]),
    "tb200_acquire_grad_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "tb200_acquire_grad": (C.c_int, [_P, C.c_int, _P, _P, _P, C.c_int, C.POINTER(KernelSpecC), _P, C.POINTER(AcqSpecC),
                                     _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "tb200_timing_enable": (C.c_int, [C.c_int]),
    "tb200_trace_enable": (C.c_int, [C.c_int]),
    "tb200_trace_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_int]),
    "tb200_timing_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "tb200_dgemm": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, _P, C.c_int, _P, C.c_int,
                              C.c_double, _P, C.c_int, _P]),
    "tb200_launch_count": (C.c_int64, []),
}
EXPORTS = tuple(_SIGNATURES)

_lib = None


def load():
    """dlopen the in-tree library; raises (never falls back) when it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(turbo_b200 has no CPU fallback)" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the header and the library disagree
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(status):
    if status != 0:
        raise RuntimeError("turbo_b200: " + load().tb200_last_error().decode())


def ptr(t):
    """device pointer of a torch tensor (or None)"""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def cand_spec(mode, seed=0, m_total=0, lo=(), hi=()):
    c = CandSpecC()
    c.mode, c.seed, c.m_total = int(mode), int(seed) & (2 ** 64 - 1), int(m_total)
    for k, (a, b) in enumerate(zip(lo, hi)):
        c.lo[k], c.hi[k] = float(a), float(b)
    return c


# ---------------------------------------------------------------------------------------------
# host mirror of the device candidate generator (Philox4x32-10), vectorised NumPy.  Used to turn a
# winning global index back into its row without a device round trip, and by the tests.
# ---------------------------------------------------------------------------------------------
def _philox(c0, c1, c2, c3, k0, k1):
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    c0, c1, c2, c3 = (np.asarray(v, dtype=np.uint64) & np.uint64(0xFFFFFFFF) for v in (c0, c1, c2, c3))
    k0, k1 = np.uint64(k0), np.uint64(k1)
    lo32 = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        n0 = ((p1 >> np.uint64(32)) ^ c1 ^ k0) & lo32
        n1 = p1 & lo32
        n2 = ((p0 >> np.uint64(32)) ^ c3 ^ k1) & lo32
        n3 = p0 & lo32
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + np.uint64(0x9E3779B9)) & lo32
        k1 = (k1 + np.uint64(0xBB67AE85)) & lo32
    return c0, c1, c2, c3


def _uniform01(seed, idx, dim):
    idx = np.asarray(idx, dtype=np.uint64)
    o0, o1, _, _ = _philox(idx & np.uint64(0xFFFFFFFF), idx >> np.uint64(32), np.uint64(dim), np.uint64(0),
                           seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    bits = ((o0 << np.uint64(21)) ^ (o1 >> np.uint64(11))) & np.uint64((1 << 53) - 1)
    return bits.astype(np.float64) * (1.0 / 9007199254740992.0)


def _permute_index(seed, dim, idx, m):
    bits = 2
    while (1 << bits) < m:
        bits += 1
    bits += bits & 1
    half = bits // 2
    mask = np.uint64((1 << half) - 1)
    x = np.asarray(idx, dtype=np.uint64).copy()
    todo = np.ones(x.shape, dtype=bool)
    while todo.any():
        v = x[todo]
        l, r = v >> np.uint64(half), v & mask
        for rnd in range(4):
            o0, o1, _, _ = _philox(r & np.uint64(0xFFFFFFFF), r >> np.uint64(32), np.uint64(dim),
                                   np.uint64(0x5eed0000 + rnd), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
            f = ((o0 << np.uint64(32)) | o1) & mask
            l, r = r, l ^ f
        v = (l << np.uint64(half)) | r
        x[todo] = v
        todo[todo] = v >= np.uint64(m)
    return x


def host_candidates(mode, seed, m_total, lo, hi, indices):
    """rows of the generated candidate set at the given global indices (bit-identical to the device)."""
    with np.errstate(over="ignore"):
        idx = np.asarray(indices, dtype=np.uint64).reshape(-1)
        out = np.empty((idx.size, len(lo)))
        for k, (a, b) in enumerate(zip(lo, hi)):
            u = _uniform01(int(seed), idx, k)
            if mode == CAND_UNIFORM:
                out[:, k] = a + (b - a) * u
            else:
                s = _permute_index(int(seed), k, idx, int(m_total)).astype(np.float64)
                out[:, k] = a + (b - a) * (s + u) / float(m_total)
    return out
