"""tb200_alpha_lml_static: the allocation-free variant of tb200_alpha_lml (scratch in a file-scope device array of the
library; no counterpart in the reference).  Collected LAST (file name) so
that a failure here cannot hide another test under `pytest -x`.  Passes alone in a fresh process and after the other
tests: it builds its own model and shares nothing with them."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

STATIC_MAX_N = 16384  # include/turbo_b200.h: the limit of the built-in scratch


def _model(n, d, seed):
    import turbo_b200 as tb
    rng = np.random.default_rng(seed)
    X = rng.random((n, d))
    y = np.sin(3.0 * X.sum(axis=1)) + 0.1 * rng.standard_normal(n)
    spec = tb.KernelSpec(3, d, True)
    theta = np.log(np.r_[1.3, np.full(d, 0.6), 1e-3])
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0)
    model, _ = sur.construct_model(0, X, y)
    return model, X, spec, theta, rng


def _both(lib, f, yn_dev, n):
    """(alpha, lml) from the existing entry point (caller workspace) and from the allocation-free one"""
    from turbo_b200 import _lib, device
    dev = yn_dev.device
    out = []
    nbytes = lib.tb200_append_workspace_bytes(n)
    work = device._workspace(nbytes)
    for static in (False, True):
        alpha = torch.full((f.NP,), float("nan"), dtype=torch.float64, device=dev)
        lml = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
        if static:
            _lib.check(lib.tb200_alpha_lml_static(_lib.ptr(yn_dev), n, _lib.ptr(f.Linv), _lib.ptr(f.Ldiag),
                                                  _lib.ptr(alpha), _lib.ptr(lml), _lib.stream_ptr()))
        else:
            _lib.check(lib.tb200_alpha_lml(_lib.ptr(yn_dev), n, _lib.ptr(f.Linv), _lib.ptr(f.Ldiag), _lib.ptr(alpha),
                                           _lib.ptr(lml), _lib.ptr(work), nbytes, _lib.stream_ptr()))
        torch.cuda.synchronize()
        out.append((alpha.cpu().numpy()[:n].copy(), float(lml.item())))
    return out


def test_alpha_lml_static_matches_workspace_entry_point_two_calls_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from oracle import gp_oracle as O
    from turbo_b200 import _lib, device
    lib = _lib.load()
    n, d = 700, 5
    model, X, spec, theta, rng = _model(n, d, 11)
    f = model.factor
    # two calls in a row with DIFFERENT targets: the shared scratch keeps the first call's contents, the second call
    # must not see them
    for call in range(2):
        y = np.cos(2.0 * X[:, 0] + call) + 0.05 * rng.standard_normal(n)
        yn = (y - y.mean()) / y.std()
        yn_dev = device.to_device(yn)
        (a_ref, l_ref), (a_st, l_st) = _both(lib, f, yn_dev, n)
        assert np.isfinite(a_st).all() and np.isfinite(l_st)
        np.testing.assert_allclose(a_st, a_ref, rtol=1e-8, atol=1e-8 * np.abs(a_ref).max())
        assert l_st == pytest.approx(l_ref, rel=1e-6)
        # and against the oracle (the checker), as the existing entry point is tested
        og = O.FittedGP(O.KernelSpec(3, d, True), theta, X, y)
        assert l_st == pytest.approx(og.lml(), rel=1e-6)
    # the allocation-free call on its own changes nothing in the device's free memory
    alpha = torch.empty(f.NP, dtype=torch.float64, device=yn_dev.device)
    lml = torch.empty(1, dtype=torch.float64, device=yn_dev.device)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    _lib.check(lib.tb200_alpha_lml_static(_lib.ptr(yn_dev), n, _lib.ptr(f.Linv), _lib.ptr(f.Ldiag), _lib.ptr(alpha),
                                          _lib.ptr(lml), _lib.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before
    assert float(lml.item()) == pytest.approx(l_ref, rel=1e-6)


def test_alpha_lml_static_rejects_sizes_beyond_its_scratch():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from turbo_b200 import _lib
    lib = _lib.load()
    t = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = lib.tb200_alpha_lml_static(_lib.ptr(t), STATIC_MAX_N + 1, _lib.ptr(t), _lib.ptr(t), _lib.ptr(t), _lib.ptr(t),
                                    _lib.stream_ptr())
    assert rc != 0
    msg = lib.tb200_last_error().decode()
    assert "16384" in msg and "tb200_alpha_lml_static" in msg
    rc = lib.tb200_alpha_lml_static(_lib.ptr(t), 0, _lib.ptr(t), _lib.ptr(t), _lib.ptr(t), _lib.ptr(t),
                                    _lib.stream_ptr())
    assert rc != 0
