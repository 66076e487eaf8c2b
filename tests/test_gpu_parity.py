"""GPU parity tests: the CUDA path (through the C ABI) against the oracle and the golden vectors
generated from the reference.  Tolerances (BASELINE.json north_star): mean / variance / acquisition
1e-8 relative, LML 1e-6 relative, argmax index and selected point exact."""
import math

import numpy as np
import pytest

from oracle import gp_oracle as O

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
EPS = np.finfo(float).eps


@pytest.fixture(scope="module", autouse=True)
def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def _tb():
    import turbo_b200
    return turbo_b200


def _model(g, theta=None):
    tb = _tb()
    d = g["X"].shape[1]
    spec = tb.KernelSpec(g["kind"], d, g["ard"], g["has_const"], g["has_white"])
    theta = g["theta"] if theta is None else theta
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=bool(g["normalize_y"])),
                            training_iterations=0)
    model, info = sur.construct_model(0, g["X"], g["y"])
    assert info["iterations"] == 0
    return model


def _var_floor(g, model):
    # var = k** - v.v cancels near training points: absolute floor tied to the prior variance
    prior = np.exp(g["theta"][0]) if g["has_const"] else 1.0
    return 64 * EPS * g["X"].shape[0] * prior * model.y_std ** 2  # conftest.VAR_FLOOR_C


def test_dgemm_all_layouts():
    from turbo_b200 import device
    gen = torch.Generator(device="cuda").manual_seed(0)
    M, N, K = 256, 384, 144
    for ta in (0, 1):
        for tb_ in (0, 1):
            A = torch.randn((K, M) if ta else (M, K), dtype=torch.float64, device="cuda", generator=gen)
            B = torch.randn((N, K) if tb_ else (K, N), dtype=torch.float64, device="cuda", generator=gen)
            C0 = torch.randn((M, N), dtype=torch.float64, device="cuda", generator=gen)
            ref = 1.5 * (A.T if ta else A) @ (B.T if tb_ else B) - 0.5 * C0
            C = device.dgemm(ta, tb_, 1.5, A, B, -0.5, C0.clone())
            torch.cuda.synchronize()
            err = (C - ref).abs().max().item()
            assert err < 1e-11, (ta, tb_, err)


def test_factor_matches_reference(golden):
    g = golden
    model = _model(g)
    f = model.factor
    n = g["X"].shape[0]
    np.testing.assert_allclose(f.Ldiag.cpu().numpy()[:n], g["L_diag"], rtol=1e-9)
    assert np.all(f.Ldiag.cpu().numpy()[n:] == 1.0)
    assert model.get_log_likelihood() == pytest.approx(g["lml"], rel=1e-6)
    alpha = f.alpha.cpu().numpy()
    assert np.all(alpha[n:] == 0.0)
    np.testing.assert_allclose(alpha[:n], g["alpha_"], rtol=1e-5, atol=1e-7 * np.abs(g["alpha_"]).max())
    # Linv * K * Linv^T = I
    Ko = O.gram_train(O.KernelSpec(g["kind"], g["X"].shape[1], g["ard"], g["has_const"], g["has_white"]),
                      g["theta"], g["X"])
    Li = f.Linv.cpu().numpy()[:n, :n]
    assert np.abs(np.triu(Li, 1)).max() == 0.0
    resid = Li @ Ko @ Li.T - np.eye(n)
    assert np.abs(resid).max() < 1e-7
    assert list(model.get_hyper_param_names()) == [str(s) for s in g["hyper_names"]]
    np.testing.assert_allclose(model.get_hyper_params(), g["hyper_params"], rtol=1e-12)


def test_predict_matches_reference(golden):
    g = golden
    model = _model(g)
    mu, sigma = model.predict(g["Xc"], return_std_dev=True)
    assert mu.shape == sigma.shape == (g["Xc"].shape[0],)
    np.testing.assert_allclose(mu, g["mu"], rtol=1e-8, atol=1e-8 * np.abs(g["y"]).max())
    from conftest import check_variance
    prior = (np.exp(g["theta"][0]) if g["has_const"] else 1.0) * model.y_std ** 2
    check_variance("golden/%s/auto" % g["name"], sigma ** 2, g["sigma"] ** 2, g["X"].shape[0], prior)
    mu_only = model.predict(g["Xc"][:5])
    np.testing.assert_array_equal(mu_only, mu[:5])


def test_acquisition_matches_reference(golden):
    g = golden
    tb = _tb()
    model = _model(g)
    ext, inc = g["extremum"], g["incumbent"]
    fns = {
        "ei": tb.GpuEI(g["xi"]).construct_function(0, model, ext, inc)[0],
        "pi": tb.GpuPI(g["xi"]).construct_function(0, model, ext, inc)[0],
        "ucb": tb.GpuUCB(g["beta"]).construct_function(0, model, ext)[0],
        "sig_only": tb.GpuUCB(float("inf")).construct_function(0, model, ext)[0],
    }
    assert fns["ei"].get_name() == "EI" and fns["pi"].get_name() == "PI"
    assert fns["ucb"].get_name() == ("UCB" if ext == "max" else "-LCB")
    sd_floor = math.sqrt(_var_floor(g, model))
    for key, fn in fns.items():
        got = fn(g["Xc"])
        ref = g[key]
        # the acquisition value inherits sigma's cancellation floor where sigma ~ 0 (duplicates of training points)
        np.testing.assert_allclose(got, ref, rtol=1e-8, atol=max(1e-8 * np.abs(ref).max(), 4 * sd_floor), err_msg=key)
        val, idx, x = fn.argmax_over(candidates=g["Xc"])
        ref_idx = O.argmax_first(ref)
        if key + "_argmax" in g:
            assert ref[ref_idx] == ref[g[key + "_argmax"]]
        # exact index unless the reference's top two are closer than the parity tolerance
        top2 = np.sort(ref)[-2:]
        if top2[1] - top2[0] > 1e-7 * max(abs(top2[1]), 1e-300):
            assert idx == ref_idx, key
            np.testing.assert_array_equal(x, g["Xc"][ref_idx:ref_idx + 1])
        assert val == pytest.approx(got[idx], rel=0, abs=0)
        assert val == got.max()


def test_lml_and_gradient_match_reference(golden):
    g = golden
    if "lml_thetas" not in g:
        pytest.skip("no gradient vectors in this fixture")
    model = _model(g)
    for t, v, gr in zip(g["lml_thetas"], g["lml_vals"], g["lml_grads"]):
        lml, grad = model.log_marginal_likelihood(t, eval_gradient=True)
        assert lml == pytest.approx(float(v), rel=1e-6)
        np.testing.assert_allclose(grad, gr, rtol=1e-6, atol=1e-6 * np.abs(gr).max())


def test_lml_batched_members_are_independent(golden):
    g = golden
    if "lml_thetas" not in g or g["X"].shape[0] > 400:
        pytest.skip("small gradient fixtures only")
    from turbo_b200 import device
    model = _model(g)
    rng = np.random.default_rng(3)
    T = np.vstack([g["lml_thetas"], g["theta"] + rng.normal(0, 0.2, (5, g["theta"].size))])
    X_dev, y_dev = device.to_device(model.X), device.to_device(model.y_n)
    lml, grad, info = device.lml_batched(X_dev, y_dev, model.X.shape[0], model.spec, T, model.jitter, True)
    assert not info.any()
    for r in range(T.shape[0]):
        l1, g1, _ = device.lml_batched(X_dev, y_dev, model.X.shape[0], model.spec, T[r:r + 1], model.jitter, True)
        assert l1[0] == lml[r] and np.array_equal(g1[0], grad[r])  # bit-identical in or out of a batch
    ospec = O.KernelSpec(g["kind"], g["X"].shape[1], g["ard"], g["has_const"], g["has_white"])
    for r in range(T.shape[0]):
        ol, og = O.log_marginal_likelihood(ospec, T[r], model.X, model.y_n, model.jitter, True)
        assert lml[r] == pytest.approx(ol, rel=1e-6)
        np.testing.assert_allclose(grad[r], og, rtol=1e-6, atol=1e-6 * np.abs(og).max())


def test_left_looking_schedule_matches_right_looking_and_oracle():
    """many restarts switch potrf/trtri to the left-looking (long-K) schedule (fit.cu use_left_looking):
    same numbers as the right-looking schedule used for a single matrix, and as the oracle"""
    from turbo_b200 import device
    tb = _tb()
    rng = np.random.default_rng(11)
    n, d, R = 520, 4, 120            # NB = 5 -> R*NB = 600 >= 4*148: left-looking
    X = rng.random((n, d))
    y = np.sin(3 * X).sum(axis=1) + 0.05 * rng.standard_normal(n)
    y = (y - y.mean()) / y.std()
    spec = tb.KernelSpec(3, d, True)
    base = np.log(np.r_[1.0, np.full(d, 0.4), 1e-3])
    T = base + rng.normal(0, 0.3, (R, base.size))
    X_dev, y_dev = device.to_device(X), device.to_device(y)
    lml, grad, info = device.lml_batched(X_dev, y_dev, n, spec, T, 1e-10, True)
    assert not info.any()
    ospec = O.KernelSpec(3, d, True)
    for r in (0, 7, R - 1):
        l1, g1, _ = device.lml_batched(X_dev, y_dev, n, spec, T[r:r + 1], 1e-10, True)   # right-looking
        assert lml[r] == pytest.approx(l1[0], rel=1e-10)
        np.testing.assert_allclose(grad[r], g1[0], rtol=1e-8, atol=1e-8 * np.abs(g1[0]).max())
        ol, og = O.log_marginal_likelihood(ospec, T[r], X, y, 1e-10, True)
        assert lml[r] == pytest.approx(ol, rel=1e-6)
        np.testing.assert_allclose(grad[r], og, rtol=1e-6, atol=1e-6 * np.abs(og).max())
    # a non-PD member inside a left-looking batch is flagged and does not disturb the others
    T2 = T.copy()
    T2[5, -1] = -80.0                # noise ~ 0
    T2[5, 1:-1] = 6.0                # huge length scales -> numerically singular with jitter 0
    lml2, grad2, info2 = device.lml_batched(X_dev, y_dev, n, spec, T2, 0.0, True)
    assert info2[5] > 0 and lml2[5] == -np.inf and not grad2[5].any()
    keep = np.arange(R) != 5
    lml0, _, info0 = device.lml_batched(X_dev, y_dev, n, spec, T, 0.0, True)
    assert not info2[keep].any() and np.array_equal(lml2[keep], lml0[keep])


def test_non_positive_definite_member_is_flagged():
    from turbo_b200 import device
    tb = _tb()
    X = np.zeros((6, 2))
    X[3:] = 1.0  # two clusters of identical points, no noise, no jitter -> singular
    y = np.arange(6.0)
    spec = tb.KernelSpec(0, 2, False, has_const=True, has_white=False)
    good = tb.KernelSpec(0, 2, False, has_const=True, has_white=True)
    lml, grad, info = device.lml_batched(device.to_device(X), device.to_device(y), 6, spec, np.zeros((2, 2)), 0.0, True)
    assert (info > 0).all() and np.all(lml == -np.inf) and not grad.any()
    lml, grad, info = device.lml_batched(device.to_device(X), device.to_device(y), 6, good, np.zeros((1, 3)), 0.0, True)
    assert info[0] == 0 and np.isfinite(lml[0])
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, np.zeros(2)), alpha=0.0), training_iterations=0)
    with pytest.raises(np.linalg.LinAlgError):
        sur.construct_model(0, X, y)
