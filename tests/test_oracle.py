"""The oracle (oracle/gp_oracle.py) against outputs of the reference itself
(tests/golden/*.npz, made by oracle/make_golden.py) and against scikit-learn live."""
import numpy as np
import pytest

from oracle import gp_oracle as O


def _spec(g):
    return O.KernelSpec(g["kind"], g["X"].shape[1], g["ard"], g["has_const"], g["has_white"])


def _fit(g):
    return O.FittedGP(_spec(g), g["theta"], g["X"], g["y"], normalize_y=bool(g["normalize_y"]))


def test_oracle_predict_matches_reference(golden):
    g = golden
    gp = _fit(g)
    mu, sigma = gp.predict(g["Xc"])
    np.testing.assert_allclose(mu, g["mu"], rtol=1e-10, atol=1e-10 * np.abs(g["y"]).max())
    # variance is k** - v.v: absolute floor tied to the prior variance
    floor = 64 * np.finfo(float).eps * g["X"].shape[0] * O.prior_diag(gp.spec, gp.theta) * gp.y_std ** 2
    np.testing.assert_allclose(sigma ** 2, g["sigma"] ** 2, rtol=1e-9, atol=floor)
    np.testing.assert_allclose(gp.alpha, g["alpha_"], rtol=1e-7, atol=1e-9 * np.abs(g["alpha_"]).max())
    np.testing.assert_allclose(np.diag(gp.L), g["L_diag"], rtol=1e-12)


def test_oracle_acquisition_matches_reference(golden):
    g = golden
    # from the reference's own mu/sigma: the acquisition formulas alone, bit-level
    for acq, key, par in ((O.ACQ_EI, "ei", g["xi"]), (O.ACQ_PI, "pi", g["xi"]), (O.ACQ_UCB, "ucb", g["beta"]),
                          (O.ACQ_UCB, "sig_only", float("inf"))):
        got = O.acquisition(acq, g["mu"], g["sigma"], g["sign"], g["incumbent"], par)
        np.testing.assert_array_equal(got, g[key])
    assert O.argmax_first(g["ei"]) == g["ei_argmax"] or g["ei"][O.argmax_first(g["ei"])] == g["ei"][g["ei_argmax"]]
    assert O.argmax_first(g["ucb"]) == g["ucb_argmax"]


def test_oracle_lml_matches_reference(golden):
    g = golden
    gp = _fit(g)
    assert gp.lml() == pytest.approx(g["lml"], rel=1e-10)
    if "lml_thetas" in g:
        for t, v, gr in zip(g["lml_thetas"], g["lml_vals"], g["lml_grads"]):
            lml, grad = gp.lml(t, eval_gradient=True)
            assert lml == pytest.approx(float(v), rel=1e-10)
            np.testing.assert_allclose(grad, gr, rtol=1e-7, atol=1e-9 * np.abs(gr).max())


def test_oracle_matches_sklearn_live():
    sk = pytest.importorskip("sklearn.gaussian_process")
    K = sk.kernels
    rng = np.random.default_rng(11)
    X, Xc = rng.random((80, 3)), rng.random((200, 3))
    y = np.sin(X.sum(1) * 3)
    kern = K.ConstantKernel(1.7) * K.Matern(length_scale=[0.3, 0.5, 0.9], nu=2.5) + K.WhiteKernel(1e-4)
    gp = sk.GaussianProcessRegressor(kernel=kern, optimizer=None, normalize_y=True).fit(X, y)
    o = O.FittedGP(O.KernelSpec(O.KIND_MATERN52, 3, True), gp.kernel_.theta, X, y)
    mu, sd = gp.predict(Xc, return_std=True)
    omu, osd = o.predict(Xc)
    np.testing.assert_allclose(omu, mu, rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(osd, sd, rtol=1e-8, atol=1e-10)
    lml, grad = gp.log_marginal_likelihood(gp.kernel_.theta, eval_gradient=True)
    olml, ograd = o.lml(eval_gradient=True)
    assert olml == pytest.approx(lml, rel=1e-11)
    np.testing.assert_allclose(ograd, grad, rtol=1e-8, atol=1e-10)


def test_oracle_lml_nonpd_is_minus_inf():
    X = np.zeros((4, 2))  # identical points, no noise -> singular K
    spec = O.KernelSpec(O.KIND_RBF, 2, False, has_white=False)
    lml, grad = O.log_marginal_likelihood(spec, np.zeros(2), X, np.ones(4), alpha_jitter=0.0, eval_gradient=True)
    assert lml == -np.inf and not grad.any()


def test_candidate_generators():
    rng = np.random.default_rng(0)
    U = O.uniform_candidates(rng, 100, [-5, 0], [10, 15])
    assert U.shape == (100, 2) and U[:, 0].min() >= -5 and U[:, 1].max() <= 15
    L = O.lhs_candidates(rng, 50, [0, 0, 0], [1, 2, 3])
    for k, hi in enumerate((1, 2, 3)):  # one point per stratum per dim
        assert sorted(np.floor(L[:, k] / hi * 50).astype(int)) == list(range(50))
