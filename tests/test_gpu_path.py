"""GPU tests of the wider path: candidate generation, fused argmax, hyper-parameter fitting, the
Branin-Hoo demo (config 1) replayed against the reference's recorded run, full-size properties
(config 2-4 shapes) and pickling for the unmodified Recorder."""
import os
import pickle

import numpy as np
import pytest

from oracle import gp_oracle as O

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module", autouse=True)
def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def _fixed_model(spec, theta, X, y, normalize_y=True):
    import turbo_b200 as tb
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=normalize_y), training_iterations=0)
    return sur.construct_model(0, X, y)[0]


@pytest.mark.parametrize("mode", ["uniform", "lhs"])
def test_device_generator_matches_host_mirror(mode):
    from turbo_b200 import _lib, device
    lo, hi = [-5.0, 0.0, 2.0], [10.0, 15.0, 2.5]
    m = 1000
    code = {"uniform": _lib.CAND_UNIFORM, "lhs": _lib.CAND_LHS}[mode]
    c = _lib.cand_spec(code, 1234, m, lo, hi)
    full = device.generate_candidates(c, 3, 0, m).cpu().numpy()
    part = device.generate_candidates(c, 3, 400, 200).cpu().numpy()
    np.testing.assert_array_equal(part, full[400:600])  # keyed by the GLOBAL index
    np.testing.assert_array_equal(full, _lib.host_candidates(code, 1234, m, lo, hi, np.arange(m)))
    assert (full >= lo).all() and (full <= hi).all()
    if mode == "lhs":  # one point per stratum in every dimension (naive_selectors.py:75-80)
        for k in range(3):
            strata = np.floor((full[:, k] - lo[k]) / (hi[k] - lo[k]) * m).astype(int)
            assert sorted(strata) == list(range(m))


def test_fused_argmax_equals_evaluate_then_argmax():
    import turbo_b200 as tb
    from turbo_b200 import _lib
    rng = np.random.default_rng(5)
    X = rng.random((200, 4))
    y = np.sin(3 * X.sum(1))
    spec = tb.KernelSpec(3, 4, True)
    theta = np.log([1.2, 0.4, 0.5, 0.6, 0.7, 1e-4])
    model = _fixed_model(spec, theta, X, y)
    lo, hi = [0.0] * 4, [1.0] * 4
    m = 5000  # not a multiple of the batch or the tile
    for acq in (tb.GpuEI(0.01).construct_function(0, model, "min", float(y.min()))[0],
                tb.GpuUCB(2.0).construct_function(0, model, "max")[0]):
        for gen, code in (("uniform", _lib.CAND_UNIFORM), ("lhs", _lib.CAND_LHS)):
            val, idx, x = acq.argmax_over(generator=gen, m=m, lo=lo, hi=hi, seed=9, batch=1024)
            cands = _lib.host_candidates(code, 9, m, lo, hi, np.arange(m))
            vals = acq(cands)
            assert idx == int(np.argmax(vals)) and val == vals.max()
            np.testing.assert_array_equal(x, cands[idx:idx + 1])
            mu, sd = O.FittedGP(O.KernelSpec(3, 4, True), theta, X, y).predict(cands)
            code_acq, par = (O.ACQ_EI, 0.01) if acq.get_name() == "EI" else (O.ACQ_UCB, 2.0)
            ref = O.acquisition(code_acq, mu, sd, acq.scale_factor, float(y.min()), par)
            np.testing.assert_allclose(vals, ref, rtol=1e-8, atol=1e-8 * np.abs(ref).max())
            assert idx == O.argmax_first(ref)


def test_host_candidates_streamed_in_chunks_through_two_buffers():
    """argmax_over(candidates=host array) streams chunks through two reusable device buffers (Acquirer.run_host): ragged
    chunk sizes, more chunks than buffers, a pinned torch tensor or a NumPy array -- always the oracle's argmax."""
    import turbo_b200 as tb
    rng = np.random.default_rng(8)
    X = rng.random((150, 3))
    y = np.cos(4 * X.sum(1))
    spec = tb.KernelSpec(3, 3, True)
    theta = np.log([0.9, 0.3, 0.4, 0.5, 1e-4])
    model = _fixed_model(spec, theta, X, y)
    acq = tb.GpuEI(0.01).construct_function(0, model, "min", float(y.min()))[0]
    m = 3001
    cands = rng.random((m, 3))
    mu, sd = O.FittedGP(O.KernelSpec(3, 3, True), theta, X, y).predict(cands)
    ref = O.acquisition(O.ACQ_EI, mu, sd, acq.scale_factor, float(y.min()), 0.01)
    want = O.argmax_first(ref)
    pinned = torch.from_numpy(cands.copy()).pin_memory()
    for src in (cands, pinned):
        for chunk in (1 << 18, 1000, 777, 128, m):
            val, idx, x = acq.argmax_over(candidates=src, chunk=chunk, batch=256)
            assert idx == want, (chunk, idx, want)
            assert val == pytest.approx(ref[want], rel=1e-8)
            np.testing.assert_array_equal(x, cands[want:want + 1])
    # the top list of the polish stage is collected per chunk
    val, idx, x, (tv, ti, tx) = acq.argmax_over(candidates=cands, chunk=500, batch=256, keep_top=4)
    assert idx == want and ti[0] == want and len(set(ti.tolist())) == len(ti)
    np.testing.assert_array_equal(tx, cands[ti])


def test_argmax_edge_cases():
    import turbo_b200 as tb
    X = np.array([[0.1], [0.5], [0.9]])
    y = np.array([1.0, 0.0, 2.0])
    model = _fixed_model(tb.KernelSpec(3, 1, False), np.log([1.0, 0.3, 1e-6]), X, y)
    acq = tb.GpuUCB(0.0).construct_function(0, model, "max")[0]
    # ties: identical candidates -> lowest index wins
    C = np.array([[0.3], [0.7], [0.7], [0.3]])
    val, idx, x = acq.argmax_over(candidates=C)
    vals = acq(C)
    assert vals[1] == vals[2] and idx == int(np.argmax(vals))
    # a single candidate, and an empty set
    assert acq.argmax_over(candidates=C[:1])[1] == 0
    assert acq.argmax_over(candidates=np.zeros((0, 1))) == (-np.inf, -1, None)
    assert acq(np.zeros((0, 1))).shape == (0,)
    # NaN candidates never win
    C2 = np.array([[np.nan], [0.2], [np.nan]])
    assert acq.argmax_over(candidates=C2)[1] == 1
    aux = tb.GpuCandidateArgmax(num_random=0)
    assert aux.num_random == 0


def test_fit_reaches_sklearn_optimum():
    sk = pytest.importorskip("sklearn.gaussian_process")
    import turbo_b200 as tb
    K = sk.kernels
    rng = np.random.default_rng(0)
    X = rng.random((120, 2))
    y = np.sin(5 * X[:, 0]) + np.cos(3 * X[:, 1]) + 0.05 * rng.standard_normal(120)
    kern = 1.0 * K.Matern(length_scale=[1.0, 1.0], nu=2.5) + K.WhiteKernel()
    ref = sk.GaussianProcessRegressor(kernel=kern, normalize_y=True, n_restarts_optimizer=4, random_state=0).fit(X, y)
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=kern, normalize_y=True), training_iterations=5, seed=0)
    model, info = sur.construct_model(10, X, y)
    assert info["iterations"] == 5 and info["restarts"] == 5
    # same optimum up to optimiser tolerance (trajectories differ)
    assert model.get_log_likelihood() >= ref.log_marginal_likelihood_value_ - 1e-3 * abs(ref.log_marginal_likelihood_value_)
    # and the LML we report for our theta is the reference's LML at that theta
    assert model.get_log_likelihood() == pytest.approx(ref.log_marginal_likelihood(model.theta), rel=1e-6)
    # warm start: next trial begins from the fitted hyper-parameters (surrogates.py:302-304)
    np.testing.assert_allclose(sur._last_model_params, np.exp(model.theta))
    model2, info2 = sur.construct_model(11, X, y)
    assert model2.get_log_likelihood() >= model.get_log_likelihood() - 1e-6 * abs(model.get_log_likelihood())


def test_branin_demo_replay_matches_reference_run():
    """config 1: every Bayesian trial of the reference's recorded Branin-Hoo run, replayed on the device
    at the reference's fitted theta with the reference's candidates -> same chosen point."""
    import turbo_b200 as tb
    g = np.load(os.path.join(GOLDEN, "branin_run.npz"))
    xs, ys, cands, thetas = g["xs"], g["ys"], g["cands"], g["thetas"]
    spec = tb.KernelSpec(3, 2, False)
    agree = 0
    for k, t in enumerate(g["bayes_trials"]):
        model = _fixed_model(spec, thetas[k], xs[:t], ys[:t])
        assert model.get_log_likelihood() == pytest.approx(float(g["lmls"][k]), rel=1e-6)
        acq = tb.GpuEI(0.01).construct_function(int(t), model, "min", float(ys[:t].min()))[0]
        val, idx, x = acq.argmax_over(candidates=cands[k])
        assert val == pytest.approx(float(g["max_acq"][k]), rel=1e-7, abs=1e-300)
        if np.array_equal(x[0], xs[t]):  # xs[t] differs only when the reference's fallback replaced the point
            agree += 1
        else:
            assert not any(np.array_equal(xs[t], c) for c in cands[k])
    assert agree >= len(g["bayes_trials"]) - 5


def test_model_pickles_without_device_state():
    import turbo_b200 as tb
    rng = np.random.default_rng(1)
    X, y = rng.random((50, 3)), rng.random(50)
    model = _fixed_model(tb.KernelSpec(0, 3, True), np.log([1.0, 0.5, 0.5, 0.5, 1e-3]), X, y)
    Xc = rng.random((10, 3))
    mu, sd = model.predict(Xc, return_std_dev=True)
    blob = pickle.dumps(model)
    assert len(blob) < 20000  # no O(n^2) state inside
    try:
        import dill
        blob = dill.dumps(model)
    except ImportError:
        pass
    clone = pickle.loads(blob)
    assert clone._factor is None
    mu2, sd2 = clone.predict(Xc, return_std_dev=True)  # rebuilt lazily on first use
    np.testing.assert_array_equal(mu, mu2)
    np.testing.assert_array_equal(sd, sd2)
    model.release()
    np.testing.assert_array_equal(model.predict(Xc), mu)


@pytest.mark.parametrize("n,d,kind", [(1024, 6, 3), (4096, 10, 3), (2000, 20, 0)])
def test_full_size_properties(n, d, kind):
    """BASELINE config shapes: size-independent properties + oracle comparison on a candidate sample."""
    import turbo_b200 as tb
    rng = np.random.default_rng(n)
    X = rng.random((n, d))
    y = np.sin(2 * np.pi * X).sum(1) + 0.01 * rng.standard_normal(n)
    spec = tb.KernelSpec(kind, d, True)
    theta = np.log(np.r_[1.0, np.full(d, 0.5 if kind == 3 else 1.5), 1e-4 if kind == 3 else 1e-2])
    model = _fixed_model(spec, theta, X, y)
    f = model.factor
    # (1) Linv really is the inverse factor: Linv K Linv^T = I on a random probe
    Ko = O.gram_train(O.KernelSpec(kind, d, True), theta, X)
    Li = f.Linv.cpu().numpy()[:n, :n]
    v = rng.standard_normal(n)
    w = Li.T @ (Li @ (Ko @ v))
    np.testing.assert_allclose(w, v, rtol=0, atol=1e-7 * np.abs(v).max())
    # (2) exact identity at training points: K (K + s I)^-1 y = y - s alpha  (s = noise + jitter)
    mu, sd = model.predict(X[:256], return_std_dev=True)
    s_tot = np.exp(theta[-1]) + model.jitter
    expect = model.y_mean + model.y_std * (model.y_n[:256] - s_tot * f.alpha.cpu().numpy()[:256])
    np.testing.assert_allclose(mu, expect, rtol=0, atol=1e-7 * np.abs(y).max())
    assert np.isfinite(sd).all() and (sd >= 0).all()
    # (3) oracle parity on a sample
    Xc = rng.random((512, d))
    og = O.FittedGP(O.KernelSpec(kind, d, True), theta, X, y)
    omu, osd = og.predict(Xc)
    mu, sd = model.predict(Xc, return_std_dev=True)
    np.testing.assert_allclose(mu, omu, rtol=1e-8, atol=1e-8 * np.abs(y).max())
    from conftest import check_variance
    check_variance("full_size/n%d_d%d_kind%d" % (n, d, kind), sd ** 2, osd ** 2, n,
                   (1.0 + np.exp(theta[-1])) * model.y_std ** 2)
    assert model.get_log_likelihood() == pytest.approx(og.lml(), rel=1e-6)
    # (4) sharding the candidate index range does not change the winner (what multi-GPU relies on)
    from turbo_b200 import _lib, device, dist
    acq = tb.GpuEI(0.01).construct_function(0, model, "min", float(y.min()))[0]
    m = 3000
    val, idx, x = acq.argmax_over(generator="lhs", m=m, lo=[0.0] * d, hi=[1.0] * d, seed=3)
    pairs = []
    for r in range(4):
        a = device.Acquirer(model.factor, acq._spec_c())
        lo_i, hi_i = dist.shard_range(m, r, 4)
        a.run(None, hi_i - lo_i, lo_i, _lib.cand_spec(_lib.CAND_LHS, 3, m, [0.0] * d, [1.0] * d))
        pairs.append(a.best())
    assert dist.reduce_argmax_pairs(*zip(*pairs)) == (val, idx)


@pytest.mark.parametrize("kind,acq_name", [(3, "EI"), (0, "PI"), (2, "UCB"), (3, "LCB")])
def test_acquisition_gradient_matches_oracle_finite_differences(kind, acq_name):
    """tb200_acquire_grad: the value equals the fused path's, the analytic gradient equals central
    finite differences of the ORACLE's acquisition (sklearn predict + acquisition_functions.py formulas)"""
    import turbo_b200 as tb
    rng = np.random.default_rng(5)
    n, d, P = 300, 4, 9
    X = rng.random((n, d))
    y = np.sin(3 * X).sum(axis=1) + 0.05 * rng.standard_normal(n)
    theta = np.log(np.r_[1.3, 0.3 + 0.3 * rng.random(d), 1e-3])
    spec = tb.KernelSpec(kind, d, True)
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0)
    model = sur.construct_model(0, X, y)[0]
    og = O.FittedGP(O.KernelSpec(kind, d, True), theta, X, y)
    inc = float(np.median(y))  # an incumbent in the bulk of the data keeps PI / EI away from the 1e-100 tail
    if acq_name == "EI":
        acq = tb.GpuEI(0.01).construct_function(0, model, 'min', inc)[0]
        code, sign, param = O.ACQ_EI, -1.0, 0.01
    elif acq_name == "PI":
        acq = tb.GpuPI(0.01).construct_function(0, model, 'min', inc)[0]
        code, sign, param = O.ACQ_PI, -1.0, 0.01
    else:
        ext = 'max' if acq_name == "UCB" else 'min'
        acq = tb.GpuUCB(2.0).construct_function(0, model, ext)[0]
        code, sign, param = O.ACQ_UCB, (1.0 if ext == 'max' else -1.0), 2.0

    def oracle_acq(Z):
        mu, sd = og.predict(Z)
        return O.acquisition(code, mu, sd, sign, inc, param)

    Xp = rng.random((P, d))
    val, grad = acq.value_and_grad(Xp)
    np.testing.assert_allclose(val, acq(Xp), rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(val, oracle_acq(Xp), rtol=1e-8, atol=1e-10)
    h = 1e-5
    fd = np.zeros_like(grad)
    for k in range(d):
        e = np.zeros(d); e[k] = h
        fd[:, k] = (oracle_acq(Xp + e) - oracle_acq(Xp - e)) / (2 * h)
    np.testing.assert_allclose(grad, fd, rtol=2e-5, atol=2e-6 * np.abs(fd).max())


def test_gradient_polish_improves_on_random_stage():
    """GpuCandidateArgmax(grad_restarts>0): the polished point is inside the box and its acquisition value (re-evaluated
    through the plain __call__ path) is >= the random-stage maximum it started from"""
    import turbo_b200 as tb

    class B:  # the two attributes of turbo's Bounds the aux-optimiser reads
        def __init__(self, lo, hi):
            self.ordered = [("x%d" % i, a, b) for i, (a, b) in enumerate(zip(lo, hi))]

    rng = np.random.default_rng(9)
    n, d = 200, 3
    X = rng.random((n, d))
    y = ((X - 0.3) ** 2).sum(axis=1) + 0.01 * rng.standard_normal(n)
    spec = tb.KernelSpec(3, d, True)
    theta = np.log(np.r_[1.0, np.full(d, 0.4), 1e-4])
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0)
    model = sur.construct_model(0, X, y)[0]
    acq = tb.GpuEI(0.01).construct_function(0, model, 'min', float(y.min()))[0]
    aux = tb.GpuCandidateArgmax(num_random=4096, seed=3, grad_restarts=6, start_from_best=3, release_model=False)
    x, info = aux(B([0.0] * d, [1.0] * d), acq)
    assert x.shape == (1, d) and (x >= 0).all() and (x <= 1).all()
    assert info['max_acq'] >= info['random_stage_max_acq']
    assert acq(x)[0] == pytest.approx(info['max_acq'], rel=1e-8, abs=1e-12)
    assert info['polish_starts'] == 6 and info['polish_evaluations'] >= 1


def test_incremental_refit_equals_full_refactorisation():
    """trials that append rows at fixed hyper-parameters extend the previous factor (tb200_factor_append);
    predictions, LML and the selected candidate equal those of a model factorised from scratch"""
    import turbo_b200 as tb
    rng = np.random.default_rng(21)
    n0, d = 250, 5                     # NP = 256: six appends fit the padding, the seventh needs a rebuild
    X = rng.random((n0 + 8, d))
    y = np.cos(2 * X).sum(axis=1) + 0.02 * rng.standard_normal(n0 + 8)
    spec = tb.KernelSpec(3, d, True)
    theta = np.log(np.r_[0.8, np.full(d, 0.45), 1e-3])
    mp = dict(kernel=(spec, theta), normalize_y=True)
    inc = tb.GpuGPSurrogate(model_params=mp, training_iterations=0)
    Xc = rng.random((3000, d))
    prev = None
    for n in (n0, n0 + 1, n0 + 4, n0 + 6, n0 + 8):
        model, info = inc.construct_model(0, X[:n], y[:n])
        expect_inc = prev is not None and n <= 256
        assert bool(info.get('incremental', False)) == expect_inc, (n, info)
        if prev is not None and expect_inc:
            assert prev._factor is None  # the factor moved to the new model
        fresh = tb.GpuGPSurrogate(model_params=mp, training_iterations=0, incremental=False)
        ref, _ = fresh.construct_model(0, X[:n], y[:n])
        mu, sd = model.predict(Xc, return_std_dev=True)
        mu0, sd0 = ref.predict(Xc, return_std_dev=True)
        np.testing.assert_allclose(mu, mu0, rtol=1e-9, atol=1e-10)
        np.testing.assert_allclose(sd ** 2, sd0 ** 2, rtol=1e-8, atol=64 * np.finfo(float).eps * n * model.y_std ** 2)
        assert model.get_log_likelihood() == pytest.approx(ref.get_log_likelihood(), rel=1e-10)
        a1 = tb.GpuEI(0.01).construct_function(0, model, 'min', float(y[:n].min()))[0]
        a0 = tb.GpuEI(0.01).construct_function(0, ref, 'min', float(y[:n].min()))[0]
        assert a1.argmax_over(candidates=Xc)[1] == a0.argmax_over(candidates=Xc)[1]
        prev = model
    # the oracle agrees with an incrementally built model as well
    sur = tb.GpuGPSurrogate(model_params=mp, training_iterations=0)
    sur.construct_model(0, X[:n0], y[:n0])
    model, info = sur.construct_model(1, X[:n0 + 5], y[:n0 + 5])
    assert info['incremental']
    og = O.FittedGP(O.KernelSpec(3, d, True), theta, X[:n0 + 5], y[:n0 + 5])
    mu_o, sd_o = og.predict(Xc[:500])
    mu, sd = model.predict(Xc[:500], return_std_dev=True)
    np.testing.assert_allclose(mu, mu_o, rtol=1e-8, atol=1e-8 * np.abs(mu_o).max())
    np.testing.assert_allclose(sd ** 2, sd_o ** 2, rtol=1e-8, atol=64 * np.finfo(float).eps * (n0 + 5) * model.y_std ** 2)
    assert model.get_log_likelihood() == pytest.approx(og.lml(), rel=1e-8)


# ---------------------------------------------------------------------------------------------
# the fallback's "too close to an existing sample" test over the whole candidate set
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("d", [1, 2, 5, 7, 8, 9, 10, 16, 17, 20, 33, 64])
def test_min_sqdist_is_bit_exact_with_numpy(d):
    from turbo_b200 import device
    rng = np.random.default_rng(d)
    for n, m in ((1, 1), (63, 127), (64, 128), (65, 129), (300, 1000)):
        X = rng.random((n, d)) * rng.choice([1e-3, 1.0, 1e3], size=(1, d))
        Xq = rng.random((m, d)) * rng.choice([1e-3, 1.0, 1e3], size=(1, d))
        Xq[::7] = X[rng.integers(0, n, size=len(Xq[::7]))]          # exact duplicates: distance 0
        Xq[1::7] = X[rng.integers(0, n, size=len(Xq[1::7]))] + 1e-6  # near duplicates around the tolerance
        got, arg = device.min_sqdist(Xq, X, want_arg=True)
        want, warg = O.min_sqdist(Xq, X)
        np.testing.assert_array_equal(got, want)   # bit for bit: the comparison with tol must agree
        np.testing.assert_array_equal(arg, warg)   # first of the nearest rows
        np.testing.assert_array_equal(device.min_sqdist(Xq, X), want)  # the split-over-points schedule (no arg)
        for tol in (0.0, 1e-10, 1e-12 * d, 1e-5):
            close = np.array([O.close_to_any(Xq[i:i + 1], X, tol) for i in range(m)])
            np.testing.assert_array_equal(got <= tol, close)


def test_min_sqdist_empty_sets_and_nan():
    from turbo_b200 import device
    Xq = np.random.default_rng(0).random((5, 3))
    got, arg = device.min_sqdist(Xq, np.zeros((0, 3)), want_arg=True)
    assert np.all(np.isinf(got)) and np.all(arg == -1)  # turbo/utils.py:119-121: nothing to be close to
    assert device.min_sqdist(np.zeros((0, 3)), Xq).shape == (0,)
    X = np.random.default_rng(1).random((4, 3))
    X[2, 1] = np.nan
    got, arg = device.min_sqdist(Xq, X, want_arg=True)
    want, warg = O.min_sqdist(Xq, X)
    np.testing.assert_array_equal(got, want)  # a NaN row is never the nearest (NaN <= tol is False)
    np.testing.assert_array_equal(arg, warg)


@pytest.mark.parametrize("engine", ["dmma", "i8"])
def test_argmax_skips_candidates_too_close_to_a_sample(engine):
    import turbo_b200 as tb
    rng = np.random.default_rng(11)
    n, d = 150, 3
    X = rng.random((n, d))
    y = np.sin(3 * X.sum(1))
    spec = tb.KernelSpec(3, d, True)
    theta = np.log([1.0, 0.5, 0.5, 0.5, 1e-6])
    model = _fixed_model(spec, theta, X, y)
    # optimism with beta = 0 on 'max': the acquisition is the posterior mean, whose best candidates are the samples
    acq = tb.GpuUCB(0.0).construct_function(0, model, "max")[0]
    cand = np.vstack([rng.random((3000, d)), X, X + 3e-6, rng.random((777, d))])
    values = acq(cand)
    v0, i0, x0 = acq.argmax_over(candidates=cand, engine=engine)
    assert i0 == O.argmax_first(values)
    for tol in (1e-10, 1e-9, 0.01):
        mind, _ = O.min_sqdist(cand, X)
        masked = np.where(mind <= tol, -np.inf, values)
        v, i, x = acq.argmax_over(candidates=cand, engine=engine, exclude=(X, tol), batch=512)
        assert i == O.argmax_first(masked)
        assert not O.close_to_any(x, X, tol)
        np.testing.assert_array_equal(x[0], cand[i])
    # disabled (fallback.py:51-52) and empty sets change nothing
    assert acq.argmax_over(candidates=cand, engine=engine, exclude=(X, -1.0))[1] == i0
    assert acq.argmax_over(candidates=cand, engine=engine, exclude=(np.zeros((0, d)), 1e-3))[1] == i0
    # everything excluded: nothing selectable (auxiliary_optimisers.py:117-118)
    v, i, x = acq.argmax_over(candidates=X, engine=engine, exclude=(X, 1e-10))
    assert i == -1 and x is None and v == -np.inf


def test_aux_optimiser_close_tolerance_never_returns_a_known_point():
    import turbo_b200 as tb
    from types import SimpleNamespace
    rng = np.random.default_rng(3)
    n, d = 60, 2
    X = rng.random((n, d))
    y = np.cos(4 * X[:, 0]) + X[:, 1]
    model = _fixed_model(tb.KernelSpec(3, d, True), np.log([1.0, 0.4, 0.4, 1e-6]), X, y)
    acq = tb.GpuUCB(0.0).construct_function(0, model, "max")[0]
    bounds = SimpleNamespace(ordered=[("a", 0.0, 1.0), ("b", 0.0, 1.0)])
    pool = np.vstack([X, rng.random((500, d))])
    plain = tb.GpuCandidateArgmax(num_random=len(pool), candidates=lambda m, lb: pool, release_model=False)
    x_plain, _ = plain(bounds, acq)
    guarded = tb.GpuCandidateArgmax(num_random=len(pool), candidates=lambda m, lb: pool, release_model=False,
                                    close_tolerance=1e-10)
    x_g, info = guarded(bounds, acq)
    assert O.close_to_any(x_plain, X, 1e-10)      # the reference would now discard the trial (optimiser.py:349-352)
    assert not O.close_to_any(x_g, X, 1e-10)
    assert info["argmax_index"] >= n


# ---------------------------------------------------------------------------------------------
# the GPy-parametrised surrogate (turbo/modules/surrogates.py:84-222).  GPy is absent, so the checker is the
# oracle's restatement of the same model (variance inside the kernel + Gaussian noise, no jitter, Standardize).
# ---------------------------------------------------------------------------------------------
def _gpy_problem(n=120, d=3, seed=2):
    rng = np.random.default_rng(seed)
    X = rng.random((n, d))
    y = 3.0 + 2.0 * np.sin(3 * X.sum(1)) + 0.05 * rng.normal(size=n)
    return rng, X, y


def test_gpy_surrogate_fixed_parameters_match_the_restated_model():
    import turbo_b200 as tb
    rng, X, y = _gpy_problem()
    params = np.array([1.3, 0.4, 0.5, 0.6, 1e-3])
    k = tb.gpy_kernel('Matern52', ARD=True, variance=1.3, lengthscale=[0.4, 0.5, 0.6])
    sur = tb.GpuGPySurrogate(model_params=dict(kernel=k, normalizer=True, noise_var=1e-3), training_iterations=0)
    model, info = sur.construct_model(7, X, y)
    assert info["iterations"] == 0
    np.testing.assert_allclose(info["fixed"], params, rtol=1e-15)  # surrogates.py:176-177
    assert model.get_hyper_param_names() == ["Mat52.variance", "Mat52.lengthscale[0]", "Mat52.lengthscale[1]",
                                             "Mat52.lengthscale[2]", "Gaussian_noise.variance"]
    np.testing.assert_allclose(model.get_hyper_params(), params, rtol=1e-14)
    gp = O.FittedGP(O.KernelSpec(O.KIND_MATERN52, 3, True), np.log(params), X, y, normalize_y=True, alpha_jitter=0.0)
    Xc = rng.random((1000, 3))
    mu, sd = model.predict(Xc, return_std_dev=True)
    omu, osd = gp.predict(Xc)
    np.testing.assert_allclose(mu, omu, rtol=1e-8, atol=1e-8 * np.abs(omu).max())
    np.testing.assert_allclose(sd, osd, rtol=1e-8, atol=1e-8 * osd.max())
    assert sd.min() >= np.sqrt(1e-3) * gp.y_std * (1 - 1e-9)  # the noise term is part of the predictive variance
    assert abs(model.get_log_likelihood() - gp.lml()) <= 1e-6 * abs(gp.lml())
    assert model.predict(Xc).shape == (1000,)
    # the default model: isotropic RBF, unit parameters, GPy's names
    m2, _ = tb.GpuGPySurrogate(training_iterations=0).construct_model(0, X, y)
    assert m2.get_hyper_param_names() == ["rbf.variance", "rbf.lengthscale", "Gaussian_noise.variance"]
    np.testing.assert_allclose(m2.get_hyper_params(), [1.0, 1.0, 1.0], rtol=1e-15)
    import pickle
    m3 = pickle.loads(pickle.dumps(model))  # recorder.py:141-146 keeps every model
    np.testing.assert_allclose(m3.predict(Xc[:10]), mu[:10], rtol=1e-12)
    with pytest.raises(NotImplementedError):
        tb.GpuGPySurrogate(sparse=True, model_params=dict(kernel=k))


def test_gpy_surrogate_restarts_and_parameter_continuity():
    import turbo_b200 as tb
    rng, X, y = _gpy_problem(n=80)
    sur = tb.GpuGPySurrogate(training_iterations=lambda t: 4 if t == 0 else 1, seed=3)
    m0, info0 = sur.construct_model(0, X[:70], y[:70])
    assert info0["iterations"] == 4 and info0["restarts"] == 4  # restarts == iterations (surrogates.py:185-186)
    start = tb.GpuGPySurrogate(training_iterations=0).construct_model(0, X[:70], y[:70])[0]
    assert m0.get_log_likelihood() > start.get_log_likelihood()  # better than GPy's unit starting parameters
    # the optimum is a stationary point of the restated objective
    gp = O.FittedGP(O.KernelSpec(O.KIND_RBF, 3, False), np.log(m0.get_hyper_params()), X[:70], y[:70], True, 0.0)
    lml, grad = gp.lml(eval_gradient=True)
    assert abs(lml - m0.get_log_likelihood()) <= 1e-6 * abs(lml)
    assert np.abs(grad).max() < 1e-2
    # next trial: one restart, started from the stored parameter array (surrogates.py:171-173)
    m1, info1 = sur.construct_model(1, X, y)
    assert info1["restarts"] == 1
    np.testing.assert_allclose(sur._last_model_params, m1.get_hyper_params(), rtol=1e-14)
    assert np.all(np.abs(np.log(m1.get_hyper_params() / m0.get_hyper_params())) < 1.0)
