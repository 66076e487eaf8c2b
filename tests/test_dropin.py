"""The drop-in, dropped in: the UNMODIFIED reference driver (turbo.Optimiser._select_trial,
/root/reference/turbo/optimiser.py:318-357), recorder (turbo/recorder.py:117-155, :249-265) and preset
mechanism (turbo/optimiser_presets.py:8-37) running with GpuGPSurrogate + GpuEI + GpuCandidateArgmax.

`turbo` comes from the normal import path, or from baseline/_ref (the unmodified reference installed by
baseline/install_reference.py, which ships to the GPU box).  Without it these tests SKIP -- they never
fall back to the stand-in classes, whose only purpose is to keep `import turbo_b200` working.
"""
import os
import warnings

import numpy as np
import pytest

from conftest import GOLDEN

import turbo_b200 as tg
from turbo_b200 import _reference_api

if not _reference_api.HAVE_TURBO:
    pytest.skip("mbway/turbo is not importable (run baseline/install_reference.py where /root/reference exists)",
                allow_module_level=True)

import turbo as tb  # noqa: E402
import turbo.modules as tm  # noqa: E402

from oracle import gp_oracle as O  # noqa: E402  (checker only)

BRANIN_BOUNDS = [("x", -5.0, 10.0), ("y", 0.0, 15.0)]
BRANIN_OPT = 0.397887


class Branin:
    def __call__(self, x, y):
        return O.branin(x, y)


def _demo_model_params():
    import sklearn.gaussian_process as sk_gp
    K = sk_gp.kernels
    # demos/Branin-Hoo.ipynb cell 7
    return dict(kernel=1.0 * K.Matern(nu=2.5) + K.WhiteKernel(), normalize_y=True)


# ---------------------------------------------------------------------------------------------
# no GPU needed: type identity and preset wiring against the REAL classes
# ---------------------------------------------------------------------------------------------
def test_plugins_subclass_the_real_reference_classes():
    assert issubclass(tg.GpuGPSurrogate, tm.Surrogate)
    assert issubclass(tg.GpuGPSurrogate.ModelInstance, tm.SciKitGPSurrogate.ModelInstance)
    assert issubclass(tg.GpuGPySurrogate.ModelInstance, tm.GPySurrogate.ModelInstance)
    for ours, theirs in ((tg.GpuEI, tm.EI), (tg.GpuPI, tm.PI), (tg.GpuUCB, tm.UCB)):
        assert issubclass(ours, theirs)
        assert issubclass(ours.FunctionInstance, theirs.FunctionInstance)


def test_load_gpu_preset_fills_the_six_module_slots_of_a_real_optimiser():
    op = tb.Optimiser(Branin(), "min", BRANIN_BOUNDS, pre_phase_trials=10, settings_preset=None)
    tg.load_gpu_preset(op, num_random=1000, training_iterations=5, candidates="numpy", model_params=_demo_model_params())
    # the six slots of turbo/optimiser_presets.py:8-37, nothing else (Optimiser.__setattr__ would have raised)
    assert isinstance(op.latent_space, tm.NoLatentSpace)
    assert isinstance(op.pre_phase_select, tm.LHS_selector)
    assert isinstance(op.fallback, tm.Fallback)
    assert isinstance(op.aux_optimiser, tg.GpuCandidateArgmax)
    assert isinstance(op.surrogate, tm.Surrogate) and isinstance(op.surrogate, tg.GpuGPSurrogate)
    assert isinstance(op.acquisition, tm.EI) and op.acquisition.get_type() == "improvement"
    with pytest.raises(ValueError):
        tg.load_gpu_preset(op, "no such preset")
    # the configured optimiser pickles the way Recorder.save_compressed pickles it (no device state anywhere)
    import dill
    clone = dill.loads(dill.dumps(op))
    assert isinstance(clone.surrogate, tg.GpuGPSurrogate) and clone.aux_optimiser.num_random == 1000


# ---------------------------------------------------------------------------------------------
# on the GPU: the loop itself
# ---------------------------------------------------------------------------------------------
def _check_types(op, rec):
    bayes = [t for _n, t in rec.get_sorted_trials() if t.selection_info["type"] == "bayes"]
    assert bayes
    for t in bayes:
        m = t.selection_info["model"]
        # plotting/surrogates.py:67-75 (model type), plotting/overview.py:53-61 (acquisition type)
        assert isinstance(m, tm.SciKitGPSurrogate.ModelInstance)
        assert isinstance(m, tm.Surrogate.ModelInstance)
        assert len(m.get_hyper_params()) == len(m.get_hyper_param_names())
        assert "iterations" in t.selection_info["fitting_info"]          # plotting/overview.py:117-118
        assert np.isfinite(t.selection_info["maximisation_info"]["max_acq"])  # plotting/trials.py:195
    assert isinstance(op.acquisition, tm.EI)
    return bayes


@pytest.mark.gpu
def test_real_optimiser_runs_the_gpu_preset_and_the_recorder_round_trips(tmp_path):
    """config 1 through the unmodified driver: LHS pre-phase of 10, then 40 Bayesian trials."""
    np.random.seed(42)
    op = tb.Optimiser(Branin(), "min", BRANIN_BOUNDS, pre_phase_trials=10, settings_preset=None)
    tg.load_gpu_preset(op, num_random=1000, training_iterations=5, candidates="numpy", seed=0,
                       model_params=_demo_model_params())
    rec = tb.Recorder(op)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        op.run(max_trials=50)
    assert len(rec.trials) == 50 and not rec.has_unfinished_trials()
    bayes = _check_types(op, rec)
    assert len(bayes) >= 35  # a few may have been replaced by the Fallback's too-close rule
    for t in bayes:
        assert t.selection_info["fitting_info"]["iterations"] == 5
        assert t.selection_info["fitting_info"]["restarts"] == 5
    # the tolerance the reference's own end-to-end tests use (old_library/tests/optimisation_tests.py:1426-1490)
    _n, best = rec.get_incumbent()
    assert abs(best.y - BRANIN_OPT) <= 0.1, best.y

    # Recorder.save_compressed dills the recorder WITH the optimiser and every trial's model (recorder.py:117-155)
    path = str(tmp_path / "run.pkl.gz")
    rec.save_compressed(path)
    assert os.path.getsize(path) < 2_000_000  # theta + X + y per model, no O(n^2) state
    loaded = tb.Recorder.load_compressed(path, quiet=True)
    assert len(loaded.trials) == 50
    Xq = np.random.default_rng(3).uniform([-5.0, 0.0], [10.0, 15.0], size=(257, 2))
    for n_t in (10, 23, 49):
        live, back = rec.trials[n_t], loaded.trials[n_t]
        np.testing.assert_array_equal(live.x, back.x)
        m_live, m_back = live.selection_info["model"], back.selection_info["model"]
        assert isinstance(m_back, tm.SciKitGPSurrogate.ModelInstance)
        mu0, sd0 = m_live.predict(Xq, return_std_dev=True)
        mu1, sd1 = m_back.predict(Xq, return_std_dev=True)   # device state rebuilt lazily from theta + X + y
        np.testing.assert_array_equal(mu0, mu1)
        np.testing.assert_array_equal(sd0, sd1)
        # the acquisition function rebuilt from the stored model (recorder.py:249-265) == the live one == the oracle
        a_live = rec.get_acquisition_function(n_t)
        a_back = loaded.get_acquisition_function(n_t)
        assert isinstance(a_live, tm.EI.FunctionInstance) and a_live.get_name() == "EI"
        v_live, v_back = a_live(Xq), a_back(Xq)
        np.testing.assert_array_equal(v_live, v_back)
        finished = [rec.trials[k] for k in range(n_t)]
        X = np.vstack([f.x for f in finished])
        y = np.array([f.y for f in finished])
        og = O.FittedGP(O.KernelSpec(O.KIND_MATERN52, 2, False), m_live.theta, X, y)
        mu, sd = og.predict(Xq)
        ref = O.acquisition(O.ACQ_EI, mu, sd, -1.0, float(y.min()), 0.01)
        np.testing.assert_allclose(v_live, ref, rtol=1e-7, atol=1e-9 * np.abs(ref).max())
        assert m_live.get_log_likelihood() == pytest.approx(og.lml(), rel=1e-6)


@pytest.mark.gpu
def test_real_optimiser_with_pinned_theta_repeats_the_reference_run():
    """The reference's recorded Branin-Hoo run (tests/golden/branin_run.npz: the unmodified Optimiser with
    SciKitGPSurrogate + EI + RandomAndQuasiNewton's random stage) repeated through the SAME unmodified driver with the
    device plugins, hyper-parameters pinned per trial to the ones scikit-learn fitted there and the reference's own
    candidate draws: every trial must select the same point."""
    g = np.load(os.path.join(GOLDEN, "branin_run.npz"))
    xs, thetas, cands = g["xs"], g["thetas"], g["cands"]
    first = int(g["bayes_trials"][0])
    assert list(g["bayes_trials"]) == list(range(first, 50))
    spec = tg.KernelSpec(3, 2, False)

    class Pinned(tg.GpuGPSurrogate):
        def construct_model(self, trial_num, X, y, start_thetas=None):
            self.model_params["kernel"] = (spec, thetas[trial_num - first])
            return super().construct_model(trial_num, X, y)

    calls = []

    def recorded_candidates(num_points, latent_bounds):
        k = len(calls)
        calls.append(k)
        assert num_points == cands.shape[1]
        return cands[k]

    np.random.seed(42)
    op = tb.Optimiser(Branin(), "min", BRANIN_BOUNDS, pre_phase_trials=10, settings_preset=None)
    op.latent_space = tm.NoLatentSpace()
    op.pre_phase_select = tm.LHS_selector(num_total=10)
    op.fallback = tm.Fallback(selector=tm.random_selector())
    op.aux_optimiser = tg.GpuCandidateArgmax(num_random=1000, candidates=recorded_candidates)
    op.surrogate = Pinned(model_params=dict(kernel=(spec, thetas[0]), normalize_y=True), training_iterations=0)
    op.acquisition = tg.GpuEI(xi=0.01)
    rec = tb.Recorder(op)
    op.run(max_trials=50)
    bayes = _check_types(op, rec)
    got = np.vstack([t.x for _n, t in rec.get_sorted_trials()])
    np.testing.assert_array_equal(got[:first], xs[:first])   # the pre-phase is the reference's own code and stream
    np.testing.assert_array_equal(got, xs)                    # ... and every Bayesian trial picks the reference's point
    assert len(bayes) == 50 - first
    for k, t in enumerate(bayes):
        mi = t.selection_info["maximisation_info"]
        assert mi["max_acq"] == pytest.approx(float(g["max_acq"][k]), rel=1e-7, abs=1e-300)
        assert t.selection_info["model"].get_log_likelihood() == pytest.approx(float(g["lmls"][k]), rel=1e-6)
        assert "fixed" in t.selection_info["fitting_info"]


@pytest.mark.gpu
def test_incremental_refit_happens_inside_the_default_drop_in_loop():
    """with the aux-optimiser's default release_model=True the per-trial model drops its factor, and the
    O(n^2) incremental refit must still run: the surrogate keeps the latest factor alive itself."""
    np.random.seed(7)
    op = tb.Optimiser(Branin(), "min", BRANIN_BOUNDS, pre_phase_trials=6, settings_preset=None)
    # hyper-parameters are fitted on the first Bayesian trial and every 5th after it, fixed in between
    # (the reference's `training_iterations` callable, surrogates.py:250-269)
    tg.load_gpu_preset(op, num_random=4096, candidates="uniform", seed=1, model_params=_demo_model_params(),
                       training_iterations=lambda trial_num: 4 if (trial_num - 6) % 5 == 0 else 0)
    assert op.aux_optimiser.release_model is True
    rec = tb.Recorder(op)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        op.run(max_trials=18)
    flags = []
    for n_t, t in rec.get_sorted_trials():
        if t.selection_info["type"] != "bayes":
            continue
        fi = t.selection_info["fitting_info"]
        flags.append((n_t, fi["iterations"], bool(fi.get("incremental", False))))
        assert t.selection_info["model"]._factor is None  # released after the argmax
    fixed = [f for f in flags if f[1] == 0]
    assert fixed and all(f[2] for f in fixed), flags       # every fixed-theta trial extended the previous factor
    assert not any(f[2] for f in flags if f[1] > 0), flags  # a refit never does
    # and the incrementally built models agree with a from-scratch factorisation / the oracle
    n_t = max(f[0] for f in fixed)
    t = rec.trials[n_t]
    m = t.selection_info["model"]
    X = np.vstack([rec.trials[k].x for k in range(n_t)])
    y = np.array([rec.trials[k].y for k in range(n_t)])
    og = O.FittedGP(O.KernelSpec(O.KIND_MATERN52, 2, False), m.theta, X, y)
    assert t.selection_info["fitting_info"].get("lml", m.get_log_likelihood()) == pytest.approx(og.lml(), rel=1e-6)
    Xq = np.random.default_rng(5).uniform([-5.0, 0.0], [10.0, 15.0], size=(64, 2))
    mu, sd = m.predict(Xq, return_std_dev=True)
    mu_o, sd_o = og.predict(Xq)
    np.testing.assert_allclose(mu, mu_o, rtol=1e-8, atol=1e-8 * np.abs(mu_o).max())
    np.testing.assert_allclose(sd, sd_o, rtol=1e-6, atol=1e-9 * np.abs(sd_o).max())
