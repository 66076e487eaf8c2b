"""Generate tests/golden/*.npz from the REFERENCE ITSELF (run in the build container only).

    PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden.py

Imports mbway/turbo from /root/reference (read-only) and scikit-learn, runs the
reference's own objects on seeded inputs, and stores inputs + outputs:

* ``SciKitGPSurrogate.ModelInstance(GaussianProcessRegressor(kernel=fixed,
  optimizer=None).fit(X, y))`` (turbo/modules/surrogates.py:328-338; the
  ``training_iterations=0`` branch at :306-310 raises UnboundLocalError, so fixed
  hyper-parameters are reached by building the sklearn model directly),
* ``EI/PI/UCB(...).construct_function(...)`` (acquisition_functions.py),
* ``log_marginal_likelihood(theta, eval_gradient=True)`` (sklearn:_gpr.py:541-656),
* the candidate argmax of auxiliary_optimisers.py:60-66,
* the Branin-Hoo demo loop (config 1) through the unmodified ``turbo.Optimiser``
  with ``SciKitGPSurrogate`` + EI, every trial's X, y, fitted theta and the
  candidates / chosen point of that trial.

/root/reference does not exist on the GPU box; the tests read only the .npz files.
"""
import os
import sys
import warnings

import numpy as np

sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

if not hasattr(np, "asscalar"):  # removed in NumPy>=1.23; auxiliary_optimisers.py:125 needs it
    np.asscalar = lambda a: np.asarray(a).item()

import sklearn  # noqa: E402
import sklearn.gaussian_process as sk_gp  # noqa: E402
import scipy  # noqa: E402
import turbo as tb  # noqa: E402
import turbo.modules as tm  # noqa: E402

from oracle import gp_oracle as O  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
VERSIONS = "sklearn %s scipy %s numpy %s" % (sklearn.__version__, scipy.__version__, np.__version__)
K = sk_gp.kernels


def make_kernel(kind, sf2, ls, noise, has_const=True, has_white=True):
    base = K.RBF(length_scale=ls) if kind == O.KIND_RBF else K.Matern(length_scale=ls, nu={1: 0.5, 2: 1.5, 3: 2.5}[kind])
    k = K.ConstantKernel(sf2) * base if has_const else base
    return k + K.WhiteKernel(noise) if has_white else k


def fixed_case(name, kind, ard, n, d, m, seed, fn, sf2, ls, noise, normalize_y=True,
               extremum="min", xi=0.01, beta=2.0, has_const=True, has_white=True, grad=True, dup=0):
    rng = np.random.default_rng(seed)
    X = rng.random((n, d))
    y = fn(X, rng)
    Xc = rng.random((m, d))
    if dup:  # candidates equal to training points: sigma ~ 0 edge
        Xc[:dup] = X[:dup]
    ls_arr = np.full(d, ls) if ard else ls
    kern = make_kernel(kind, sf2, ls_arr, noise, has_const, has_white)
    with warnings.catch_warnings(record=True) as ws:
        warnings.simplefilter("always")
        gp = sk_gp.GaussianProcessRegressor(kernel=kern, optimizer=None, normalize_y=normalize_y).fit(X, y)
        model = tm.SciKitGPSurrogate.ModelInstance(gp)
        mu, sigma = model.predict(Xc, return_std_dev=True)
        inc = float(y.min() if extremum == "min" else y.max())
        ei = tm.EI(xi).construct_function(0, model, extremum, inc)[0](Xc)
        pi = tm.PI(xi).construct_function(0, model, extremum, inc)[0](Xc)
        ucb = tm.UCB(beta).construct_function(0, model, extremum)[0](Xc)
        sig_only = tm.UCB(float("inf")).construct_function(0, model, extremum)[0](Xc)
    theta = gp.kernel_.theta.copy()
    out = dict(X=X, y=y, Xc=Xc, theta=theta, mu=mu, sigma=sigma, ei=ei, pi=pi, ucb=ucb, sig_only=sig_only,
               incumbent=inc, xi=xi, beta=beta, kind=kind, ard=int(ard), has_const=int(has_const),
               has_white=int(has_white), normalize_y=int(normalize_y), extremum=extremum,
               ei_argmax=int(np.argsort(-ei, axis=0).flatten()[0]),
               ucb_argmax=int(np.argsort(-ucb, axis=0).flatten()[0]),
               pi_argmax=int(np.argsort(-pi, axis=0).flatten()[0]),
               lml=gp.log_marginal_likelihood(theta), alpha_=gp.alpha_, L_diag=np.diag(gp.L_),
               n_warnings=len(ws), versions=VERSIONS, hyper_names=np.array(model.get_hyper_param_names()),
               hyper_params=model.get_hyper_params())
    if grad:
        thetas = [theta, theta + rng.normal(0, 0.3, theta.shape)]
        lg = [gp.log_marginal_likelihood(t, eval_gradient=True) for t in thetas]
        out.update(lml_thetas=np.array(thetas), lml_vals=np.array([v for v, _ in lg]),
                   lml_grads=np.array([g for _, g in lg]))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, "n=%d d=%d m=%d theta=%s lml=%.6f ei_argmax=%d warnings=%d" % (
        n, d, m, np.round(theta, 3), out["lml"], out["ei_argmax"], len(ws)))


def f_hartmann(X, rng):
    return O.hartmann6(X)


def f_sines(X, rng):
    return np.sin(2 * np.pi * X).sum(axis=1) + 0.01 * rng.standard_normal(X.shape[0])


def branin_run():
    """config 1: the demo workload through the unmodified turbo.Optimiser."""
    np.random.seed(42)

    class BH:
        bounds = [("x", -5.0, 10.0), ("y", 0.0, 15.0)]

        def __call__(self, x, y):
            return O.branin(x, y)

    cands = []
    gen = tm.random_selector()

    class RecordingAux(tm.RandomAndQuasiNewton):
        def __call__(self, latent_bounds, acq):
            X = gen(self.num_random, latent_bounds)
            cands.append(X)
            self.gen_random = lambda n, lb: X  # feed the recorded candidates to the stock code
            return super().__call__(latent_bounds, acq)

    op = tb.Optimiser(BH(), "min", BH.bounds, pre_phase_trials=10, settings_preset=None)
    op.latent_space = tm.NoLatentSpace()
    op.pre_phase_select = tm.LHS_selector(num_total=10)
    op.fallback = tm.Fallback(selector=tm.random_selector())
    op.aux_optimiser = RecordingAux(num_random=1000, grad_restarts=0, start_from_best=0)
    op.surrogate = tm.SciKitGPSurrogate(model_params=dict(
        kernel=1.0 * K.Matern(nu=2.5) + K.WhiteKernel(), normalize_y=True), training_iterations=5)
    op.acquisition = tm.EI(xi=0.01)
    rec = tb.Recorder(op)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        op.run(max_trials=50)
    trials = sorted(rec.trials.values(), key=lambda t: t.trial_num)
    xs = np.vstack([t.x for t in trials])
    ys = np.array([t.y for t in trials])
    bayes = [t for t in trials if t.selection_info["type"] == "bayes"]
    thetas = np.array([t.selection_info["model"].model.kernel_.theta for t in bayes])
    max_acq = np.array([t.selection_info["maximisation_info"]["max_acq"] for t in bayes])
    lmls = np.array([t.selection_info["model"].get_log_likelihood() for t in bayes])
    nums = np.array([t.trial_num for t in bayes])
    assert len(cands) >= len(bayes)
    np.savez_compressed(os.path.join(OUT, "branin_run.npz"), xs=xs, ys=ys, bayes_trials=nums, thetas=thetas,
                        max_acq=max_acq, lmls=lmls, cands=np.array(cands[:len(bayes)]), versions=VERSIONS)
    print("branin_run: %d trials (%d bayes), best y=%.6f" % (len(trials), len(bayes), ys.min()))


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    # config 2 shape at reduced n (full-size properties are tested separately)
    fixed_case("hartmann6_n256", O.KIND_MATERN52, True, 256, 6, 2048, 0, f_hartmann, 1.0, 0.5, 1e-6)
    fixed_case("hartmann6_n1024", O.KIND_MATERN52, True, 1024, 6, 4096, 1, f_hartmann, 1.0, 0.5, 1e-6, grad=False)
    # config 3/4 shape at reduced n: ARD Matern-5/2, d=10
    fixed_case("sines_m52_n300", O.KIND_MATERN52, True, 300, 10, 1500, 2, f_sines, 1.3, 0.7, 1e-3, extremum="max")
    # config 5 shape at reduced n: ARD RBF, d=20
    fixed_case("sines_rbf_n200", O.KIND_RBF, True, 200, 20, 1000, 3, f_sines, 0.8, 1.5, 1e-2)
    # other kernel families / layouts / edges
    fixed_case("iso_m32_n97", O.KIND_MATERN32, False, 97, 3, 777, 4, f_sines, 2.0, 0.4, 1e-4, normalize_y=False)
    fixed_case("iso_m12_n64", O.KIND_MATERN12, False, 64, 2, 500, 5, f_sines, 1.0, 0.3, 1e-5, extremum="max")
    fixed_case("iso_rbf_noconst_n50", O.KIND_RBF, False, 50, 4, 300, 6, f_sines, 1.0, 0.6, 1e-3, has_const=False)
    fixed_case("ard_m52_nowhite_n40", O.KIND_MATERN52, True, 40, 5, 300, 7, f_sines, 1.5, 0.9, 0.0, has_white=False)
    fixed_case("dup_candidates_n128", O.KIND_MATERN52, True, 128, 4, 512, 8, f_sines, 1.0, 0.5, 1e-8, dup=64)
    fixed_case("tiny_n3", O.KIND_MATERN52, False, 3, 1, 64, 9, f_sines, 1.0, 0.5, 1e-2)
    branin_run()
