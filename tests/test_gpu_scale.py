"""GPU parity at the BASELINE.json configuration SIZES against the oracle:
config 5's n=16384, d=20 ARD RBF predict + EI, and config 3's batched LML + gradient at n=4096 with 64 restarts on the
INT8 left-looking schedule.  The oracle needs tens of seconds at these sizes on the box's host cores; each case
computes it once."""
import numpy as np
import pytest

from conftest import check_variance
from oracle import gp_oracle as O

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module", autouse=True)
def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def test_cfg5_shape_predict_and_ei_match_oracle():
    """n=16384, d=20, ARD RBF (bench.py --config cfg5's model): rows of Linv are longer than one exact INT32 accumulation
    pass (8192 terms), so the variance kernel runs its multi-pass path; 384 candidates, 64 of them ON training points."""
    import bench
    import turbo_b200 as tb
    cfg = bench.CONFIGS["cfg5"]
    n, d = cfg["n"], cfg["d"]
    X, y, theta = bench.synthetic_problem(n, d, 0, cfg["problem"], cfg["sf2"], cfg["ls"], cfg["noise"])
    rng = np.random.default_rng(5)
    Xc = rng.random((384, d))
    Xc[:64] = X[rng.choice(n, 64, replace=False)]  # sigma^2 -> the noise level: the cancellation case
    spec = tb.KernelSpec(cfg["kind"], d, True)
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0)
    model, _ = sur.construct_model(0, X, y)
    assert tb.device.resolve_engine(None, model.factor.NP) == "i8a"
    mu, sd = model.predict(Xc, return_std_dev=True)
    acq = tb.GpuEI(0.01).construct_function(0, model, "min", float(y.min()))[0]
    vals = acq(Xc)
    val, idx, x = acq.argmax_over(candidates=Xc)

    og = O.FittedGP(O.KernelSpec(cfg["kind"], d, True), theta, X, y)
    mu_o, sd_o = og.predict(Xc)
    ref = O.acquisition(O.ACQ_EI, mu_o, sd_o, -1.0, float(y.min()), 0.01)
    np.testing.assert_allclose(mu, mu_o, rtol=1e-8, atol=1e-8 * np.abs(y).max())
    prior = (np.exp(theta[0]) + np.exp(theta[-1])) * og.y_std ** 2
    check_variance("cfg5/n16384_d20_rbf", sd ** 2, sd_o ** 2, n, prior)
    np.testing.assert_allclose(vals, ref, rtol=1e-8, atol=1e-8 * np.abs(ref).max())
    assert idx == O.argmax_first(ref)
    np.testing.assert_array_equal(x[0], Xc[idx])
    assert val == vals[idx]
    assert model.get_log_likelihood() == pytest.approx(og.lml(), rel=1e-6)


def test_cfg3_batched_lml_and_gradient_on_the_int8_schedule_match_oracle():
    """n=4096, d=10, R=64 (bench.py's `fit` block): potrf column updates, the inverse-factor recurrence and K^-1 all run
    on the INT8 tensor cores here (28/28/21 slice products).  Three members are checked against the oracle's
    restatement of sklearn's log_marginal_likelihood(eval_gradient=True): the first restart, an ill-conditioned one
    (tiny noise, long length scales) and the last."""
    import bench
    import turbo_b200 as tb
    from turbo_b200 import device
    n, d, R = 4096, 10, 64
    X, y, theta = bench.synthetic_problem(n, d)
    y_n = (y - y.mean()) / y.std()
    rng = np.random.default_rng(2)
    T = np.clip(theta + rng.normal(0, 0.5, (R, theta.size)), np.log(1e-2) * 2, np.log(1e2))
    T[1, -1], T[1, 1:-1] = np.log(1e-6), np.log(2.0)   # cond(K) ~ 1e9
    spec = tb.KernelSpec(3, d, True)
    lml, grad, info = device.lml_batched(device.to_device(X), device.to_device(y_n), n, spec, T, 1e-10, True)
    assert not info.any() and np.isfinite(lml).all() and np.isfinite(grad).all()
    ospec = O.KernelSpec(3, d, True)
    for r in (0, 1, R - 1):
        ol, og = O.log_marginal_likelihood(ospec, T[r], X, y_n, 1e-10, True)
        assert lml[r] == pytest.approx(ol, rel=1e-6), r
        np.testing.assert_allclose(grad[r], og, rtol=1e-6, atol=1e-6 * np.abs(og).max(), err_msg=str(r))


def test_two_devices_in_one_process():
    """VERDICT r1 / ADVICE: function attributes (dynamic shared-memory opt-in, carve-out) are per device; a process that
    moves to a second GPU must get them there too.  Skipped on a single-GPU box."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import turbo_b200 as tb
    rng = np.random.default_rng(0)
    n, d = 700, 5
    X, y = rng.random((n, d)), rng.standard_normal(n)
    Xc = rng.random((1000, d))
    spec = tb.KernelSpec(3, d, True)
    theta = np.log(np.r_[1.0, np.full(d, 0.5), 1e-3])
    outs = []
    for dev in (0, 1, 0):
        torch.cuda.set_device(dev)
        sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=3,
                                max_iter=5, seed=0)
        model, _ = sur.construct_model(0, X, y)           # batched fit: DMMA + diag kernels with > 48 KB of shared memory
        mu, sd = model.predict(Xc, return_std_dev=True)   # i8 variance kernel: 173 KB
        acq = tb.GpuEI(0.01).construct_function(0, model, "min", float(y.min()))[0]
        outs.append((mu, sd, acq.argmax_over(candidates=Xc)[:2], model.theta))
    torch.cuda.set_device(0)
    for o in outs[1:]:
        np.testing.assert_array_equal(o[0], outs[0][0])
        np.testing.assert_array_equal(o[1], outs[0][1])
        assert o[2] == outs[0][2]
    tb.device.shutdown()  # streams/events of both devices are released and re-created on demand
    torch.cuda.set_device(1)
    mu2 = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, outs[0][3]), normalize_y=True),
                            training_iterations=0).construct_model(0, X, y)[0].predict(Xc)
    np.testing.assert_array_equal(mu2, outs[0][0])
    torch.cuda.set_device(0)
