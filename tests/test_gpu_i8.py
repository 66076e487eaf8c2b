"""INT8 tensor-core variance engine (csrc/ozaki_i8.cuh) against the DMMA engine, the oracle and the golden
vectors: same tolerances as the FP64 path (mean/variance/acquisition 1e-8 relative with the documented
absolute variance floor; argmax exact)."""
import os

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


def _model(g, engine):
    import turbo_b200 as tb
    d = g["X"].shape[1]
    spec = tb.KernelSpec(g["kind"], d, g["ard"], g["has_const"], g["has_white"])
    sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, g["theta"]), normalize_y=bool(g["normalize_y"])),
                            training_iterations=0, variance_engine=engine)
    return sur.construct_model(0, g["X"], g["y"])[0]


def test_i8l_precision_is_recorded(golden):
    """21 slice products (6 digit planes, truncation ~2^-46 per term): NOT the default engine.  The variance error it
    achieves on every fixture is recorded next to the other engines' (gpurun_out/variance_error.json) and must stay
    inside the 64-unit floor; 'auto' uses this engine only for models whose probe shows <= 8 units."""
    from conftest import check_variance
    g = golden
    model = _model(g, "i8l")
    mu, sd = model.predict(g["Xc"], return_std_dev=True)
    prior = np.exp(g["theta"][0]) if g["has_const"] else 1.0
    np.testing.assert_allclose(mu, g["mu"], rtol=1e-8, atol=1e-8 * np.abs(g["mu"]).max())
    check_variance("golden/%s/i8l" % g["name"], sd ** 2, g["sigma"] ** 2, g["X"].shape[0], prior * model.y_std ** 2)


@pytest.mark.parametrize("engine", ["i8", "i8x"])
def test_i8_predict_and_acquisition_match_reference(golden, engine):
    import turbo_b200 as tb
    g = golden
    model = _model(g, engine)
    mu, sd = model.predict(g["Xc"], return_std_dev=True)
    prior = np.exp(g["theta"][0]) if g["has_const"] else 1.0
    floor = 4 * EPS * g["X"].shape[0] * prior * model.y_std ** 2
    np.testing.assert_allclose(mu, g["mu"], rtol=1e-8, atol=1e-8 * np.abs(g["mu"]).max())
    from conftest import check_variance, VAR_FLOOR_EXACT
    check_variance("golden/%s/%s" % (g["name"], engine), sd ** 2, g["sigma"] ** 2, g["X"].shape[0],
                   prior * model.y_std ** 2, c=VAR_FLOOR_EXACT)
    acq, _ = tb.GpuEI(g["xi"]).construct_function(0, model, g["extremum"], g["incumbent"])
    vals = acq(g["Xc"])
    ref = g["ei"]
    tol = 1e-8 * np.abs(ref).max() + np.sqrt(floor)  # EI moves by at most |d sigma| = sqrt(var floor) near sigma = 0
    np.testing.assert_allclose(vals, ref, rtol=1e-8, atol=tol)
    val, idx, x = acq.argmax_over(candidates=g["Xc"])
    order = np.argsort(-ref, kind="stable")
    gap = ref[order[0]] - ref[order[1]]
    if gap > 1e-9 * abs(ref[order[0]]):
        assert idx == g["ei_argmax"]
        np.testing.assert_array_equal(x[0], g["Xc"][idx])


# 8200 -> NP = 8320 = 130 row tiles: more than one INT32 accumulation pass (PASS_KB = 128 k-blocks) for the last rows
@pytest.mark.parametrize("n,d,kind,noise", [(1000, 6, 3, 1e-6), (2048, 3, 0, 1e-8), (4096, 10, 3, 1e-4),
                                            (8200, 8, 3, 1e-4)])
def test_i8_matches_dmma_at_size(n, d, kind, noise):
    """full-size, including an ill-conditioned factor (large rows of L^-1): the three engines agree on the
    variance to the parity floor, and on the selected candidate"""
    import turbo_b200 as tb
    rng = np.random.default_rng(n)
    X = rng.random((n, d))
    y = np.sin(2 * np.pi * X).sum(axis=1) + 0.01 * rng.standard_normal(n)
    spec = tb.KernelSpec(kind, d, True)
    theta = np.log(np.r_[1.0, np.full(d, 0.5), noise])
    m = 5000
    Xc = rng.random((m, d))
    Xc[:16] = X[:16]  # duplicates of training points: var cancels to ~noise
    out = {}
    for engine in ("dmma", "i8", "i8x"):
        sur = tb.GpuGPSurrogate(model_params=dict(kernel=(spec, theta), normalize_y=True), training_iterations=0,
                                variance_engine=engine)
        model = sur.construct_model(0, X, y)[0]
        mu, sd = model.predict(Xc, return_std_dev=True)
        acq, _ = tb.GpuEI(0.01).construct_function(0, model, 'min', float(y.min()))
        val, idx, _ = acq.argmax_over(candidates=Xc)
        out[engine] = (mu, sd ** 2, idx, model.y_std)
    floor = 4 * EPS * n * out["dmma"][3] ** 2  # conftest.VAR_FLOOR_EXACT: engines with >= 28 slice products
    og = O.FittedGP(O.KernelSpec(kind, d, True), theta, X, y)
    # mu = y_std * sum_j K*_j alpha_j + y_mean: the engines add the n terms in different orders, so they agree to the
    # rounding of that sum, eps * sum_j |K*_j alpha_j| <= eps * sf2 * |alpha|_1 (1e-8 absolute for the ill-conditioned
    # noise = 1e-8 case where |alpha|_1 ~ 1e8), not to 1e-10 of the (cancelled) result
    mu_floor = 8 * EPS * np.abs(og.alpha).sum() * og.y_std
    for engine in ("i8", "i8x"):
        np.testing.assert_allclose(out[engine][0], out["dmma"][0], rtol=1e-10, atol=1e-12 + mu_floor)
        np.testing.assert_allclose(out[engine][1], out["dmma"][1], rtol=1e-8, atol=floor)
        assert out[engine][2] == out["dmma"][2]
    # i8x (34 slice products) is as accurate as the FP64 GEMM: compare both to the oracle's TRSM form
    _, sd_o = og.predict(Xc[:512])
    e_dmma = np.abs(out["dmma"][1][:512] - sd_o ** 2).max()
    e_i8x = np.abs(out["i8x"][1][:512] - sd_o ** 2).max()
    assert e_i8x <= 4 * e_dmma + 64 * EPS * out["dmma"][3] ** 2


def test_i8_rejects_large_n():
    from turbo_b200 import device
    with pytest.raises(ValueError):
        device.resolve_engine("i8", 65536)
    assert device.resolve_engine("auto", 65536) == "dmma" and device.resolve_engine("auto", 16384) == "i8a"


def test_max_dimension_and_slot_reuse_over_many_small_batches():
    """d = TB200_MAX_D (64) and a candidate set that is 11 ragged batches of 256 (every buffer slot is reused several
    times, the last batch is partial): mean/variance match the oracle, the fused argmax equals evaluate-then-argmax,
    for both engines; generated (uniform, LHS) sets give the same winner on both engines"""
    import turbo_b200 as tb
    from turbo_b200 import device
    rng = np.random.default_rng(64)
    n, d, m = 200, 64, 2700
    X = rng.random((n, d))
    y = np.sin(X[:, :5]).sum(axis=1) + 0.01 * rng.standard_normal(n)
    theta = np.log(np.r_[1.0, np.full(d, 3.0), 1e-3])
    Xc = rng.random((m, d))
    og = O.FittedGP(O.KernelSpec(3, d, True), theta, X, y)
    mu_o, sd_o = og.predict(Xc)
    ei_o = O.acquisition(O.ACQ_EI, mu_o, sd_o, -1.0, float(y.min()), 0.01)
    winners = {}
    for engine in ("dmma", "i8"):
        sur = tb.GpuGPSurrogate(model_params=dict(kernel=(tb.KernelSpec(3, d, True), theta), normalize_y=True),
                                training_iterations=0, variance_engine=engine)
        model = sur.construct_model(0, X, y)[0]
        mu, sd = model.predict(Xc, return_std_dev=True)
        np.testing.assert_allclose(mu, mu_o, rtol=1e-8, atol=1e-8 * np.abs(mu_o).max())
        np.testing.assert_allclose(sd ** 2, sd_o ** 2, rtol=1e-8, atol=4 * EPS * n * model.y_std ** 2)
        acq = tb.GpuEI(0.01).construct_function(0, model, 'min', float(y.min()))[0]
        val, idx, x = acq.argmax_over(candidates=Xc, batch=256)
        assert idx == O.argmax_first(ei_o) and np.array_equal(x[0], Xc[idx])
        assert val == pytest.approx(ei_o[idx], rel=1e-8)
        for gen in ("uniform", "lhs"):
            v, i, xg = acq.argmax_over(generator=gen, m=3001, lo=[0.0] * d, hi=[1.0] * d, seed=11, batch=256)
            winners.setdefault(gen, []).append((i, xg.tobytes()))
            assert acq(xg)[0] == pytest.approx(v, rel=1e-8)
    assert winners["uniform"][0] == winners["uniform"][1] and winners["lhs"][0] == winners["lhs"][1]


def test_max_dimension_lml_and_gradient():
    """the fit kernels at d = 64 (64-70 KB of dynamic shared memory): LML and gradient against the oracle"""
    from turbo_b200 import device
    import turbo_b200 as tb
    rng = np.random.default_rng(65)
    n, d = 150, 64
    X = rng.random((n, d))
    y = np.sin(X[:, :5]).sum(axis=1) + 0.01 * rng.standard_normal(n)
    y = (y - y.mean()) / y.std()
    T = np.log(np.r_[1.0, np.full(d, 3.0), 1e-3]) + rng.normal(0, 0.2, (3, d + 2))
    lml, grad, info = device.lml_batched(device.to_device(X), device.to_device(y), n, tb.KernelSpec(3, d, True), T, 1e-10)
    assert not info.any()
    for r in range(3):
        ol, og = O.log_marginal_likelihood(O.KernelSpec(3, d, True), T[r], X, y, 1e-10, True)
        assert lml[r] == pytest.approx(ol, rel=1e-6)
        np.testing.assert_allclose(grad[r], og, rtol=1e-6, atol=1e-6 * np.abs(og).max())


_FIT_CHILD = r"""
import json, sys
import os

import numpy as np
sys.path.insert(0, %(root)r)
import turbo_b200 as tb
from turbo_b200 import device
rng = np.random.default_rng(%(seed)d)
n, d, R = %(n)d, %(d)d, %(R)d
X = rng.random((n, d))
y = np.sin(3 * X.sum(1)) + 0.01 * rng.standard_normal(n)
y = (y - y.mean()) / y.std()
spec = tb.KernelSpec(3, d, True)
T = np.log(np.r_[1.0, np.full(d, 0.3), 1e-3]) + rng.normal(0, 0.2, (R, d + 2))
X_dev, y_dev = device.to_device(X), device.to_device(y)
out = {}
if R > 1:
    lml, grad, info = device.lml_batched(X_dev, y_dev, n, spec, T, 1e-10)
    out.update(lml=lml.tolist(), grad=grad.tolist(), info=info.tolist())
else:
    f = device.Factor(X_dev, y_dev, n, spec, T[0], 1e-10)
    acq_c = device.acq_spec(2, 1.0, 0.0, 0.0, 0.0, 1.0)
    import torch
    Xc = device.to_device(rng.random((256, d)))
    mu = torch.empty(256, dtype=torch.float64, device=Xc.device)
    sg = torch.empty(256, dtype=torch.float64, device=Xc.device)
    device.Acquirer(f, acq_c, engine="dmma").run(Xc, 256, 0, None, mu=mu, sigma=sg)
    out.update(lml=float(f.lml), info=int(f.info), mu=mu.cpu().tolist(), sigma=sg.cpu().tolist())
print("RESULT" + json.dumps(out))
"""


def _fit_in_child(engine, **kw):
    import json
    import subprocess
    import sys
    env = dict(os.environ)
    env.pop("TB200_DEBUG", None)
    if engine == "dmma":
        env["TB200_DEBUG"] = "fit_engine=dmma"  # the library's one debug hook (csrc/common.cuh), read once per process
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = subprocess.run([sys.executable, "-c", _FIT_CHILD % dict(root=root, **kw)], env=env, capture_output=True,
                         text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    line = [ln for ln in res.stdout.splitlines() if ln.startswith("RESULT")][-1]
    return json.loads(line[len("RESULT"):])


def test_batched_fit_beyond_one_int32_window_matches_dmma():
    """n > 8192: the fit GEMMs cut their k-ranges into exact INT32 windows of 8192 terms (ozaki_gemm.cuh); the whole
    batched LML + gradient must agree with the FP64 DMMA engine (each engine in its own process: the switch is read
    once)"""
    kw = dict(n=8300, d=4, R=4, seed=7)
    a, b = _fit_in_child("i8", **kw), _fit_in_child("dmma", **kw)
    assert a["info"] == b["info"] == [0, 0, 0, 0]
    np.testing.assert_allclose(a["lml"], b["lml"], rtol=1e-9)
    ga, gb = np.array(a["grad"]), np.array(b["grad"])
    assert np.abs(ga - gb).max() <= 1e-7 * np.abs(gb).max()


def test_single_large_factor_on_the_int8_schedule_matches_dmma():
    """one matrix with n >= 12288 takes the batched INT8 schedule too (a block column alone fills the machine)"""
    kw = dict(n=12400, d=3, R=1, seed=9)
    a, b = _fit_in_child("i8", **kw), _fit_in_child("dmma", **kw)
    assert a["info"] == b["info"] == 0
    assert a["lml"] == pytest.approx(b["lml"], rel=1e-9)
    np.testing.assert_allclose(a["mu"], b["mu"], rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(a["sigma"], b["sigma"], rtol=1e-7, atol=1e-9)
