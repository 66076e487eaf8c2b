"""CPU tests (-m "not gpu"): the C-ABI library loads and exports every symbol the header declares,
and the host logic (kernel parsing, batched L-BFGS-B, sharding + reductions over gloo, candidate
generator mirror, plugin class identities).  No compute call into the library happens here."""
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from turbo_b200 import _lib
    header = open(os.path.join(ROOT, "include", "turbo_b200.h")).read()
    declared = set(re.findall(r"\b(tb200_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    lib = _lib.load()  # binds every symbol; AttributeError if one is missing
    assert lib.tb200_version() >= 100
    assert lib.tb200_padded_n(1) == 128 and lib.tb200_padded_n(128) == 128 and lib.tb200_padded_n(4097) == 4224
    assert lib.tb200_default_batch(4096) % 128 == 0
    assert lib.tb200_fit_workspace_bytes(4096, 10, 1, 1) > 2 * 4096 * 4096 * 8
    assert lib.tb200_acquire_workspace_bytes(4096, 10, 2048) > 2048 * 4096 * 8


def test_product_path_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "turbo_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "gp_oracle" not in src and "from oracle" not in src and "import oracle" not in src, f


def test_no_cpu_fallback_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    import turbo_b200 as tb
    sur = tb.GpuGPSurrogate(training_iterations=0)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sur.construct_model(0, np.zeros((3, 2)), np.zeros(3))


def test_kernel_spec_from_sklearn_matches_theta_layout():
    K = pytest.importorskip("sklearn.gaussian_process.kernels")
    from turbo_b200 import kernels, _lib
    k = 1.0 * K.Matern(nu=2.5) + K.WhiteKernel()  # the reference default (surrogates.py:237)
    spec, theta, bounds = kernels.from_sklearn(k, 2)
    assert (spec.kind, spec.ard, spec.has_const, spec.has_white, spec.p) == (_lib.KIND_MATERN52, False, True, True, 3)
    np.testing.assert_array_equal(theta, k.theta)
    np.testing.assert_array_equal(bounds, k.bounds)
    assert spec.names() == ["k1__k1__constant_value", "k1__k2__length_scale", "k2__noise_level"]
    k = K.ConstantKernel(2.0) * K.RBF(length_scale=[1.0, 2.0, 3.0])
    spec, theta, _ = kernels.from_sklearn(k, 3)
    assert (spec.kind, spec.ard, spec.has_white, spec.p) == (_lib.KIND_RBF, True, False, 4)
    assert spec.names() == ["k1__constant_value", "k2__length_scale_0", "k2__length_scale_1", "k2__length_scale_2"]
    spec, _, _ = kernels.from_sklearn(K.Matern(nu=1.5) + K.WhiteKernel(), 4)
    assert spec.names() == ["k1__length_scale", "k2__noise_level"] and spec.kind == _lib.KIND_MATERN32
    spec, _, _ = kernels.from_sklearn(K.RBF(), 1)
    assert spec.names() == ["length_scale"]
    for bad in (K.RationalQuadratic(), K.Matern(nu=0.7), K.RBF() + K.RBF(), K.WhiteKernel() + K.RBF(),
                K.ConstantKernel(1.0, "fixed") * K.RBF()):
        with pytest.raises(ValueError):
            kernels.from_sklearn(bad, 2)
    with pytest.raises(ValueError):
        kernels.from_sklearn(K.RBF(length_scale=[1.0, 2.0]), 3)


def test_training_iterations_semantics():
    import turbo_b200 as tb
    s = tb.GpuGPSurrogate(training_iterations=lambda t: 4 if t % 3 == 0 else 1)
    assert [s._get_training_iterations(t) for t in range(4)] == [4, 1, 1, 4]
    assert tb.GpuGPSurrogate(model_params=dict(kernel='default', n_restarts_optimizer=3))._get_training_iterations(0) == 3
    with pytest.raises(AssertionError):
        tb.GpuGPSurrogate(model_params=dict(kernel='default', n_restarts_optimizer=3), training_iterations=2)
    with pytest.raises(AssertionError):
        tb.GpuGPSurrogate()._get_training_iterations(0)
    with pytest.raises(ValueError):
        tb.GpuGPSurrogate(model_params=dict(kernel='default', copy_X_train=False))


def test_batched_lbfgs_converges_on_independent_problems():
    from turbo_b200 import lbfgs
    rng = np.random.default_rng(0)
    R, p = 6, 5
    A = [np.diag(rng.uniform(0.5, 20, p)) for _ in range(R)]
    c = rng.uniform(-3, 3, (R, p))
    bounds = np.tile([-2.0, 2.0], (p, 1))
    calls = []

    def fun(Xs, members):
        calls.append(len(members))
        f = np.array([0.5 * (x - c[r]) @ A[r] @ (x - c[r]) for x, r in zip(Xs, members)])
        g = np.array([A[r] @ (x - c[r]) for x, r in zip(Xs, members)])
        return f, g

    x, f, iters, n_eval = lbfgs.minimize_batched(fun, rng.uniform(-2, 2, (R, p)), bounds)
    np.testing.assert_allclose(x, np.clip(c, -2, 2), atol=1e-4)  # separable: the box optimum is the clipped centre
    assert n_eval == len(calls) and calls[0] == R and min(calls) < R  # converged members drop out of the batch


def test_batched_lbfgs_rosenbrock_and_infeasible_regions():
    from turbo_b200 import lbfgs

    def fun(Xs, members):
        f, g = [], []
        for x in Xs:
            if x[0] > 1.5:  # "Cholesky failed" region: the line search must back off
                f.append(np.inf); g.append(np.zeros(2)); continue
            f.append(100 * (x[1] - x[0] ** 2) ** 2 + (1 - x[0]) ** 2)
            g.append(np.array([-400 * x[0] * (x[1] - x[0] ** 2) - 2 * (1 - x[0]), 200 * (x[1] - x[0] ** 2)]))
        return np.array(f), np.array(g)

    x0 = np.array([[-1.2, 1.0], [0.0, 0.0], [1.4, -1.0], [1.8, 0.0]])
    x, f, iters, _ = lbfgs.minimize_batched(fun, x0, np.array([[-2.0, 2.0], [-2.0, 2.0]]), max_iter=500)
    np.testing.assert_allclose(x[:3], 1.0, atol=1e-3)
    assert np.isinf(f[3]) and iters[3] == 0  # started infeasible: reported, not crashed


def test_shard_ranges_tile_the_index_space():
    from turbo_b200 import dist
    for m in (0, 1, 7, 1000, 2 ** 24 + 3):
        for world in (1, 2, 3, 8):
            r = [dist.shard_range(m, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == m
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(h - l for l, h in r) - min(h - l for l, h in r) <= 1


def test_argmax_pair_reduction_rule():
    from turbo_b200.dist import reduce_argmax_pairs as red
    assert red([1.0, 3.0, 3.0, 2.0], [10, 30, 20, 5]) == (3.0, 20)      # tie -> lowest index
    assert red([float("nan"), 1.0], [0, 9]) == (1.0, 9)                  # NaN never wins
    assert red([-np.inf, -np.inf], [-1, -1]) == (-np.inf, -1)            # empty shards
    assert red([-np.inf, 5.0], [-1, 3]) == (5.0, 3)
    assert red([-np.inf], [4]) == (-np.inf, 4)                           # a real -inf value is selectable


def test_merge_top_rule():
    from turbo_b200 import dist
    v, i = dist.merge_top([1.0, 3.0, 3.0, 2.0, np.nan, 9.0], [5, 9, 4, 6, 7, -1], 3)
    assert v.tolist() == [3.0, 3.0, 2.0] and i.tolist() == [4, 9, 6]      # NaN and index -1 never enter
    v, i = dist.merge_top([], [], 2)
    assert len(v) == 0 and len(i) == 0


def _gloo_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as td
    sys.path.insert(0, ROOT)
    from turbo_b200 import dist
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        out = {}
        out["rw"] = dist.rank_world("world")
        # each rank found a local best in its shard of 1000 candidates; ties across ranks -> lowest index
        local = [(2.5, 700), (2.5, 120)][rank]
        out["argmax"] = dist.allreduce_argmax(local[0], local[1], "world")
        out["empty"] = dist.allreduce_argmax(-np.inf, -1, "world")
        out["one_sided"] = dist.allreduce_argmax(*[(-np.inf, -1), (0.25, 999)][rank], "world")
        thetas = np.array([[0.1, 0.2, 0.3], [1.0, 2.0, 3.0]]) + rank
        negl = np.array([[5.0, 4.0], [np.inf, 3.5]][rank])
        out["best"] = dist.select_best_restart(thetas, negl, 3, "world")
        out["best_none_local"] = dist.select_best_restart(np.zeros((0, 3)), np.zeros(0), 3, "world") if rank == 1 \
            else dist.select_best_restart(thetas, negl, 3, "world")
        try:
            dist.select_best_restart(thetas, np.array([np.inf, np.inf]), 3, "world")
            out["all_failed"] = "no error"
        except np.linalg.LinAlgError:
            out["all_failed"] = "LinAlgError"
        # polish starts: every rank contributes local top candidates, every rank gets the same global top-k
        tv, ti = [([3.0, 1.0, 2.0], [40, 7, 9]), ([3.0, 2.5], [12, 900])][rank]
        out["top"] = dist.allgather_top(np.array(tv), np.array(ti), 3, "world")
        out["top_short"] = dist.allgather_top(np.array(tv[:rank]), np.array(ti[:rank]), 4, "world")
        # host-drawn candidate sets: every rank draws its own, rank 0's is the set everybody shards
        mine = np.random.default_rng(100 + rank).random((7 + rank, 3))
        out["bcast"] = dist.broadcast_array(mine, "world", src=0)
        q.put((rank, out))
    finally:
        td.destroy_process_group()


def test_world_size_2_reductions_over_gloo():
    import torch.multiprocessing as mp
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = dict(q.get(timeout=120) for _ in range(2))
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank in (0, 1):
        o = res[rank]
        assert o["rw"] == (rank, 2)
        assert o["argmax"] == (2.5, 120)
        assert o["empty"] == (-np.inf, -1)
        assert o["one_sided"] == (0.25, 999)
        theta, f = o["best"]
        assert f == 3.5 and np.array_equal(theta, [2.0, 3.0, 4.0])  # rank 1's second restart
        theta, f = o["best_none_local"]
        assert f == 4.0 and np.array_equal(theta, [1.0, 2.0, 3.0])
        assert o["all_failed"] == "LinAlgError"
        assert o["top"][0].tolist() == [3.0, 3.0, 2.5] and o["top"][1].tolist() == [12, 40, 900]
        assert o["top_short"][0].tolist() == [3.0] and o["top_short"][1].tolist() == [12]
        np.testing.assert_array_equal(o["bcast"], np.random.default_rng(100).random((7, 3)))  # rank 0's, shape included


def test_host_candidate_mirror_properties():
    from turbo_b200 import _lib
    lo, hi = [-5.0, 0.0], [10.0, 15.0]
    U = _lib.host_candidates(_lib.CAND_UNIFORM, 42, 0, lo, hi, np.arange(20000))
    assert U.shape == (20000, 2) and (U >= lo).all() and (U < hi).all()
    assert abs(U[:, 0].mean() - 2.5) < 0.15 and abs(U[:, 1].mean() - 7.5) < 0.15
    assert abs(np.corrcoef(U.T)[0, 1]) < 0.03
    np.testing.assert_array_equal(U[5000:5010], _lib.host_candidates(_lib.CAND_UNIFORM, 42, 0, lo, hi, range(5000, 5010)))
    assert not np.array_equal(U[:10], _lib.host_candidates(_lib.CAND_UNIFORM, 43, 0, lo, hi, range(10)))
    for m in (1, 2, 7, 1000, 4097):
        L = _lib.host_candidates(_lib.CAND_LHS, 3, m, [0.0, 0.0, 0.0], [1.0, 2.0, 4.0], np.arange(m))
        for k, h in enumerate((1.0, 2.0, 4.0)):
            assert sorted(np.floor(L[:, k] / h * m).astype(int)) == list(range(m))
    big = 2 ** 33 + 5  # 64-bit global indices
    _lib.host_candidates(_lib.CAND_UNIFORM, 1, 0, lo, hi, [big])


def test_plugins_subclass_the_reference_classes_when_turbo_is_importable():
    import turbo_b200 as tb
    from turbo_b200 import _reference_api as api
    assert issubclass(tb.GpuEI, api.EI) and issubclass(tb.GpuPI, api.PI) and issubclass(tb.GpuUCB, api.UCB)
    assert issubclass(tb.GpuGPSurrogate, api.Surrogate)
    assert issubclass(tb.GpuGPSurrogate.ModelInstance, api.SciKitModelInstance)
    # the GPy flavour answers plotting's GPySurrogate.ModelInstance check (plotting/surrogates.py:71), never the sklearn one
    assert issubclass(tb.GpuGPySurrogate, api.Surrogate)
    assert issubclass(tb.GpuGPySurrogate.ModelInstance, api.GPyModelInstance)
    assert tb.GpuGPySurrogate.ModelInstance.__mro__[1] is tb.GpuGPSurrogate.ModelInstance.__mro__[1]  # one device model
    assert tb.GpuEI(0.01).get_type() == "improvement" and tb.GpuUCB(2.0).get_type() == "optimism"
    assert tb.GpuEI(lambda t: 0.1 * t).xi(3) == pytest.approx(0.3)
    if os.path.isdir("/root/reference"):  # build container only: the real classes
        sys.path.insert(0, "/root/reference")
        sys.dont_write_bytecode = True
        import importlib
        import turbo.modules as tm
        importlib.reload(api)
        assert api.HAVE_TURBO and api.EI is tm.EI and api.SciKitModelInstance is tm.SciKitGPSurrogate.ModelInstance
        assert api.GPyModelInstance is tm.GPySurrogate.ModelInstance
        assert not issubclass(api.GPyModelInstance, api.SciKitModelInstance)


def test_gpy_surrogate_arguments_follow_the_reference():
    """turbo/modules/surrogates.py:101-158: same constructor arguments, same assertions, no device needed"""
    import turbo_b200 as tb
    from turbo_b200 import gpy_surrogate as G
    s = tb.GpuGPySurrogate()
    assert s.model_params == {'normalizer': True} and s.training_iterations == 10 and s.param_continuity
    assert s._get_training_iterations(3) == 10
    assert tb.GpuGPySurrogate(training_iterations=lambda t: [10, 5, 2][t % 3])._get_training_iterations(4) == 5
    assert tb.GpuGPySurrogate(training_iterations=None,
                              optimise_params={'num_restarts': 7})._get_training_iterations(0) == 7
    with pytest.raises(AssertionError):  # :138-139
        tb.GpuGPySurrogate(training_iterations=3, optimise_params={'num_restarts': 7})
    with pytest.raises(AssertionError):  # :156
        tb.GpuGPySurrogate(training_iterations=None)._get_training_iterations(0)
    with pytest.raises(NotImplementedError):
        tb.GpuGPySurrogate(sparse=True)
    with pytest.raises(ValueError):
        tb.GpuGPySurrogate(model_params={'inference_method': None})
    spec, prefix, params = G._describe_kernel(None, 4)  # GPRegression's default kernel: isotropic RBF, unit parameters
    assert (spec.kind, spec.ard, spec.has_const, spec.has_white, prefix) == (0, False, True, True, 'rbf')
    np.testing.assert_array_equal(params, [1.0, 1.0])
    spec, prefix, params = G._describe_kernel(tb.gpy_kernel('Mat32', ARD=True, variance=2.0, lengthscale=0.3), 3)
    assert (spec.kind, spec.ard, prefix) == (2, True, 'Mat32')
    np.testing.assert_array_equal(params, [2.0, 0.3, 0.3, 0.3])

    class Exponential:  # duck-typed GPy kernel object
        ARD, variance, lengthscale = False, np.array([0.5]), np.array([2.0])
    spec, prefix, params = G._describe_kernel(Exponential(), 2)
    assert (spec.kind, prefix) == (1, 'Exponential') and list(params) == [0.5, 2.0]


def test_reference_arm_uses_all_host_cores_under_torchrun_env():
    """torchrun exports OMP_NUM_THREADS=1; bench.py's CPU leg lifts the BLAS pools back to every core it may run on"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import bench, os, numpy as np; cfg = dict(bench.CONFIGS['headline'], n=128, d=3); "
            "X, y, th = bench.synthetic_problem(128, 3); arm = bench.CpuArm(cfg, X, y, th); "
            "r, top = arm.rate(bench.step_candidates_host(cfg, 1 << 20, 256)); "
            "print('THREADS', arm.threads, len(os.sched_getaffinity(0)), r > 0)")
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    m = re.search(r"THREADS (\d+) (\d+) True", out.stdout + out.stderr)  # bench.py keeps stdout for its one JSON line
    assert m, out.stdout + out.stderr
    assert int(m.group(1)) == int(m.group(2))


def test_lbfgs_errors_from_the_objective_propagate_and_leave_no_threads():
    """the guard around SciPy's private reverse-communication routine must cover the signature probe only --
    a TypeError / ValueError raised by the objective (the device call) is an error, not a reason to re-run the fit on
    the threads driver; and the threads driver must not leave workers blocked when the coordinator fails."""
    import threading
    from turbo_b200 import lbfgs
    bounds = np.array([[0.0, 1.0]] * 3)

    def good(X, members):
        return ((X - 0.3) ** 2).sum(1), 2 * (X - 0.3)
    x0 = np.random.default_rng(0).random((5, 3))
    for drv in ("auto", "threads"):
        x, f, it, n_eval = lbfgs.minimize_batched(good, x0, bounds, driver=drv)
        assert np.abs(x - 0.3).max() < 1e-6 and n_eval < 12
    assert lbfgs._reverse_communication_available() in (True, False)
    for drv, exc in (("auto", TypeError), ("threads", ValueError)):
        calls = []

        def bad(X, members):
            calls.append(len(members))
            if len(calls) > 2:
                raise exc("from the device call")
            return good(X, members)
        before = threading.active_count()
        with pytest.raises(exc):
            lbfgs.minimize_batched(bad, x0, bounds, driver=drv)
        assert len(calls) == 3              # not re-run on another driver
        assert threading.active_count() == before


def test_reference_install_recipe_is_verbatim():
    """baseline/_ref (git-ignored, shipped to the GPU box) holds the reference package byte for byte"""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref, "MANIFEST.json")):
        pytest.skip("baseline/_ref not installed on this machine")
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    try:
        import install_reference
        install_reference.verify()
    finally:
        sys.path.pop(0)
    from turbo_b200 import _reference_api
    assert _reference_api.HAVE_TURBO
