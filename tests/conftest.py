import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                  if not p.endswith("branin_run.npz"))


def load_golden(name):
    g = dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))
    for k in ("kind", "ard", "has_const", "has_white", "normalize_y", "ei_argmax", "ucb_argmax", "pi_argmax"):
        if k in g:
            g[k] = int(g[k])
    for k in ("incumbent", "xi", "beta", "lml"):
        if k in g:
            g[k] = float(g[k])
    g["extremum"] = str(g["extremum"])
    g["sign"] = 1.0 if g["extremum"] == "max" else -1.0
    return g


@pytest.fixture(params=golden_names())
def golden(request):
    g = load_golden(request.param)
    g["name"] = request.param
    return g


# ---------------------------------------------------------------------------------------------
# predictive-variance tolerance.  var = k** - v.v cancels near training points, so "1e-8 relative" needs an absolute
# floor of the FORM c * eps * n * prior_var.  Every GPU test that compares a variance with the oracle /
# the reference goes through check_variance(): it asserts the tolerance AND records the error actually achieved, in
# units of eps * n * prior_var; the session summary (gpurun_out/variance_error.json, committed per round under
# profiles/) is what the constant is tightened to.
# Round 2 (profiles/r2_variance_error_first.json, profiles/r2_predict_precision.md), error where 1e-8 relative does not
# already hold: FP64 DMMA and the 28- / 34-product INT8 engines <= 0.24 units on every fixture and size (VAR_FLOOR_EXACT
# = 4 leaves 16x); the 21-product engine, which 'auto' uses where its probe allows, <= 5.2 units at the BASELINE shapes
# and 35 on an ill-conditioned factor -- inside VAR_FLOOR_C = 64, the constant the ORACLE needs against the reference
# (tests/test_oracle.py); round 1 asserted 256.
# ---------------------------------------------------------------------------------------------
VAR_FLOOR_C = 64.0
VAR_FLOOR_EXACT = 4.0
_VAR_RECORDS = []


def check_variance(tag, var, var_ref, n, prior_var, rtol=1e-8, c=None):
    """assert |var - var_ref| <= rtol * var_ref + c * eps * n * prior_var element-wise; record the achieved error"""
    c = VAR_FLOOR_C if c is None else c
    var, var_ref = np.asarray(var, dtype=np.float64), np.asarray(var_ref, dtype=np.float64)
    unit = np.finfo(np.float64).eps * n * prior_var
    dv = np.abs(var - var_ref)
    outside = dv > rtol * var_ref  # candidates that need the floor at all
    _VAR_RECORDS.append({"case": tag, "n": int(n), "candidates": int(var.size),
                         "max_abs_err_over_eps_n_prior": float(dv.max() / unit) if var.size else 0.0,
                         "max_rel_err": float((dv / np.maximum(var_ref, 1e-300)).max()) if var.size else 0.0,
                         "need_floor": int(outside.sum()),
                         "max_floor_units_where_needed": float((dv[outside] / unit).max()) if outside.any() else 0.0})
    np.testing.assert_allclose(var, var_ref, rtol=rtol, atol=c * unit, err_msg=tag)


def pytest_sessionfinish(session, exitstatus):
    if not _VAR_RECORDS:
        return
    import json
    out = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        worst = max(r["max_floor_units_where_needed"] for r in _VAR_RECORDS)
        with open(os.path.join(out, "variance_error.json"), "w") as f:
            json.dump({"floor_constant_asserted": VAR_FLOOR_C, "worst_floor_units_where_needed": worst,
                       "records": _VAR_RECORDS}, f, indent=1)
    except OSError:
        pass
