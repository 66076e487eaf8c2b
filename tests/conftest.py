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
