"""Install the UNMODIFIED reference (mbway/turbo) into baseline/_ref so that it travels to the GPU box.

    python baseline/install_reference.py [--source /root/reference]

baseline/_ref/ is git-ignored (the reference's sources never enter this repository's history) but NOT
gpurun-ignored, so the drop-in tests (`tests/test_dropin.py`) and `bench.py --impl reference` can drive
the reference's own `turbo.Optimiser`, `turbo.Recorder`, `SciKitGPSurrogate.ModelInstance` and `EI` objects
on the GPU box, where /root/reference does not exist.

Recipe, in the order the task statement prescribes:
1. `pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref <copy>`.
   Outcome on this reference: FAILS -- "Neither 'setup.py' nor 'pyproject.toml' found" (the reference is a
   plain package directory plus a requirements.txt; it is meant to be put on PYTHONPATH).
2. Fallback = what its README asks of a user: put the package directory on the path.  The `turbo/` package
   is copied byte for byte; MANIFEST.json records the sha256 of every file so a reader can verify that
   nothing was edited (`python baseline/install_reference.py --verify`).
"""
import argparse
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
TARGET = os.path.join(HERE, "_ref")


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def _manifest(root):
    out = {}
    for base, _dirs, files in os.walk(os.path.join(root, "turbo")):
        for fn in sorted(files):
            if fn.endswith((".pyc", ".pyo")):
                continue
            p = os.path.join(base, fn)
            out[os.path.relpath(p, root)] = _sha(p)
    return out


def install(source="/root/reference", quiet=False):
    """-> dict(outcome=..., files=N).  No-op (outcome 'source absent') when the reference is not on this machine."""
    if not os.path.isdir(os.path.join(source, "turbo")):
        return {"outcome": "source absent", "files": 0}
    pip_msg = None
    with tempfile.TemporaryDirectory() as tmp:
        copy = os.path.join(tmp, "reference")
        shutil.copytree(source, copy, ignore=shutil.ignore_patterns(".git", "__pycache__"))
        r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation",
                            "--find-links", "/opt/wheelhouse", "--target", TARGET, copy],
                           capture_output=True, text=True)
        if r.returncode == 0 and os.path.isdir(os.path.join(TARGET, "turbo")):
            outcome = "pip install ok"
        else:
            pip_msg = (r.stderr.strip().splitlines() or ["pip failed"])[-1]
            dst = os.path.join(TARGET, "turbo")
            if os.path.isdir(dst):
                shutil.rmtree(dst)
            os.makedirs(TARGET, exist_ok=True)
            shutil.copytree(os.path.join(source, "turbo"), dst, ignore=shutil.ignore_patterns("__pycache__"))
            outcome = "pip not possible (%s); package directory copied unmodified" % pip_msg
    man = _manifest(TARGET)
    src_man = _manifest(source)
    assert man == src_man, "installed files differ from the source"
    with open(os.path.join(TARGET, "MANIFEST.json"), "w") as f:
        json.dump({"source": source, "outcome": outcome, "sha256": man}, f, indent=1, sort_keys=True)
    if not quiet:
        print("baseline/_ref: %s (%d files)" % (outcome, len(man)))
    return {"outcome": outcome, "files": len(man)}


def verify():
    with open(os.path.join(TARGET, "MANIFEST.json")) as f:
        man = json.load(f)["sha256"]
    now = _manifest(TARGET)
    bad = sorted(k for k in set(man) | set(now) if man.get(k) != now.get(k))
    if bad:
        raise SystemExit("baseline/_ref differs from its manifest: %s" % bad)
    print("baseline/_ref matches its manifest (%d files)" % len(man))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--source", default="/root/reference")
    ap.add_argument("--verify", action="store_true")
    a = ap.parse_args()
    if a.verify:
        verify()
    else:
        install(a.source)
