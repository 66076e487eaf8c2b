"""NCCL path on real GPUs: needs at least two devices (skipped on the single-GPU box; run with `gpurun --gpus 2`).
The checks themselves live in tools/nccl_argmax_check.py (one process per GPU under torchrun)."""
import json
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def test_sharded_argmax_and_fit_over_nccl_match_single_process():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    n = min(torch.cuda.device_count(), 8)
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n),
                          "--master-addr", "127.0.0.1", "--master-port", str(port),
                          os.path.join(root, "tools", "nccl_argmax_check.py")], cwd=root, capture_output=True, text=True,
                         timeout=900)
    assert res.returncode == 0, (res.stdout + res.stderr)[-4000:]
    line = [ln for ln in res.stdout.splitlines() if ln.startswith("NCCL_CHECK ")][-1]
    out = json.loads(line[len("NCCL_CHECK "):])
    assert out["world"] == n and len(out["cases"]) == 4
    for c in out["cases"]:
        assert c["owners"]["rank_order"] != c["owners"]["reverse_order"]
