"""Parity gate T3: 1 GPU vs N GPUs give identical results.  Needs >= 2 visible GPUs."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_n_gpus() < 2, reason="needs at least 2 GPUs")
def test_multi_gpu_equals_single_gpu():
    n = min(_n_gpus(), 8)
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_multi_gpu_worker.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
                        "--master-port", "29655", worker], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("identical") == 2 and "DIFFERENT" not in r.stdout
