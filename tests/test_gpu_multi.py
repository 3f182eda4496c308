"""Multi-GPU parity (needs >= 2 GPUs; skipped otherwise): launches tests/multigpu_check.py under torchrun."""
import os
import subprocess
import sys
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharded_ensemble_on_two_gpus():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU: the N>1 path is covered on CPU by tests/test_sharding_gloo.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29611", os.path.join(ROOT, "tests", "multigpu_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MULTIGPU_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
