"""Multi-GPU parity (needs >= 2 GPUs; skipped on a 1-GPU box): torchrun, one process per GPU, NCCL inside the
library; state, compliance, mass and adjoint thickness gradient against the oracle's global solve."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("name", ["plate_4x20", "plate_varh", "plate_dg0", "warped"])
def test_two_gpu_solve_matches_oracle(built_lib, name):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    n = min(_ngpu(), 4) if name == "plate_4x20" else 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "dist_gpu_worker.py"), name]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    sys.stdout.write(r.stdout[-4000:]); sys.stderr.write(r.stderr[-4000:])
    assert r.returncode == 0
