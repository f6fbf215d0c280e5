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


@pytest.mark.parametrize("name,sw", [("plate_varh", 0), ("plate_dg0", 1)])
def test_two_gpu_time_stepping_matches_oracle(built_lib, name, sw):
    """SURVEY 8 f1 on 2 GPUs: history, energies, time adjoint and thickness gradient (tests/dist_gpu_dyn_worker.py)."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29534", os.path.join(ROOT, "tests", "dist_gpu_dyn_worker.py"), name, str(sw)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    sys.stdout.write(r.stdout[-4000:]); sys.stderr.write(r.stderr[-4000:])
    assert r.returncode == 0


@pytest.mark.parametrize("n", [2, 3])
def test_multi_gpu_hexahedral_solid_matches_oracle(built_lib, n):
    """SURVEY 8 f4 on N GPUs (DistributedSolid): state, compliance, volume, adjoint density gradient
    (tests/dist_gpu_hex_worker.py)."""
    if _ngpu() < n:
        pytest.skip("needs %d GPUs" % n)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29535", os.path.join(ROOT, "tests", "dist_gpu_hex_worker.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    sys.stdout.write(r.stdout[-4000:]); sys.stderr.write(r.stderr[-4000:])
    assert r.returncode == 0


@pytest.mark.parametrize("n", [2, 3])
def test_reverse_halo_exchange(built_lib, n):
    """owners += ghost copies (tests/dist_gpu_accumulate_worker.py)"""
    if _ngpu() < n:
        pytest.skip("needs %d GPUs" % n)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29536", os.path.join(ROOT, "tests", "dist_gpu_accumulate_worker.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    sys.stdout.write(r.stdout[-3000:]); sys.stderr.write(r.stderr[-3000:])
    assert r.returncode == 0


@pytest.mark.parametrize("n,n_dist", [(2, 1), (2, 3), (3, 2), (4, 3)])
def test_distributed_vcycle(built_lib, n, n_dist):
    """V-cycle preconditioned state + adjoint solves with `n_dist` distributed levels over replicated coarser ones
    (tests/dist_gpu_mg_worker.py): parity with the oracle, iteration count of the single-GPU hierarchy."""
    if _ngpu() < n:
        pytest.skip("needs %d GPUs" % n)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29537", os.path.join(ROOT, "tests", "dist_gpu_mg_worker.py"), str(n_dist)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    sys.stdout.write(r.stdout[-3000:]); sys.stderr.write(r.stderr[-3000:])
    assert r.returncode == 0
