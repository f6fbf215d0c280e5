#!/bin/bash
# 2-GPU session: refinement test on one GPU, then the distributed V-cycle tests that fit two GPUs
set -u
mkdir -p gpurun_out
T=${TAG:-r02x}
timeout 600 python -m pytest tests/test_gpu_multigrid.py -m gpu -q --tb=short -s -k "refinement" > gpurun_out/${T}_pytest_mg.log 2>&1; echo "pytest mg rc=$?"
grep -E "passed|failed|plain:" gpurun_out/${T}_pytest_mg.log | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q --tb=short -rA -s -k "vcycle" > gpurun_out/${T}_pytest_dist.log 2>&1; echo "pytest dist rc=$?"
grep -E "^rank|passed|failed|^FAILED|^PASSED|^SKIPPED" gpurun_out/${T}_pytest_dist.log | cut -c1-330 | tail -n 20
