#!/bin/bash
set -u
mkdir -p gpurun_out
T=${TAG:-r01f}
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/${T}_pytest_gpu.log
