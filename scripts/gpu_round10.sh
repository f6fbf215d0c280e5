#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_golden.py -m gpu -q -s > gpurun_out/pytest_golden.log 2>&1; echo "pytest golden rc=$?"; grep -v "^$" gpurun_out/pytest_golden.log | tail -n 40
