#!/bin/bash
# dynamic path: GPU parity tests + config-4 timings
set -u
mkdir -p gpurun_out
T=${TAG:-r01h}
timeout 900 python -m pytest tests/test_gpu_dynamic.py -m gpu -q --tb=short > gpurun_out/${T}_pytest_dynamic.log 2>&1; echo "pytest rc=$?"; tail -n 60 gpurun_out/${T}_pytest_dynamic.log
timeout 600 python scripts/dyn_bench.py --large --out gpurun_out/${T}_dyn_bench.json > gpurun_out/${T}_dyn_bench.log 2>&1; echo "dyn bench rc=$?"; tail -n 20 gpurun_out/${T}_dyn_bench.log
