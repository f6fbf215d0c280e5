#!/bin/bash
set -u
mkdir -p gpurun_out
T=${TAG:-r01e}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/${T}_pytest_gpu.log
FEMO_VERBOSE=1 timeout 900 python bench.py --krylov-iters 50 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"; cat gpurun_out/${T}_bench.json; tail -n 6 gpurun_out/${T}_bench.err
