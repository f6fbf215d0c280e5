#!/bin/bash
# multigrid: parity tests, then small and config-3-scale solves
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multigrid.py -m gpu -x -q > gpurun_out/pytest_mg.log 2>&1; rc=$?; echo "pytest mg rc=$rc"; tail -n 25 gpurun_out/pytest_mg.log
timeout 300 python scripts/mg_bench.py 13 3 3 > gpurun_out/mg_small.json 2> gpurun_out/mg_small.err; echo "mg small rc=$?"; cat gpurun_out/mg_small.json; tail -3 gpurun_out/mg_small.err
if [ $rc -ne 0 ]; then exit 0; fi
timeout 900 python scripts/mg_bench.py 51 5 3 > gpurun_out/mg_cfg3.json 2> gpurun_out/mg_cfg3.err; echo "mg cfg3 rc=$?"; cat gpurun_out/mg_cfg3.json; tail -3 gpurun_out/mg_cfg3.err
