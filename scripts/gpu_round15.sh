#!/bin/bash
# r01j: full GPU suite after the ordered tear-down of hierarchy / dynamic handles
set -u
mkdir -p gpurun_out
T=${TAG:-r01j}
timeout 1500 python -m pytest tests -m gpu -q --tb=short > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 30 gpurun_out/${T}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/${T}_smoke.log
