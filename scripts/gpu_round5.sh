#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tiled or assembly or parity" > gpurun_out/pytest_parity7.log 2>&1; echo "pytest parity rc=$?"; tail -n 12 gpurun_out/pytest_parity7.log
FEMO_VERBOSE=1 timeout 300 python scripts/quick_bench.py 816 20 > gpurun_out/quick7.log 2>&1; echo "quick rc=$?"; cat gpurun_out/quick7.log
SHORT="python scripts/quick_bench.py 816 2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_asm_tiles -s 2 -c 1 -o gpurun_out/prof_tiles7 $SHORT > gpurun_out/ncu_full7.log 2>&1
echo "ncu full rc=$?"
