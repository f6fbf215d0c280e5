#!/bin/bash
# tiled assembly kernel: parity, timing, ncu
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity4.log 2>&1; echo "pytest parity rc=$?"; tail -n 25 gpurun_out/pytest_parity4.log
timeout 300 python scripts/quick_bench.py 816 20 > gpurun_out/quick4.log 2>&1; echo "quick rc=$?"; cat gpurun_out/quick4.log
FEMO_TILE_ORDER=identity timeout 300 python scripts/quick_bench.py 816 5 > gpurun_out/quick4_identity.log 2>&1; echo "quick identity rc=$?"; grep assemble gpurun_out/quick4_identity.log | tail -2
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_parity.py > gpurun_out/pytest_rest4.log 2>&1; echo "pytest rest rc=$?"; tail -n 8 gpurun_out/pytest_rest4.log
SHORT="python scripts/quick_bench.py 816 2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_asm_tiles -s 2 -c 1 -o gpurun_out/prof_tiles $SHORT > gpurun_out/ncu_full4.log 2>&1
echo "ncu full rc=$?"
