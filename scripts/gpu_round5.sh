#!/bin/bash
set -u
mkdir -p gpurun_out
T=${TAG:-9}
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tiled" > gpurun_out/pytest_parity$T.log 2>&1; echo "pytest parity rc=$?"; tail -n 3 gpurun_out/pytest_parity$T.log
FEMO_VERBOSE=1 timeout 300 python scripts/quick_bench.py 816 20 > gpurun_out/quick$T.log 2>&1; echo "quick rc=$?"; grep -v "^peaks\|launches" gpurun_out/quick$T.log | tail -4
FEMO_TILE_SORT=1 timeout 300 python scripts/quick_bench.py 816 2 > gpurun_out/quick${T}_sort1.log 2>&1; echo "sort by type only:"; grep assemble gpurun_out/quick${T}_sort1.log | tail -1
SHORT="python scripts/quick_bench.py 816 2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_asm_tiles -s 2 -c 1 -o gpurun_out/prof_tiles$T $SHORT > gpurun_out/ncu_full$T.log 2>&1
echo "ncu full rc=$?"
