#!/bin/bash
set -u
mkdir -p gpurun_out
T=${TAG:-9}
FEMO_TILE_WS=2 timeout 90 python scripts/quick_bench.py 64 2 > gpurun_out/quick${T}_small.log 2>&1; rc=$?; echo "small ws2 rc=$rc"; grep assemble gpurun_out/quick${T}_small.log | tail -1
if [ $rc -ne 0 ]; then echo "ABORT: ws2 failed on the small case"; tail -5 gpurun_out/quick${T}_small.log; exit 0; fi
FEMO_TILE_WS=2 FEMO_VERBOSE=1 timeout 120 python scripts/quick_bench.py 816 20 > gpurun_out/quick$T.log 2>&1; rc=$?; echo "quick ws2 rc=$rc"; grep -v "^peaks\|launches" gpurun_out/quick$T.log | tail -4
if [ $rc -ne 0 ]; then echo "ABORT: ws2 failed at cfg 3"; exit 0; fi
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tiled" > gpurun_out/pytest_parity$T.log 2>&1; echo "pytest parity rc=$?"; tail -n 3 gpurun_out/pytest_parity$T.log
SHORT="python scripts/quick_bench.py 816 2"
FEMO_TILE_WS=2 timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_asm_tiles -s 2 -c 1 -o gpurun_out/prof_tiles$T $SHORT > gpurun_out/ncu_full$T.log 2>&1
echo "ncu full rc=$?"
