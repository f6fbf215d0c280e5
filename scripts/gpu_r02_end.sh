#!/bin/bash
# Last 1-GPU session of round 2: full GPU suite, smoke, both bench arms, launch list of the bench command, two variants of
# the tiled kernel for the record (staged halo records, 6+6 warp roles) -- the least important last
set -u
mkdir -p gpurun_out
T=${TAG:-r02end}
timeout 900 python -m pytest tests -m gpu -q --tb=short > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/${T}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/${T}_smoke.log | cut -c1-200
timeout 600 python bench.py --krylov-iters 50 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"; cut -c1-420 gpurun_out/${T}_bench.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference.json 2>> gpurun_out/${T}_bench.err; echo "bench ref rc=$?"; cut -c1-200 gpurun_out/${T}_bench_reference.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --krylov-iters 3 --no-mg > gpurun_out/${T}_ncu_list.log 2>&1; echo "ncu list rc=$?"
FEMO_TILE_HREC=1 timeout 120 python scripts/asm_experiments.py --masks 0 > gpurun_out/${T}_asm_hrec1.log 2>&1; grep mask gpurun_out/${T}_asm_hrec1.log
FEMO_TILE_WS=1 timeout 120 python scripts/asm_experiments.py --masks 0 > gpurun_out/${T}_asm_ws1.log 2>&1; grep mask gpurun_out/${T}_asm_ws1.log
